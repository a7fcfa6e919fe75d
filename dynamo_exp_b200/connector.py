"""KV transfer connector over libkvdp.so — the counterpart of the patched vLLM `DynamoNixlConnector`
(container/deps/vllm/vllm_v0.8.4-dynamo-kv-disagg-patch.patch:1081-1492) with the same method surface:

    register_kv_caches, get_agent_metadata, add_remote_agent, write_blocks, read_blocks, get_done_tranfers,
    get_new_notifs, shutdown                                  (method names as in the reference, typo included)

What changes underneath:
  * no Triton rearrange pass and no staging blocks: the TP head reshard happens in the same kernel that moves the
    bytes, so `staging_block_ids` is accepted and ignored, and block ids need not be contiguous (the reference
    coalesces ranges, patch:1201-1294, only to cut its descriptor and launch counts);
  * no NIXL/UCX: remote caches are CUDA-IPC peer mappings (or raw pointers when the "remote" agent lives in the same
    process); WRITE = kernel on this GPU storing into the peer, READ = kernel on this GPU loading from the peer;
  * decode TP smaller than prefill TP is supported (the reference asserts against it, patch:1437);
  * notifications: after a write the sender drops `notify_msg` into a mailbox ring in the receiver's device memory
    and bumps the ring head with a system-scope release store, all in stream order behind the data. Every sending
    agent (engine id + rank) claims a lane of its own in each receiver it talks to, so any number of prefill engines
    can feed one decode engine and one prefill engine can feed several; a sender never overwrites a slot the receiver
    has not read (it waits on the receiver's consumed counter instead).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import msgpack
import numpy as np

from . import _lib
from .storage import KvError
from .transfer import CopyPlan, TransferEngine

MAILBOX_LANES = 64          # senders one receiver rank can hear from at a time (each claims its own lane)
MAILBOX_SLOTS = 64          # ring depth per lane
SLOT_BYTES = 256            # u32 msg_len | u32 seq | u32 name_len | u32 0 | 48 bytes sender agent name | 192 bytes message
NAME_BYTES = 48
MAX_MSG = SLOT_BYTES - 16 - NAME_BYTES
CLAIMS_BYTES = MAILBOX_LANES * 8                       # u64 key of the sender that owns the lane, 0 = free
HEADS_OFF = CLAIMS_BYTES
HEADS_BYTES = MAILBOX_LANES * 64                       # one 64-byte line per lane head (u32 messages delivered)
CONSUMED_OFF = HEADS_OFF + HEADS_BYTES
CONSUMED_BYTES = MAILBOX_LANES * 4                     # u32 messages the receiver has read, per lane (flow control)
SLOTS_OFF = CONSUMED_OFF + CONSUMED_BYTES
MAILBOX_BYTES = SLOTS_OFF + MAILBOX_LANES * MAILBOX_SLOTS * SLOT_BYTES
STAGE_SLOTS = 64            # sender-side pinned staging ring, shared by every peer of this connector


def _head_off(lane: int) -> int:
    return HEADS_OFF + lane * 64


def _slot_off(lane: int, slot: int) -> int:
    return SLOTS_OFF + (lane * MAILBOX_SLOTS + slot) * SLOT_BYTES


def _agent_key(name: str) -> int:
    """Non-zero 64-bit FNV-1a of a sender's agent name: what a lane's claim word holds."""
    h = 0xCBF29CE484222325
    for b in name.encode("utf-8"):
        h = ((h ^ b) * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return h or 1


@dataclass
class KvdpMetadata:
    """What one engine publishes for its peers (the reference's `NixlMetadata`, patch:1081-1089; stored in etcd by
    examples/llm/utils/nixl.py:59-116). `agent_metadata[rank]` is opaque bytes produced by `get_agent_metadata()`."""
    engine_id: str
    agent_metadata: List[bytes]
    kv_caches_base_addr: List[List[List[int]]]   # [rank][layer][K, V] (informational; peers use the IPC handles)
    num_blocks: int

    def encode(self) -> bytes:
        return msgpack.packb({"engine_id": self.engine_id, "agent_metadata": self.agent_metadata,
                              "kv_caches_base_addr": self.kv_caches_base_addr, "num_blocks": self.num_blocks},
                             use_bin_type=True)

    @staticmethod
    def decode(raw: bytes) -> "KvdpMetadata":
        d = msgpack.unpackb(raw, raw=False)
        return KvdpMetadata(d["engine_id"], list(d["agent_metadata"]), d["kv_caches_base_addr"], d["num_blocks"])


# ---------------------------------------------------------------------------------------------- metadata exchange
class DirKvClient:
    """The two etcd calls `NixlMetadataStore` makes (`kv_create_or_validate`, `kv_get_prefix`), backed by a directory:
    enough for engines on one box (where CUDA-IPC handles are valid anyway). A deployment passes its own client."""

    def __init__(self, root: str):
        self.root = root
        os.makedirs(root, exist_ok=True)

    def _path(self, key: str) -> str:
        return os.path.join(self.root, key.replace("/", "__"))

    def kv_create_or_validate(self, key: str, value: bytes) -> None:
        path = self._path(key)
        try:
            fd = os.open(path + ".tmp%d" % os.getpid(), os.O_WRONLY | os.O_CREAT | os.O_EXCL, 0o600)
        except FileExistsError:
            raise KvError(f"key {key} is being written by someone else")
        with os.fdopen(fd, "wb") as f:
            f.write(value)
        if os.path.exists(path):
            with open(path, "rb") as f:
                existing = f.read()
            os.remove(path + ".tmp%d" % os.getpid())
            if existing != value:
                raise KvError(f"a different value exists for {key}")
            return
        os.rename(path + ".tmp%d" % os.getpid(), path)

    def kv_get_prefix(self, key: str) -> List[dict]:
        pre = os.path.basename(self._path(key))
        out = []
        for name in sorted(os.listdir(self.root)):
            if name.startswith(pre) and ".tmp" not in name:
                with open(os.path.join(self.root, name), "rb") as f:
                    out.append({"key": name.replace("__", "/"), "value": f.read()})
        return out


class KvdpMetadataStore:
    """reference: `NixlMetadataStore` (examples/llm/utils/nixl.py:59-116): engines publish their metadata under
    `<namespace>/nixl_metadata/<engine_id>` and look their peers up by engine id; lookups are cached. Synchronous; the
    client only needs the two calls of `DirKvClient`."""
    NIXL_METADATA_KEY = "nixl_metadata"

    def __init__(self, namespace: str, client) -> None:
        if client is None:
            raise KvError("Cannot be used with static workers")
        self._namespace = namespace
        self._stored: set = set()
        self._cached: Dict[str, KvdpMetadata] = {}
        self._client = client
        self._key_prefix = f"{namespace}/{KvdpMetadataStore.NIXL_METADATA_KEY}"

    def put(self, engine_id: str, metadata: KvdpMetadata) -> None:
        key = "/".join([self._key_prefix, engine_id])
        try:
            self._client.kv_create_or_validate(key, metadata.encode())
        except Exception as e:                      # the reference only warns (nixl.py:83-84)
            import warnings
            warnings.warn(f"A different metadata exists for engine {engine_id}: {e}")
        self._stored.add(engine_id)

    def get(self, engine_id: str) -> KvdpMetadata:
        if engine_id in self._cached:
            return self._cached[engine_id]
        key = "/".join([self._key_prefix, engine_id])
        for item in self._client.kv_get_prefix(key):
            meta = KvdpMetadata.decode(item["value"])
            self._cached[engine_id] = meta
            return meta
        raise KvError(f"Error retrieving metadata for engine {engine_id}")


def tp_targets(src_tp: int, dst_tp: int, src_rank: int) -> List[int]:
    """Ranks of the other engine that rank `src_rank` exchanges blocks with (patch:1427-1471 for dst_tp >= src_tp:
    `rank * tp_multiplier + i`; the dst_tp < src_tp direction is the inverse map)."""
    if src_tp < 1 or dst_tp < 1 or not (0 <= src_rank < src_tp):
        raise KvError("need src_tp, dst_tp >= 1 and 0 <= src_rank < src_tp")
    if dst_tp >= src_tp:
        if dst_tp % src_tp:
            raise KvError("dst_tp must be a multiple of src_tp")
        m = dst_tp // src_tp
        return [src_rank * m + i for i in range(m)]
    if src_tp % dst_tp:
        raise KvError("src_tp must be a multiple of dst_tp")
    return [src_rank // (src_tp // dst_tp)]


def get_ranges(block_ids: Sequence[int]) -> List[List[int]]:
    """Maximal runs of consecutive ids, `[0,1,2,4,5,6] -> [[0,2],[4,6]]` (reference: `_get_ranges`, patch:1201-1213).
    The connector here does not need it (no staging, no per-range launches); kept for callers that mirror the
    reference's bookkeeping and for the host tier's DMA coalescing."""
    ranges: List[List[int]] = []
    for i, b in enumerate(block_ids):
        if i == 0 or b != block_ids[i - 1] + 1:
            ranges.append([b, b])
        else:
            ranges[-1][1] = b
    return ranges


def get_same_length_ranges(src_ranges, dst_ranges):
    """Split two range lists so that corresponding ranges have equal length (reference: `_get_same_length_ranges`,
    patch:1247-1294; its example: src [[0,2],[4,8]], dst [[1,3],[5,7],[9,10]] ->
    ([[0,2],[4,6],[7,8]], [[1,3],[5,7],[9,10]]))."""
    src = [list(r) for r in src_ranges]
    dst = [list(r) for r in dst_ranges]
    out_s, out_d = [], []
    i = j = 0
    while i < len(src) and j < len(dst):
        ls = src[i][1] - src[i][0] + 1
        ld = dst[j][1] - dst[j][0] + 1
        if ls == ld:
            out_s.append([src[i][0], src[i][1]])
            out_d.append([dst[j][0], dst[j][1]])
            i += 1
            j += 1
        elif ls > ld:
            out_s.append([src[i][0], src[i][0] + ld - 1])
            out_d.append([dst[j][0], dst[j][1]])
            src[i] = [src[i][0] + ld, src[i][1]]
            j += 1
        else:
            out_s.append([src[i][0], src[i][1]])
            out_d.append([dst[j][0], dst[j][0] + ls - 1])
            dst[j] = [dst[j][0] + ls, dst[j][1]]
            i += 1
    return out_s, out_d


def rearrange_tensors(t1_ptr: int, t2_ptr: int, shape: Sequence[int], element_size: int, d: int, direction: str,
                      device: int = -1, engine: Optional[TransferEngine] = None) -> None:
    """Device counterpart of the reference's Triton `rearrange_tensors(t1, t2, d, direction)` (patch:934-1031) for one
    K-or-V tensor of shape [N, B, H, C]:
        write: t2 viewed as [d][N][B][H/d][C]  <-  t1[n][b][g*H/d + h'][c]
        read : t1  <-  t2 (the inverse)
    The transfer connector never calls it (the reshard is fused into the transfer); it exists so that code written
    against the reference's two-step scheme keeps working, and as a directly testable statement of the index map.
    Unlike the Triton kernels it needs no `N*B*H*C % 1024 == 0`."""
    N, B, H, Cc = [int(x) for x in shape]
    if d < 1 or H % d:
        raise KvError("H must be divisible by d")
    if direction not in ("read", "write"):
        raise KvError(f"Invalid direction: {direction}")
    run = (H // d) * Cc * element_size
    row = H * Cc * element_size
    own = engine is None
    eng = engine or TransferEngine(max(N, 1), device)
    ids = list(range(N))
    plans = []
    try:
        if direction == "write":
            desc = _lib.PlanDesc()
            src = np.array([t1_ptr], dtype=np.uint64)
            dst = np.array([t2_ptr + g * N * B * run for g in range(d)], dtype=np.uint64)
            desc.num_layers, desc.row_groups, desc.prefix_dim, desc.rows = 1, d, 1, B
            desc.run_bytes = run
            desc.src_block_stride, desc.dst_block_stride, desc.dst_row_stride = B * row, B * run, run
            desc.src_blocks = desc.dst_blocks = N
            desc.src_layer_ptrs, desc.dst_layer_ptrs, desc.device = src.ctypes.data, dst.ctypes.data, device
            plans.append(CopyPlan.from_desc(desc, keepalive=(src, dst)))
        else:
            for g in range(d):   # group g of the staging layout lands in column range g of every t1 row
                desc = _lib.PlanDesc()
                src = np.array([t2_ptr + g * N * B * run], dtype=np.uint64)
                dst = np.array([t1_ptr], dtype=np.uint64)
                desc.num_layers, desc.row_groups, desc.prefix_dim, desc.rows = 1, 1, 1, B
                desc.run_bytes = run
                desc.src_block_stride, desc.dst_block_stride = B * run, B * row
                desc.dst_row_stride, desc.dst_col_offset = row, g * run
                desc.src_blocks = desc.dst_blocks = N
                desc.src_layer_ptrs, desc.dst_layer_ptrs, desc.device = src.ctypes.data, dst.ctypes.data, device
                plans.append(CopyPlan.from_desc(desc, keepalive=(src, dst)))
        for p in plans:
            eng.run(p, ids, ids, num_layers=1)
        eng.sync()
    finally:
        for p in plans:
            p.close()
        if own:
            eng.close()


def split_roles(world: int, interleaved: bool = False):
    """Prefill:decode split used by the multi-GPU bench (north star: 2/4/8-GPU prefill:decode splits). Default: the
    first world // 2 ranks prefill, the rest decode, prefill rank r paired with decode rank P + r % D. `interleaved`
    (even world only): even ranks prefill and push to the next odd rank, so that neighbouring GPUs, which usually hang
    off the same PCIe switch, are not both reading host memory. Returns (prefill_ranks, decode_ranks, pairs)."""
    if world < 2:
        raise KvError("a prefill:decode split needs at least 2 ranks")
    if interleaved and world % 2 == 0:
        prefill, decode = list(range(0, world, 2)), list(range(1, world, 2))
        return prefill, decode, [(r, r + 1) for r in prefill]
    p = world // 2
    d = world - p
    prefill, decode = list(range(p)), list(range(p, world))
    return prefill, decode, [(r, p + r % d) for r in prefill]


@dataclass
class _Remote:
    tp: int
    num_blocks: int
    targets: List[int]
    layer_ptrs: Dict[int, List[int]] = field(default_factory=dict)    # target rank -> per-layer base pointers
    mailbox: Dict[int, int] = field(default_factory=dict)             # target rank -> mailbox base
    lane: Dict[int, int] = field(default_factory=dict)                # target rank -> the lane this sender claimed there
    mapped: List[int] = field(default_factory=list)                   # IPC mappings to close
    push: Optional[CopyPlan] = None
    pull: List[CopyPlan] = field(default_factory=list)
    heads_per_rank: int = 0


class DynamoKvdpConnector:
    def __init__(self, engine_id: str, rank: int, tp_size: int, device: int = 0, is_mla: bool = False,
                 max_blocks_per_transfer: int = 4096):
        _lib.require_gpu()
        self.engine_id = engine_id
        self.rank = int(rank)
        self.device = int(device)
        self._tp_size = {engine_id: int(tp_size)}
        self._is_mla = bool(is_mla)
        self.num_layers = None
        self.num_blocks = None
        self.num_heads = None
        self.block_len = None
        self.block_size = None
        self.kv_caches_base_addr: Dict[str, list] = {}
        self._layer_ptrs: List[int] = []
        self._remotes: Dict[str, _Remote] = {}
        self._engine = TransferEngine(max_blocks_per_transfer, device)
        L = _lib.lib()
        p = C.c_void_p()
        _lib.check(L.kvdp_device_alloc(C.byref(p), MAILBOX_BYTES, device))
        self._mailbox = p.value
        self._seen = [0] * MAILBOX_LANES
        # sender-side: pinned message staging ring (shared by all peers, recycled against the progress word) + a pinned
        # progress word the stream bumps after each transfer
        hp = C.c_void_p()
        _lib.check(L.cuda_malloc_host(C.byref(hp), STAGE_SLOTS * SLOT_BYTES + 64))
        self._host = hp.value
        C.memset(self._host, 0, STAGE_SLOTS * SLOT_BYTES + 64)
        self._progress = np.frombuffer((C.c_uint32 * 16).from_address(self._host + STAGE_SLOTS * SLOT_BYTES),
                                       dtype=np.uint32)
        self._send_seq: Dict[tuple, int] = {}     # (engine, target rank) -> messages sent on the lane claimed there
        self._peer_consumed: Dict[tuple, int] = {}  # (engine, target rank) -> receiver's consumed count, as last read
        self._stage_n = 0                         # staging slots handed out so far
        self._stage_cover = [0] * STAGE_SLOTS     # transfer sequence number whose completion frees the slot
        self._xfer_seq = 0
        self._transfers: Dict[str, int] = {}      # notify_msg -> local sequence number
        self.notify_timeout_s = 30.0              # how long a sender waits for a receiver that does not drain its ring

    # ---------------------------------------------------------------------------------------------- registration
    @property
    def agent_name(self) -> str:
        return f"{self.engine_id}:{self.rank}"

    def register_kv_caches(self, kv_caches: Sequence, *, shape: Optional[Sequence[int]] = None,
                           element_size: Optional[int] = None) -> None:
        """`kv_caches`: one entry per layer — either a tensor-like with data_ptr()/shape/element_size() (vLLM layout
        `[2, num_blocks, block_size, num_heads, head_dim]`; MLA `[num_blocks, block_size, head_dim]`, patch:1131-1186)
        or a raw device pointer together with `shape` and `element_size`."""
        if not kv_caches:
            raise KvError("kv_caches must not be empty")
        first = kv_caches[0]
        if hasattr(first, "data_ptr"):
            shape = tuple(first.shape)
            element_size = first.element_size()
            ptrs = [int(t.data_ptr()) for t in kv_caches]
        else:
            if shape is None or element_size is None:
                raise KvError("raw pointers need shape and element_size")
            ptrs = [int(p) for p in kv_caches]
        if self._is_mla:
            num_blocks, block_size, head_dim = shape
            self.num_heads, self.num_cache_entries = 1, 1
            self.block_len = head_dim * block_size * element_size
        else:
            two, num_blocks, block_size, num_heads, head_dim = shape
            if two != 2:
                raise KvError("expected a [2, num_blocks, block_size, num_heads, head_dim] cache per layer")
            self.num_heads, self.num_cache_entries = num_heads, 2
            self.block_len = block_size * num_heads * head_dim * element_size
        self.block_size = block_size
        self.num_layers = len(ptrs)
        self.num_blocks = num_blocks
        self._layer_ptrs = ptrs
        self.kv_caches_base_addr[self.engine_id] = [
            [p] if self._is_mla else [p, p + num_blocks * self.block_len] for p in ptrs]

    def get_agent_metadata(self) -> bytes:
        """Opaque per-rank blob: IPC handles (+ offsets) of every layer and of the mailbox, raw pointers for same-process
        peers, and the geometry the peer needs to validate."""
        if not self._layer_ptrs:
            raise KvError("register_kv_caches must be called first")
        L = _lib.lib()

        def export(ptr):
            h = (C.c_uint8 * _lib.IPC_HANDLE_BYTES)()
            off = C.c_uint64(0)
            _lib.check(L.kvdp_ipc_export_range(C.c_void_p(ptr), h, C.byref(off)))
            return [bytes(h), int(off.value)]
        return msgpack.packb({
            "pid": os.getpid(), "rank": self.rank, "device": self.device, "tp": self._tp_size[self.engine_id],
            "num_layers": self.num_layers, "num_blocks": self.num_blocks, "block_len": self.block_len,
            "block_size": self.block_size, "is_mla": self._is_mla,
            "layers": [export(p) for p in self._layer_ptrs], "layer_ptrs": self._layer_ptrs,
            "mailbox": export(self._mailbox), "mailbox_ptr": self._mailbox}, use_bin_type=True)

    def add_remote_agent(self, engine_id: str, agent_metadata: List[bytes], agent_tp: int,
                         kv_caches_base_addr=None, num_blocks: Optional[int] = None) -> List[str]:
        """reference: patch:1427-1471. Opens the peer caches this rank talks to and builds the push / pull plans."""
        if not self._layer_ptrs:
            raise KvError("register_kv_caches must be called first")
        my_tp = self._tp_size[self.engine_id]
        self._tp_size[engine_id] = int(agent_tp)
        if len(agent_metadata) != agent_tp:
            raise KvError(f"expected {agent_tp} agent metadata blobs, got {len(agent_metadata)}")
        targets = tp_targets(my_tp, agent_tp, self.rank)
        metas = [msgpack.unpackb(m, raw=False) for m in agent_metadata]
        rem = _Remote(tp=int(agent_tp), num_blocks=int(num_blocks if num_blocks is not None else
                                                       metas[targets[0]]["num_blocks"]), targets=targets)
        L = _lib.lib()

        def open_ptr(meta, handle_off, raw_ptr):
            if meta["pid"] == os.getpid():
                return int(raw_ptr)           # same process: the pointer is directly usable (peer access permitting)
            hb = (C.c_uint8 * _lib.IPC_HANDLE_BYTES).from_buffer_copy(handle_off[0])
            p = C.c_void_p()
            _lib.check(L.kvdp_ipc_import(hb, self.device, C.byref(p)))
            rem.mapped.append(p.value)
            return p.value + int(handle_off[1])

        for q in targets:
            m = metas[q]
            if m["num_layers"] != self.num_layers or m["block_size"] != self.block_size or m["is_mla"] != self._is_mla:
                raise KvError("remote cache geometry does not match the local one")
            if m["pid"] == os.getpid() and m["device"] != self.device:
                _lib.check(L.kvdp_peer_enable(self.device, m["device"]))
            rem.layer_ptrs[q] = [open_ptr(m, h, raw) for h, raw in zip(m["layers"], m["layer_ptrs"])]
            rem.mailbox[q] = open_ptr(m, m["mailbox"], m["mailbox_ptr"])
            rem.lane[q] = self._claim_lane(rem.mailbox[q])
            # a sender that comes back under the same name continues the lane's numbering instead of restarting at 1
            word = np.zeros(1, dtype=np.uint32)
            _lib.check(L.cuda_memcpy_sync(C.c_void_p(word.ctypes.data),
                                          C.c_void_p(rem.mailbox[q] + _head_off(rem.lane[q])), 4))
            self._send_seq[(engine_id, q)] = max(self._send_seq.get((engine_id, q), 0), int(word[0]))
        self.kv_caches_base_addr[engine_id] = kv_caches_base_addr
        self._remotes[engine_id] = rem
        self._build_plans(engine_id, rem, metas)
        return [f"{engine_id}:{q}" for q in targets]

    def _build_plans(self, engine_id: str, rem: _Remote, metas) -> None:
        my_tp = self._tp_size[self.engine_id]
        if self._is_mla:
            # latent cache: whole blocks, replicated to every target rank (patch:1440-1441,1451)
            suffix = self.block_len
            rem.push = None
            rem.pull = []
            rem._mla_push = [CopyPlan.block_copy_raw(self._layer_ptrs, rem.layer_ptrs[q], 1, self.num_blocks,
                                                     rem.num_blocks, suffix, 1, self.device) for q in rem.targets]
            rem.pull = [CopyPlan.block_copy_raw(rem.layer_ptrs[rem.targets[0]], self._layer_ptrs, 1, rem.num_blocks,
                                                self.num_blocks, suffix, 1, self.device)]
            return
        row_bytes = self.block_len // self.block_size                      # heads_here * head_dim * elem
        remote_row = metas[rem.targets[0]]["block_len"] // self.block_size
        if row_bytes * my_tp != remote_row * rem.tp:
            raise KvError("local and remote caches disagree on the model's total KV row size")
        rem.push = CopyPlan.reshard(self._layer_ptrs, [rem.layer_ptrs[q] for q in rem.targets], self.num_blocks,
                                    rem.num_blocks, self.block_size, row_bytes, my_tp, rem.tp, self.rank, self.device)
        # pull = the remote ranks' reshard toward this rank, executed here with the remote caches as the source
        rem.pull = [CopyPlan.reshard(rem.layer_ptrs[q], [self._layer_ptrs], rem.num_blocks, self.num_blocks,
                                     self.block_size, remote_row, rem.tp, my_tp, q, self.device)
                    for q in rem.targets] if rem.tp >= my_tp else [
            # remote TP smaller than ours: one remote rank holds a superset of our heads; take our sub-rows
            self._sub_row_pull_plan(rem, remote_row, row_bytes)]

    def _sub_row_pull_plan(self, rem: _Remote, remote_row: int, row_bytes: int) -> CopyPlan:
        """Remote TP smaller than ours: rank q = rank // m holds rows m times wider; read only our column range
        [j*row_bytes, (j+1)*row_bytes), j = rank % m, of every remote token row (strided source rows)."""
        q = rem.targets[0]
        m = self._tp_size[self.engine_id] // rem.tp
        j = self.rank % m
        src = np.ascontiguousarray(rem.layer_ptrs[q], dtype=np.uint64)
        dst = np.ascontiguousarray(self._layer_ptrs, dtype=np.uint64)
        d = _lib.PlanDesc()
        d.num_layers, d.row_groups, d.prefix_dim, d.rows = self.num_layers, 1, 2, self.block_size
        d.run_bytes = row_bytes
        d.src_prefix_stride = rem.num_blocks * self.block_size * remote_row
        d.src_block_stride = self.block_size * remote_row
        d.src_row_stride = remote_row
        d.src_col_offset = j * row_bytes
        d.dst_prefix_stride = self.num_blocks * self.block_len
        d.dst_block_stride = self.block_len
        d.dst_row_stride = row_bytes
        d.dst_col_offset = 0
        d.src_blocks, d.dst_blocks = rem.num_blocks, self.num_blocks
        d.src_layer_ptrs = src.ctypes.data
        d.dst_layer_ptrs = dst.ctypes.data
        d.device = self.device
        return CopyPlan.from_desc(d, keepalive=(src, dst))

    # ---------------------------------------------------------------------------------------------- transfers
    def write_blocks(self, local_block_ids, staging_block_ids, remote_block_ids, dst_engine_id: str,
                     notify_msg: str) -> None:
        """reference: patch:1361-1421. Asynchronous; completion is visible through get_done_tranfers() here and
        get_new_notifs() on the receiver."""
        rem = self._remote(dst_engine_id)
        local_block_ids = list(local_block_ids)
        remote_block_ids = list(remote_block_ids)[:len(local_block_ids)]   # patch:1368
        if len(remote_block_ids) != len(local_block_ids):
            raise KvError("fewer remote than local block ids")
        if local_block_ids:
            if self._is_mla:
                for plan in rem._mla_push:
                    self._engine.run(plan, local_block_ids, remote_block_ids, num_layers=self.num_layers)
            else:
                self._engine.run(rem.push, local_block_ids, remote_block_ids, num_layers=self.num_layers)
        for q in rem.targets:                                              # empty => notify only (patch:1373-1377)
            self._post(dst_engine_id, q, rem, notify_msg)
        self._xfer_seq += 1
        self._transfers[notify_msg] = self._xfer_seq
        self._engine.signal(self._host + STAGE_SLOTS * SLOT_BYTES, self._xfer_seq)

    def write_token_rows(self, local_slots, remote_slots, dst_engine_id: str, notify_msg: str) -> None:
        """Token-granular form of write_blocks: `slot = block * block_size + offset` on either side, every layer, K and
        V, with the same TP head resharding, in one launch per request. This is the KV half of the reference's NCCL
        connector (`send_kv_caches_and_hidden_states` / `recv_...`, patch:1602-1766: per layer a gather by
        `slot_mapping` with a head slice, a send, and `reshape_and_cache_flash` on the other side). Asynchronous,
        completion and notification as for write_blocks."""
        rem = self._remote(dst_engine_id)
        local_slots, remote_slots = list(local_slots), list(remote_slots)
        if len(local_slots) != len(remote_slots):
            raise KvError("local and remote slot lists must have the same length")
        if local_slots and self._is_mla:
            # latent cache: whole token rows, no head split, the same rows to every target rank (the reference's
            # deepseek branch: `keys.append(key_cache[current_slot_mapping])` / `kv_cache[slot_mapping] = copy_from`,
            # patch:1662-1665, 1748-1752)
            if getattr(rem, "_mla_push_tokens", None) is None:
                row = self.block_len // self.block_size
                rem._mla_push_tokens = [CopyPlan.block_copy_raw(
                    self._layer_ptrs, rem.layer_ptrs[q], 1, self.num_blocks * self.block_size,
                    rem.num_blocks * self.block_size, row, 1, self.device) for q in rem.targets]
            for plan in rem._mla_push_tokens:
                self._engine.run(plan, local_slots, remote_slots, num_layers=self.num_layers)
        elif local_slots:
            if getattr(rem, "push_tokens", None) is None:
                # the block reshard plan with one-token "blocks": rows = 1, runs of one (sub-)row
                rem.push_tokens = CopyPlan.reshard(
                    self._layer_ptrs, [rem.layer_ptrs[q] for q in rem.targets], self.num_blocks * self.block_size,
                    rem.num_blocks * self.block_size, 1, self.block_len // self.block_size,
                    self._tp_size[self.engine_id], rem.tp, self.rank, self.device)
            self._engine.run(rem.push_tokens, local_slots, remote_slots, num_layers=self.num_layers)
        for q in rem.targets:
            self._post(dst_engine_id, q, rem, notify_msg)
        self._xfer_seq += 1
        self._transfers[notify_msg] = self._xfer_seq
        self._engine.signal(self._host + STAGE_SLOTS * SLOT_BYTES, self._xfer_seq)

    def read_blocks(self, local_block_ids, staging_block_ids, remote_block_ids, dst_engine_id: str) -> None:
        """reference: patch:1296-1359. Blocking, like the reference (it polls the NIXL handles to DONE)."""
        rem = self._remote(dst_engine_id)
        local_block_ids, remote_block_ids = list(local_block_ids), list(remote_block_ids)
        if len(local_block_ids) != len(remote_block_ids):
            raise KvError("local and remote block id lists must have the same length")   # assert at patch:1299
        if not local_block_ids:
            return
        if not rem.pull:
            raise KvError("no pull plan for this peer")
        for plan in rem.pull:
            self._engine.run(plan, remote_block_ids, local_block_ids, num_layers=self.num_layers,
                             check_dst_unique=True)
        self._engine.sync()

    def _claim_lane(self, mailbox_ptr: int) -> int:
        """Take (or find again) this sender's lane in a receiver's mailbox: a system-scope CAS on the lane's claim
        word, probing from hash(agent_name). Every (engine, rank) that writes to a receiver gets a lane of its own there,
        like NIXL's per-agent notification queues, so two prefill engines with equal rank numbers never share one."""
        L = _lib.lib()
        key = _agent_key(self.agent_name)
        for probe in range(MAILBOX_LANES):
            lane = (key + probe) % MAILBOX_LANES
            old = C.c_uint64(0)
            _lib.check(L.kvdp_flag_cas64(C.c_void_p(mailbox_ptr + lane * 8), 0, key, C.byref(old), self.device))
            if old.value in (0, key):
                return lane
        raise KvError(f"all {MAILBOX_LANES} notification lanes of the receiver are taken")

    def notif_head_ptr(self, sender_agent_name: str) -> int:
        """Receiver side: device address of the u32 that counts the messages `sender_agent_name` has delivered here (0 if
        that sender has not added this agent yet). A stream can wait on it with kvdp_stream_wait_flag."""
        claims = np.zeros(MAILBOX_LANES, dtype=np.uint64)
        _lib.check(_lib.lib().cuda_memcpy_sync(C.c_void_p(claims.ctypes.data), C.c_void_p(self._mailbox), CLAIMS_BYTES))
        key = _agent_key(sender_agent_name)
        for probe in range(MAILBOX_LANES):
            lane = (key + probe) % MAILBOX_LANES
            if int(claims[lane]) == key:
                return self._mailbox + _head_off(lane)
            if int(claims[lane]) == 0:
                return 0
        return 0

    def _stage_slot(self) -> int:
        """Next pinned staging slot. A slot is free again once the transfer whose stream position covers its upload has
        completed, which the pinned progress word shows without any stream synchronisation."""
        slot = self._stage_n % STAGE_SLOTS
        self._stage_n += 1
        need = self._stage_cover[slot]
        if need and int(self._progress[0]) < need:
            import time
            deadline = time.monotonic() + self.notify_timeout_s
            while int(self._progress[0]) < need:
                if time.monotonic() > deadline:
                    raise KvError("notification staging ring did not drain (stream stalled?)")
                time.sleep(0)
        self._stage_cover[slot] = self._xfer_seq + 1     # the signal the current write call enqueues last
        return slot

    def _wait_for_room(self, key, rem: _Remote, target_rank: int, seq: int) -> None:
        """Flow control: never overwrite a ring slot the receiver has not read. The receiver publishes how many messages
        it has consumed per lane; it is re-read only when the cached value says the ring could be full."""
        if seq - self._peer_consumed.get(key, 0) <= MAILBOX_SLOTS:
            return
        import time
        L = _lib.lib()
        word = np.zeros(1, dtype=np.uint32)
        addr = rem.mailbox[target_rank] + CONSUMED_OFF + rem.lane[target_rank] * 4
        deadline = time.monotonic() + self.notify_timeout_s
        while True:
            _lib.check(L.cuda_memcpy_sync(C.c_void_p(word.ctypes.data), C.c_void_p(addr), 4))
            self._peer_consumed[key] = int(word[0])
            if seq - int(word[0]) <= MAILBOX_SLOTS:
                return
            if time.monotonic() > deadline:
                raise KvError(f"{key[0]}:{target_rank} has {MAILBOX_SLOTS} unread notifications from {self.agent_name}: "
                              "it must call get_new_notifs()")
            time.sleep(0.0005)

    def _post(self, engine_id: str, target_rank: int, rem: _Remote, msg: str) -> None:
        data = msg.encode("utf-8")
        if len(data) > MAX_MSG:
            raise KvError(f"notify message longer than {MAX_MSG} bytes")
        name = self.agent_name.encode("utf-8")[:NAME_BYTES]
        key = (engine_id, target_rank)
        seq = self._send_seq.get(key, 0) + 1
        self._wait_for_room(key, rem, target_rank, seq)
        self._send_seq[key] = seq
        slot = (seq - 1) % MAILBOX_SLOTS
        stage = self._stage_slot()
        rec = np.frombuffer((C.c_uint8 * SLOT_BYTES).from_address(self._host + stage * SLOT_BYTES), dtype=np.uint8)
        rec[:] = 0
        rec[:16].view(np.uint32)[:] = (len(data), seq, len(name), 0)
        rec[16:16 + len(name)] = np.frombuffer(name, dtype=np.uint8)
        rec[16 + NAME_BYTES:16 + NAME_BYTES + len(data)] = np.frombuffer(data, dtype=np.uint8)
        L = _lib.lib()
        mailbox_ptr, lane = rem.mailbox[target_rank], rem.lane[target_rank]
        _lib.check(L.cuda_memcpy_async(C.c_void_p(mailbox_ptr + _slot_off(lane, slot)),
                                       C.c_void_p(self._host + stage * SLOT_BYTES), SLOT_BYTES,
                                       C.c_void_p(self._engine.stream)))
        # release store of the lane head, ordered after data and message
        self._engine.signal(mailbox_ptr + _head_off(lane), seq)

    def get_done_tranfers(self) -> List[str]:
        """Sender side: request ids whose transfer (data + notification) has completed (reference: patch:1473-1492)."""
        done_upto = int(self._progress[0])
        done = [m for m, s in self._transfers.items() if s <= done_upto]
        for m in done:
            del self._transfers[m]
        return done

    def get_new_notifs(self) -> Dict[str, List[str]]:
        """Receiver side: messages that arrived since the last call, keyed by the sender's agent name
        (`<engine_id>:<rank>`), as NIXL's get_new_notifs returns {agent_name: [notifs]} (reference: patch:1421-1425)."""
        L = _lib.lib()
        out: Dict[str, List[str]] = {}
        heads = np.zeros(HEADS_BYTES // 4, dtype=np.uint32)
        _lib.check(L.cuda_memcpy_sync(C.c_void_p(heads.ctypes.data), C.c_void_p(self._mailbox + HEADS_OFF),
                                      HEADS_BYTES))
        lane_heads = heads[::16]
        for lane in range(MAILBOX_LANES):
            head = int(lane_heads[lane])
            seen = self._seen[lane]
            if head <= seen:
                continue
            if head - seen > MAILBOX_SLOTS:
                raise KvError(f"notification lane {lane} overran: {head - seen} unread messages in a ring of "
                              f"{MAILBOX_SLOTS} (a sender ignored the consumed counter)")
            ring = np.zeros(MAILBOX_SLOTS * SLOT_BYTES, dtype=np.uint8)
            _lib.check(L.cuda_memcpy_sync(C.c_void_p(ring.ctypes.data), C.c_void_p(self._mailbox + _slot_off(lane, 0)),
                                          ring.nbytes))
            for seq in range(seen + 1, head + 1):
                rec = ring[((seq - 1) % MAILBOX_SLOTS) * SLOT_BYTES:][:SLOT_BYTES]
                n, s, nn, _ = [int(x) for x in rec[:16].view(np.uint32)]
                if s != seq:
                    raise KvError(f"notification lane {lane}: slot holds message {s}, expected {seq}")
                sender = bytes(rec[16:16 + nn]).decode("utf-8")
                out.setdefault(sender, []).append(bytes(rec[16 + NAME_BYTES:16 + NAME_BYTES + n]).decode("utf-8"))
            self._seen[lane] = head
            word = np.array([head], dtype=np.uint32)      # lets the sender reuse the slots just read
            _lib.check(L.cuda_memcpy_sync(C.c_void_p(self._mailbox + CONSUMED_OFF + lane * 4),
                                          C.c_void_p(word.ctypes.data), 4))
        return out

    def sync(self) -> None:
        self._engine.sync()

    def _remote(self, engine_id: str) -> _Remote:
        if engine_id not in self._remotes:
            raise KvError(f"unknown remote engine {engine_id}: call add_remote_agent first")
        return self._remotes[engine_id]

    def shutdown(self) -> None:
        """reference: patch:1188-1199."""
        L = _lib.lib()
        try:
            self._engine.sync()
        except Exception:
            pass
        for rem in self._remotes.values():
            if rem.push:
                rem.push.close()
            for p in rem.pull:
                p.close()
            for p in getattr(rem, "_mla_push", []) + (getattr(rem, "_mla_push_tokens", None) or []):
                p.close()
            for ptr in rem.mapped:
                L.kvdp_ipc_close(C.c_void_p(ptr))
        self._remotes.clear()
        self._engine.close()
        if self._mailbox:
            L.kvdp_device_free(C.c_void_p(self._mailbox))
            self._mailbox = 0
        if self._host:
            self._progress = None
            L.cuda_free_host(C.c_void_p(self._host))
            self._host = 0
