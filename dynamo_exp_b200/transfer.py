"""V2 host API over libkvdp.so: copy plans, the transfer engine, contiguous-per-block slabs and the host tier.

These are the batched counterparts of `CopyStreamBlockMap` / `CopyStream` (lib/llm/src/kv/layer.rs:592-979): one
plan per (source set, destination set), one kernel launch per request covering all layers. The TP-reshard plan and the
peer/flag helpers replace what the patched vLLM connector does with Triton + NIXL
(container/deps/vllm/vllm_v0.8.4-dynamo-kv-disagg-patch.patch:934-1031, :1131-1186, :1361-1471).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from . import _lib
from .layer import KvBlockStorage, _ids32
from .storage import CudaPinnedMemory, DeviceStorageOwned, KvError, Storage, StorageType


def _u64(ptrs) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(ptrs, dtype=np.uint64))


class CopyPlan:
    """Owner of a `kvdp_plan*`."""

    def __init__(self, handle: C.c_void_p, keepalive=None):
        self._h = handle
        self._keep = keepalive

    @property
    def handle(self) -> C.c_void_p:
        return self._h

    def bytes_per_pair(self) -> int:
        return int(_lib.lib().kvdp_plan_bytes_per_pair(self._h))

    def close(self) -> None:
        if self._h is not None and self._h.value:
            _lib.lib().kvdp_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- constructors -------------------------------------------------------------------------------------
    @classmethod
    def block_copy(cls, src: KvBlockStorage, dst: KvBlockStorage, device: int = -1) -> "CopyPlan":
        """Same-layout multi-layer block copy; the v2 form of `CopyStreamBlockMap::new` (kv/layer.rs:648-688)."""
        if not src.block_details.is_compatible(dst.block_details):
            raise KvError("src and dst must have compatible block details")
        if len(src.layers) != len(dst.layers):
            raise KvError("src and dst must have the same number of layers")
        sp, dp = _u64(src.layer_pointers()), _u64(dst.layer_pointers())
        h = C.c_void_p()
        bd = src.block_details
        _lib.check(_lib.lib().kvdp_plan_block_copy(C.byref(h), sp.ctypes.data, dp.ctypes.data, len(sp), bd.prefix_dim(),
                                                   src.number_of_blocks(), dst.number_of_blocks(), bd.suffix_dim(),
                                                   bd.elem_size(), device))
        return cls(h, (src, dst))

    @classmethod
    def block_copy_raw(cls, src_ptrs, dst_ptrs, prefix_dim, src_block_dim, dst_block_dim, suffix_dim, elem_size,
                       device: int = -1, keepalive=None) -> "CopyPlan":
        sp, dp = _u64(src_ptrs), _u64(dst_ptrs)
        if len(sp) != len(dp):
            raise KvError("src and dst must have the same number of layers")
        h = C.c_void_p()
        _lib.check(_lib.lib().kvdp_plan_block_copy(C.byref(h), sp.ctypes.data, dp.ctypes.data, len(sp), prefix_dim,
                                                   src_block_dim, dst_block_dim, suffix_dim, elem_size, device))
        return cls(h, keepalive)

    @classmethod
    def contig(cls, layers: KvBlockStorage, contig: "KvContiguousBlocks", to_contig: bool,
               device: int = -1) -> "CopyPlan":
        """layer-separate <-> contiguous-per-block relayout."""
        bd = layers.block_details
        chunk = bd.bytes_per_token_block_per_layer() // 2
        if contig.chunk_bytes != chunk or contig.num_layers != len(layers.layers):
            raise KvError("contiguous slab geometry does not match the layer set")
        lp = _u64(layers.layer_pointers())
        h = C.c_void_p()
        _lib.check(_lib.lib().kvdp_plan_contig(C.byref(h), lp.ctypes.data, len(lp), bd.layout.value,
                                               layers.number_of_blocks(), chunk, C.c_void_p(contig.get_pointer()),
                                               contig.number_of_blocks, int(to_contig), device))
        return cls(h, (layers, contig))

    @classmethod
    def reshard(cls, src_layer_ptrs, dst_layer_ptrs_by_target, src_blocks: int, dst_blocks: int, block_size: int,
                src_row_bytes: int, src_tp: int, dst_tp: int, src_rank: int, device: int = -1,
                keepalive=None) -> "CopyPlan":
        """TP reshard push (KvFirst caches). dst_layer_ptrs_by_target: [targets][layers] raw pointers
        (local or peer-mapped)."""
        sp = _u64(src_layer_ptrs)
        dp = _u64(np.asarray(dst_layer_ptrs_by_target, dtype=np.uint64).reshape(-1))
        h = C.c_void_p()
        _lib.check(_lib.lib().kvdp_plan_reshard(C.byref(h), sp.ctypes.data, dp.ctypes.data, len(sp), src_blocks,
                                                dst_blocks, block_size, src_row_bytes, src_tp, dst_tp, src_rank,
                                                device))
        return cls(h, keepalive)

    @classmethod
    def from_desc(cls, desc: "_lib.PlanDesc", keepalive=None) -> "CopyPlan":
        h = C.c_void_p()
        _lib.check(_lib.lib().kvdp_plan_create(C.byref(h), C.byref(desc)))
        return cls(h, keepalive)


class TransferEngine:
    """Owner of a `kvdp_xfer*`: one stream + id staging; the v2 `CopyStream`."""

    def __init__(self, max_pairs: int, device: int = -1):
        _lib.require_gpu()
        h = C.c_void_p()
        _lib.check(_lib.lib().kvdp_xfer_create(C.byref(h), int(max_pairs), int(device)))
        self._h = h

    @property
    def stream(self) -> int:
        return int(_lib.lib().kvdp_xfer_stream(self._h) or 0)

    def run(self, plan: CopyPlan, src_ids: Sequence[int], dst_ids: Sequence[int], layer_begin: int = 0,
            layer_end: Optional[int] = None, check_dst_unique: bool = True, done_flag: int = 0,
            done_value: int = 0, num_layers: Optional[int] = None) -> None:
        if len(src_ids) != len(dst_ids):
            raise KvError("src_block_ids and dst_block_ids must have the same length")
        s, d = _ids32(src_ids), _ids32(dst_ids)
        if layer_end is None:
            if num_layers is None:
                raise KvError("layer_end or num_layers is required")
            layer_end = num_layers
        flags = _lib.XFER_CHECK_DST_UNIQUE if check_dst_unique else 0
        _lib.check(_lib.lib().kvdp_xfer_run(self._h, plan.handle, s.ctypes.data if len(s) else None,
                                            d.ctypes.data if len(d) else None, len(s), int(layer_begin),
                                            int(layer_end), flags, C.c_void_p(done_flag) if done_flag else None,
                                            int(done_value)))

    def signal(self, flag_ptr: int, value: int) -> None:
        _lib.check(_lib.lib().kvdp_xfer_signal(self._h, C.c_void_p(flag_ptr), int(value)))

    def sync(self) -> None:
        _lib.check(_lib.lib().kvdp_xfer_sync(self._h))

    def close(self) -> None:
        if self._h is not None and self._h.value:
            _lib.lib().kvdp_xfer_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class KvContiguousBlocks(Storage):
    """Fully-contiguous-per-block slab `[block][layer][K|V][token][head][dim]` (not in the reference)."""

    def __init__(self, number_of_blocks: int, num_layers: int, chunk_bytes: int, storage_type: StorageType):
        self.number_of_blocks = int(number_of_blocks)
        self.num_layers = int(num_layers)
        self.chunk_bytes = int(chunk_bytes)
        self.block_bytes = self.num_layers * 2 * self.chunk_bytes
        nbytes = self.number_of_blocks * self.block_bytes
        if storage_type.kind == "device":
            self.storage: Storage = DeviceStorageOwned(nbytes, storage_type.device)
        elif storage_type.kind == "pinned":
            self.storage = CudaPinnedMemory(nbytes)
        else:
            raise KvError("System memory not yet supported")

    def get_pointer(self) -> int:
        return self.storage.get_pointer()

    def storage_size(self) -> int:
        return self.storage.storage_size()

    def storage_type(self) -> StorageType:
        return self.storage.storage_type()


class HostTier:
    """Offload/onboard between layer-separate HBM blocks and a pinned contiguous-per-block slab."""

    def __init__(self, layers: KvBlockStorage, host: KvContiguousBlocks, staging_blocks: int = 32, device: int = -1):
        if host.storage_type().kind != "pinned":
            raise KvError("the host tier slab must be pinned memory")
        bd = layers.block_details
        chunk = bd.bytes_per_token_block_per_layer() // 2
        if host.chunk_bytes != chunk or host.num_layers != len(layers.layers):
            raise KvError("host slab geometry does not match the layer set")
        lp = _u64(layers.layer_pointers())
        h = C.c_void_p()
        _lib.check(_lib.lib().kvdp_host_tier_create(C.byref(h), lp.ctypes.data, len(lp), bd.layout.value,
                                                    layers.number_of_blocks(), chunk, C.c_void_p(host.get_pointer()),
                                                    host.number_of_blocks, int(staging_blocks), device))
        self._h = h
        self._keep = (layers, host)

    def offload(self, device_ids: Sequence[int], host_ids: Sequence[int]) -> None:
        if len(device_ids) != len(host_ids):
            raise KvError("device and host id lists must have the same length")
        d, h = _ids32(device_ids), _ids32(host_ids)
        _lib.check(_lib.lib().kvdp_host_tier_offload(self._h, d.ctypes.data, h.ctypes.data, len(d)))

    def onboard(self, host_ids: Sequence[int], device_ids: Sequence[int]) -> None:
        if len(device_ids) != len(host_ids):
            raise KvError("device and host id lists must have the same length")
        d, h = _ids32(device_ids), _ids32(host_ids)
        _lib.check(_lib.lib().kvdp_host_tier_onboard(self._h, h.ctypes.data, d.ctypes.data, len(d)))

    def sync(self) -> None:
        _lib.check(_lib.lib().kvdp_host_tier_sync(self._h))

    def close(self) -> None:
        if self._h is not None and self._h.value:
            _lib.lib().kvdp_host_tier_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---------------------------------------------------------------------------------------------- token-granular paths
def reshape_and_cache(key_ptr: int, value_ptr: int, key_cache_ptr: int, value_cache_ptr: int, slot_mapping,
                      num_slots: int, row_bytes: int, engine: Optional[TransferEngine] = None,
                      device: int = -1) -> None:
    """Token-granular paged write: `cache[slot_mapping[i]] = new[i]` for K and V in ONE launch — the data movement of
    vLLM's `reshape_and_cache_flash`, which the reference's NCCL connector calls on every received layer
    (container/deps/vllm/vllm_v0.8.4-dynamo-kv-disagg-patch.patch:1721-1746). `key`/`value` are `[num_tokens, H, D]`
    contiguous, the caches `[num_blocks * block_size, H, D]` (flash layout viewed per slot), `row_bytes = H*D*elem`,
    `slot = block * block_size + offset`; negative slots (padding) are skipped, as there."""
    slots = np.asarray(slot_mapping, dtype=np.int64)
    keep = np.nonzero(slots >= 0)[0]
    if keep.size == 0:
        return
    if np.any(slots[keep] >= num_slots):
        raise KvError("slot_mapping entry beyond the cache")
    plan = CopyPlan.block_copy_raw([key_ptr, value_ptr], [key_cache_ptr, value_cache_ptr], 1, int(slots.size),
                                   int(num_slots), int(row_bytes), 1, device)
    own = engine is None
    eng = engine or TransferEngine(int(keep.size), device)
    try:
        eng.run(plan, keep.astype(np.int32), slots[keep].astype(np.int32), num_layers=2)
        eng.sync()
    finally:
        plan.close()
        if own:
            eng.close()


def gather_token_rows(cache_ptr: int, out_ptr: int, slot_mapping, num_slots: int, row_bytes: int, col_offset: int = 0,
                      col_bytes: Optional[int] = None, engine: Optional[TransferEngine] = None,
                      device: int = -1) -> None:
    """Token-granular gather with an optional head slice: `out[i] = cache[slot_mapping[i]][col_offset : +col_bytes]`,
    what the reference's NCCL send path does per layer with torch indexing
    (`cache.reshape(-1, H, D)[slot_mapping, h0:h1]`, patch:1602-1672). One launch, no intermediate tensor."""
    slots = np.asarray(slot_mapping, dtype=np.int64)
    if slots.size == 0:
        return
    if np.any(slots < 0) or np.any(slots >= num_slots):
        raise KvError("slot_mapping entry outside the cache")
    col_bytes = row_bytes - col_offset if col_bytes is None else col_bytes
    if col_offset < 0 or col_bytes < 1 or col_offset + col_bytes > row_bytes:
        raise KvError("column range outside the row")
    src = np.array([cache_ptr], dtype=np.uint64)
    dst = np.array([out_ptr], dtype=np.uint64)
    d = _lib.PlanDesc()
    d.num_layers, d.row_groups, d.prefix_dim, d.rows = 1, 1, 1, 1
    d.run_bytes = col_bytes
    d.src_block_stride, d.src_row_stride, d.src_col_offset = row_bytes, row_bytes, col_offset
    d.dst_block_stride, d.dst_row_stride = col_bytes, col_bytes
    d.src_blocks, d.dst_blocks = int(num_slots), int(slots.size)
    d.src_layer_ptrs, d.dst_layer_ptrs, d.device = src.ctypes.data, dst.ctypes.data, device
    plan = CopyPlan.from_desc(d, keepalive=(src, dst))
    own = engine is None
    eng = engine or TransferEngine(int(slots.size), device)
    try:
        eng.run(plan, slots.astype(np.int32), np.arange(slots.size, dtype=np.int32), num_layers=1)
        eng.sync()
    finally:
        plan.close()
        if own:
            eng.close()


def token_rows_plan(src_layer_ptrs: Sequence[int], dst_layer_ptrs: Sequence[int], src_num_slots: int,
                    dst_num_slots: int, src_row_bytes: int, dst_row_bytes: int, col_bytes: int, src_col_offset: int = 0,
                    dst_col_offset: int = 0, device: int = -1) -> CopyPlan:
    """The copy plan behind `transfer_token_rows`: token slots play the part of blocks, K|V is the prefix, one run of
    `col_bytes` per (layer, K|V, token)."""
    src, dst = _u64(src_layer_ptrs), _u64(dst_layer_ptrs)
    d = _lib.PlanDesc()
    d.num_layers, d.row_groups, d.prefix_dim, d.rows = len(src), 1, 2, 1
    d.run_bytes = col_bytes
    d.src_prefix_stride, d.src_block_stride = int(src_num_slots) * src_row_bytes, src_row_bytes
    if src_col_offset:                      # a head slice of the source row; whole rows are plain contiguous runs
        d.src_row_stride, d.src_col_offset = src_row_bytes, src_col_offset
    d.dst_prefix_stride, d.dst_block_stride = int(dst_num_slots) * dst_row_bytes, dst_row_bytes
    d.dst_row_stride, d.dst_col_offset = dst_row_bytes, dst_col_offset
    d.src_blocks, d.dst_blocks = int(src_num_slots), int(dst_num_slots)
    d.src_layer_ptrs, d.dst_layer_ptrs, d.device = src.ctypes.data, dst.ctypes.data, device
    return CopyPlan.from_desc(d, keepalive=(src, dst))


def transfer_token_rows(src_layer_ptrs: Sequence[int], dst_layer_ptrs: Sequence[int], src_slots, dst_slots,
                        src_num_slots: int, dst_num_slots: int, src_row_bytes: int, dst_row_bytes: int,
                        col_bytes: Optional[int] = None, src_col_offset: int = 0, dst_col_offset: int = 0,
                        engine: Optional[TransferEngine] = None, device: int = -1) -> None:
    """Token-granular prefill -> decode transfer in ONE launch over all layers, K and V:

        dst[l][kv][dst_slots[i]][dst_col_offset : +col_bytes] = src[l][kv][src_slots[i]][src_col_offset : +col_bytes]

    for caches laid out `[2, num_slots, row_bytes]` per layer (vLLM's `[2, num_blocks, block_size, H, D]` viewed per
    token slot). This is the whole per-request data path of the reference's NCCL connector — per layer a torch gather
    `cache.reshape(-1, H, D)[slot_mapping, h0:h1]`, a `cat`, three NCCL sends and, on the other side,
    `reshape_and_cache_flash` (patch:1602-1672, 1676-1766) — with the head slice (TP resharding) folded into the column
    offsets. The destination pointers may be peer-mapped (kvdp_ipc_import): the stores then go over NVLink."""
    s_slots = np.asarray(src_slots, dtype=np.int64)
    d_slots = np.asarray(dst_slots, dtype=np.int64)
    if s_slots.shape != d_slots.shape:
        raise KvError("src_slots and dst_slots must have the same length")
    if s_slots.size == 0:
        return
    if np.any(s_slots < 0) or np.any(s_slots >= src_num_slots) or np.any(d_slots < 0) or np.any(d_slots >= dst_num_slots):
        raise KvError("slot outside the cache")
    if len(src_layer_ptrs) != len(dst_layer_ptrs) or not len(src_layer_ptrs):
        raise KvError("need one source and one destination pointer per layer")
    col_bytes = min(src_row_bytes - src_col_offset, dst_row_bytes - dst_col_offset) if col_bytes is None else col_bytes
    if (col_bytes < 1 or src_col_offset < 0 or dst_col_offset < 0 or src_col_offset + col_bytes > src_row_bytes
            or dst_col_offset + col_bytes > dst_row_bytes):
        raise KvError("column range outside the row")
    plan = token_rows_plan(src_layer_ptrs, dst_layer_ptrs, src_num_slots, dst_num_slots, src_row_bytes, dst_row_bytes,
                           col_bytes, src_col_offset, dst_col_offset, device)
    own = engine is None
    eng = engine or TransferEngine(int(s_slots.size), device)
    try:
        eng.run(plan, s_slots.astype(np.int32), d_slots.astype(np.int32), num_layers=len(src_layer_ptrs),
                check_dst_unique=True)
        eng.sync()
    finally:
        plan.close()
        if own:
            eng.close()


# ---------------------------------------------------------------------------------------------- raw memory helpers
def memcpy_sync(dst: int, src: int, nbytes: int) -> None:
    _lib.check(_lib.lib().cuda_memcpy_sync(C.c_void_p(dst), C.c_void_p(src), int(nbytes)))


def upload(dst_device_ptr: int, array: np.ndarray) -> None:
    a = np.ascontiguousarray(array)
    memcpy_sync(dst_device_ptr, a.ctypes.data, a.nbytes)


def download(src_device_ptr: int, nbytes: int, dtype=np.uint8) -> np.ndarray:
    out = np.empty(int(nbytes), dtype=np.uint8)
    memcpy_sync(out.ctypes.data, src_device_ptr, nbytes)
    return out.view(dtype)
