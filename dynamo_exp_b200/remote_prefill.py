"""The request / completion protocol around the transfer connector, mirroring the reference's
`vllm/remote_prefill.py` and the engine / worker code that drives `DynamoNixlConnector`
(container/deps/vllm/vllm_v0.8.4-dynamo-kv-disagg-patch.patch):

    RemotePrefillRequest, MemoryOpType, MemoryTransferRequest, RemotePrefillParams      patch:4199-4256
    plan_memory_transfers     what LLMEngine.step builds after a remote-decode prefill   patch:3188-3221
    execute_memory_transfers  Worker._read_blocks / _write_blocks                        patch:4528-4539
    collect_transfer_counters the per-step notification / done tallies of one worker     patch:4779-4800
    TransferTracker           LLMEngine's _request_notif_counter / _request_done_counter patch:2991-2992, 3251-3261

Wire format: the reference serialises these structs with msgspec (`MemoryTransferRequest` is `array_like`); here
`encode()` / `decode()` produce the same msgpack layouts with the `msgpack` package. Sampling parameters are an opaque
dict: they belong to the engine, not to the data plane.
"""
from __future__ import annotations

from collections import defaultdict
from dataclasses import dataclass, field
from enum import Enum
from typing import Callable, Dict, Iterable, List, Optional, Sequence, Set, Tuple

import msgpack


class MemoryOpType(str, Enum):
    WRITE = "WRITE"
    READ = "READ"


@dataclass
class RemotePrefillRequest:
    """What a decode engine sends to the prefill queue (patch:4199-4220); a msgpack map, defaults omitted."""
    engine_id: str
    request_id: str
    prompt_token_ids: List[int]
    sampling_params: dict
    block_ids: List[int]
    computed_block_ids: List[int]

    def encode(self) -> bytes:
        return msgpack.packb({"engine_id": self.engine_id, "request_id": self.request_id,
                              "prompt_token_ids": list(self.prompt_token_ids), "sampling_params": self.sampling_params,
                              "block_ids": list(self.block_ids), "computed_block_ids": list(self.computed_block_ids)})

    @staticmethod
    def decode(raw: bytes) -> "RemotePrefillRequest":
        d = msgpack.unpackb(raw)
        return RemotePrefillRequest(d["engine_id"], d["request_id"], list(d["prompt_token_ids"]),
                                    d.get("sampling_params", {}), list(d["block_ids"]),
                                    list(d["computed_block_ids"]))


@dataclass
class MemoryTransferRequest:
    """One block-list transfer a worker must run in a step (patch:4228-4245); a msgpack array in field order."""
    request_id: str
    local_block_ids: List[int]
    staging_block_ids: List[int]
    remote_block_ids: List[int]
    remote_engine_id: str
    notify_msg: str
    op_type: MemoryOpType

    def encode(self) -> bytes:
        return msgpack.packb([self.request_id, list(self.local_block_ids), list(self.staging_block_ids),
                              list(self.remote_block_ids), self.remote_engine_id, self.notify_msg,
                              self.op_type.value])

    @staticmethod
    def decode(raw: bytes) -> "MemoryTransferRequest":
        a = msgpack.unpackb(raw)
        return MemoryTransferRequest(a[0], list(a[1]), list(a[2]), list(a[3]), a[4], a[5], MemoryOpType(a[6]))


RemotePrefillRequestCallback = Callable[[RemotePrefillRequest], None]


@dataclass
class RemotePrefillParams:
    """patch:4250-4256."""
    is_remote_prefill: bool = False
    is_remote_decode: bool = False
    decode_block_ids: Optional[List[int]] = None
    decode_computed_block_ids: Optional[List[int]] = None
    decode_engine_id: Optional[str] = None
    remote_prefill_request_callback: Optional[RemotePrefillRequestCallback] = None


def make_remote_prefill_request(engine_id: str, request_id: str, prompt_token_ids: Sequence[int], sampling_params: dict,
                                block_table: Sequence[int], computed_block_nums: Sequence[int]
                                ) -> Optional[RemotePrefillRequest]:
    """Decode side, before model execution (patch:3159-3177): nothing to ask for when every block is already computed
    locally (the caller then marks the prefill as finished)."""
    if len(block_table) == len(computed_block_nums):
        return None
    return RemotePrefillRequest(engine_id, request_id, list(prompt_token_ids), sampling_params, list(block_table),
                                list(computed_block_nums))


def plan_memory_transfers(request_id: str, block_table: Sequence[int], staging_block_ids: Sequence[int],
                          computed_block_nums: Sequence[int], params: RemotePrefillParams
                          ) -> List[MemoryTransferRequest]:
    """Prefill side, after model execution of a remote-decode request (patch:3188-3221): blocks the prefill engine
    found in its own prefix cache are not recomputed; the decode engine's copies of them are READ first (so that the
    prefill runs on identical KV), everything after them is WRITTEN to the decode engine's blocks.

    `staging_block_ids` is the reference's shadow sequence (`block_tables[seq_id + 1]`); this data plane needs no
    staging, the list is carried for interface parity and may be empty."""
    if not params.is_remote_decode:
        return []
    n = len(computed_block_nums)
    decode_ids = list(params.decode_block_ids or [])
    staging = list(staging_block_ids)
    out: List[MemoryTransferRequest] = []
    computed_decode = decode_ids[:n]
    if computed_decode:
        out.append(MemoryTransferRequest(request_id, list(block_table[:n]), staging[:n], computed_decode,
                                         params.decode_engine_id or "", request_id, MemoryOpType.READ))
    out.append(MemoryTransferRequest(request_id, list(block_table[n:]), staging[n:], decode_ids[n:],
                                     params.decode_engine_id or "", request_id, MemoryOpType.WRITE))
    return out


def execute_memory_transfers(connector, requests: Iterable[MemoryTransferRequest],
                             run_model: Optional[Callable[[], None]] = None) -> None:
    """One worker step (patch:4528-4539, 4691, 4766): READs before the model runs, WRITEs after it."""
    requests = list(requests)
    for r in requests:
        if r.op_type == MemoryOpType.READ:
            connector.read_blocks(r.local_block_ids, r.staging_block_ids, r.remote_block_ids, r.remote_engine_id)
    if run_model is not None:
        run_model()
    for r in requests:
        if r.op_type == MemoryOpType.WRITE:
            connector.write_blocks(r.local_block_ids, r.staging_block_ids, r.remote_block_ids, r.remote_engine_id,
                                   r.notify_msg)


def collect_transfer_counters(connectors: Sequence) -> Tuple[Dict[str, int], Dict[str, int]]:
    """What the driver worker returns next to the model output (patch:4779-4800): how many notifications arrived per
    request over all TP workers' connectors, and how many of this engine's own transfers completed."""
    notif: Dict[str, int] = defaultdict(int)
    done: Dict[str, int] = defaultdict(int)
    for k, c in enumerate(connectors):
        for req_ids in c.get_new_notifs().values():
            for req_id in req_ids:
                notif[req_id.decode("utf-8") if isinstance(req_id, bytes) else req_id] += 1
        if k == 0:                       # the reference counts done transfers on the driver worker only
            for req_id in c.get_done_tranfers():
                done[req_id] += 1
    return dict(notif), dict(done)


@dataclass
class TransferTracker:
    """Engine-side completion bookkeeping (patch:2991-2992, 3251-3261). Every request starts at -tensor_parallel_size;
    it is finished once one count per TP worker has arrived."""
    tensor_parallel_size: int
    finished_prefills: Set[str] = field(default_factory=set)
    finished_transfers: Set[str] = field(default_factory=set)

    def __post_init__(self):
        self._notif = defaultdict(lambda: -self.tensor_parallel_size)
        self._done = defaultdict(lambda: -self.tensor_parallel_size)

    def update(self, request_notif_counter: Dict[str, int], request_done_counter: Dict[str, int]) -> None:
        for req_id, count in request_notif_counter.items():
            self._notif[req_id] += count
            if self._notif[req_id] > -1:
                self.finished_prefills.add(req_id)
                del self._notif[req_id]
        for req_id, count in request_done_counter.items():
            self._done[req_id] += count
            if self._done[req_id] > -1:
                self.finished_transfers.add(req_id)
                del self._done[req_id]

    def pending(self) -> Tuple[Dict[str, int], Dict[str, int]]:
        return dict(self._notif), dict(self._done)
