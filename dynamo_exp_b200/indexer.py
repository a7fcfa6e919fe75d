"""Host-side mirror of the KV router's indexer (lib/llm/src/kv_router/indexer.rs, protocols.rs) over libkvdp.so.

`RadixTree` / `KvIndexer` keep the reference's method names (apply_event, find_matches, find_matches_for_request,
remove_worker); the data types (`RouterEvent`, `KvCacheEvent`, `KvCacheStoreData`, ...) serialise to the same JSON the
reference's serde derives produce (protocols.rs:70-121: `{"stored": {...}}` / `{"removed": {...}}`, hashes as plain
u64 numbers). New here: `find_matches_batch`, which scores a whole batch of requests in one GPU launch, and
`find_matches_for_requests`, which also hashes the token blocks on the GPU first (kvdp_hash_token_blocks, seed 1337).
"""
from __future__ import annotations

import ctypes as C
import json
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Union

import numpy as np

from . import _lib
from .storage import KvError
from .tokens import _as_u32

WorkerId = int
MAX_WORKERS = 64            # default width of one index; RadixTree(max_workers=...) goes up to 512
MAX_WORKERS_LIMIT = 512


# ---------------------------------------------------------------------------------------------- protocol types
@dataclass
class KvCacheStoredBlockData:            # protocols.rs:107-113
    block_hash: int                      # ExternalSequenceBlockHash
    tokens_hash: int                     # LocalBlockHash


@dataclass
class KvCacheStoreData:                  # protocols.rs:98-104
    parent_hash: Optional[int]
    blocks: List[KvCacheStoredBlockData]


@dataclass
class KvCacheRemoveData:                 # protocols.rs:116-120
    block_hashes: List[int]


@dataclass
class KvCacheEvent:                      # protocols.rs:79-85; data is `Stored(..)` or `Removed(..)`
    event_id: int
    data: Union[KvCacheStoreData, KvCacheRemoveData]

    def to_dict(self) -> dict:
        if isinstance(self.data, KvCacheStoreData):
            d = {"stored": {"parent_hash": self.data.parent_hash,
                            "blocks": [{"block_hash": b.block_hash, "tokens_hash": b.tokens_hash}
                                       for b in self.data.blocks]}}
        else:
            d = {"removed": {"block_hashes": list(self.data.block_hashes)}}
        return {"event_id": self.event_id, "data": d}

    @staticmethod
    def from_dict(d: dict) -> "KvCacheEvent":
        data = d["data"]
        if "stored" in data:
            s = data["stored"]
            return KvCacheEvent(d["event_id"], KvCacheStoreData(
                s.get("parent_hash"), [KvCacheStoredBlockData(b["block_hash"], b["tokens_hash"]) for b in s["blocks"]]))
        if "removed" in data:
            return KvCacheEvent(d["event_id"], KvCacheRemoveData(list(data["removed"]["block_hashes"])))
        raise KvError("KvCacheEventData must be `stored` or `removed`")


@dataclass
class RouterEvent:                       # indexer.rs:137-160
    worker_id: WorkerId
    event: KvCacheEvent

    def to_json(self) -> str:
        return json.dumps({"worker_id": self.worker_id, "event": self.event.to_dict()}, separators=(",", ":"))

    @staticmethod
    def from_json(s: str) -> "RouterEvent":
        d = json.loads(s)
        return RouterEvent(d["worker_id"], KvCacheEvent.from_dict(d["event"]))


@dataclass
class KvCacheEvents:                     # protocols.rs:70-76
    events: List[KvCacheEvent]
    shutdown: bool = False

    def to_json(self) -> str:
        return json.dumps({"events": [e.to_dict() for e in self.events], "shutdown": self.shutdown},
                          separators=(",", ":"))

    @staticmethod
    def from_json(s: str) -> "KvCacheEvents":
        d = json.loads(s)
        return KvCacheEvents([KvCacheEvent.from_dict(e) for e in d["events"]], bool(d["shutdown"]))


@dataclass
class OverlapScores:                     # indexer.rs:390-436
    scores: Dict[WorkerId, int] = field(default_factory=dict)
    frequencies: List[int] = field(default_factory=list)


# ---------------------------------------------------------------------------------------------- the index
class RadixTree:
    """reference: indexer.rs:187-387. Events are applied on the host inside libkvdp.so; lookups run on the GPU."""

    def __init__(self, max_nodes: int = 1 << 20, device: int = -1, expiration_duration: Optional[float] = None,
                 max_workers: int = MAX_WORKERS):
        """`expiration_duration` (seconds) is `RadixTree::new_with_frequency` (indexer.rs:217-223): lookups then also
        report how often each matched block was visited inside that window. `max_workers` (<= 512): how many workers
        the index can hold at a time; up to 64 every walk step is one dependent read, above that one 64-worker word
        of the set is read per step and pass (the reference's set is an unbounded HashSet, indexer.rs:163-170)."""
        _lib.require_gpu()
        if not 1 <= int(max_workers) <= MAX_WORKERS_LIMIT:
            raise KvError(f"max_workers must be 1..{MAX_WORKERS_LIMIT}")
        self.expiration_duration = expiration_duration
        self.max_workers = int(max_workers)
        h = C.c_void_p()
        _lib.check(_lib.lib().kvdp_index_create(C.byref(h), int(max_nodes), self.max_workers, int(device)))
        self._h = h
        self._device = int(device)
        self._slot: Dict[WorkerId, int] = {}
        self._worker_of_slot: List[Optional[WorkerId]] = [None] * self.max_workers

    def _slot_of(self, worker_id: WorkerId, create: bool) -> Optional[int]:
        s = self._slot.get(worker_id)
        if s is None and create:
            free = [i for i, w in enumerate(self._worker_of_slot) if w is None]
            if not free:
                raise KvError(f"the prefix index holds at most {self.max_workers} workers at a time (removed workers that "
                              "left stale blocks behind keep their slot)")
            s = free[0]
            self._slot[worker_id] = s
            self._worker_of_slot[s] = worker_id
        return s

    def apply_event(self, event: RouterEvent) -> None:
        L = _lib.lib()
        slot = self._slot_of(event.worker_id, True)
        data = event.event.data
        if isinstance(data, KvCacheStoreData):
            bh = np.array([b.block_hash for b in data.blocks], dtype=np.uint64)
            th = np.array([b.tokens_hash for b in data.blocks], dtype=np.uint64)
            rc = L.kvdp_index_apply_stored(self._h, slot, 0 if data.parent_hash is None else 1,
                                           int(data.parent_hash or 0), bh.ctypes.data if bh.size else None,
                                           th.ctypes.data if th.size else None, bh.size)
            if rc < 0:
                _lib.check(rc)
            # rc == 1: parent unknown, the reference logs a warning and skips the event (indexer.rs:305-315)
        else:
            bh = np.array(list(data.block_hashes), dtype=np.uint64)
            _lib.check(L.kvdp_index_apply_removed(self._h, slot, bh.ctypes.data if bh.size else None, bh.size))

    def remove_worker(self, worker: WorkerId) -> None:
        slot = self._slot_of(worker, False)
        if slot is None:
            return
        _lib.check(_lib.lib().kvdp_index_remove_worker(self._h, slot))
        # Like the reference, a block whose lookup entry was overwritten keeps the removed worker in its set
        # (remove_worker only visits the lookup table, indexer.rs:380-386); such a stale membership must keep reporting
        # under the old id, so the slot stays bound to it. Only when no node holds the slot any more is it recycled.
        if self.stat(6, slot) == 0:
            del self._slot[worker]
            self._worker_of_slot[slot] = None

    def _frequencies(self, sequence: np.ndarray, early_exit: bool, now: Optional[float]) -> List[int]:
        import time
        now_ns = time.monotonic_ns() if now is None else int(round(now * 1e9))
        out = np.zeros(max(sequence.size, 1), dtype=np.uint64)
        n = _lib.lib().kvdp_index_touch(self._h, sequence.ctypes.data if sequence.size else None, sequence.size,
                                        int(early_exit), now_ns, int(round(self.expiration_duration * 1e9)),
                                        out.ctypes.data, out.size)
        if n < 0:
            _lib.check(int(n))
        return [int(x) for x in out[:n]]

    def find_matches_batch(self, sequences: Sequence[Sequence[int]], early_exit: bool = False,
                           now: Optional[float] = None) -> List[OverlapScores]:
        """Score many requests (each a list of LocalBlockHash) in one launch. With an expiration duration the visit
        bookkeeping runs per request in list order (`now`: clock override in seconds, for tests)."""
        lens = np.array([len(s) for s in sequences], dtype=np.int64)
        offs = np.zeros(len(sequences) + 1, dtype=np.int64)
        offs[1:] = np.cumsum(lens)
        flat = np.zeros(max(int(offs[-1]), 1), dtype=np.uint64)
        pos = 0
        for s in sequences:
            flat[pos:pos + len(s)] = np.asarray(s, dtype=np.uint64)
            pos += len(s)
        scores, _ = self.find_matches_device_arrays(flat, offs, early_exit)
        out = [self._to_overlap(row) for row in scores]
        if self.expiration_duration is not None:
            for i, o in enumerate(out):
                o.frequencies = self._frequencies(flat[offs[i]:offs[i + 1]], early_exit, now)
        return out

    def find_matches(self, sequence: Sequence[int], early_exit: bool = False, now: Optional[float] = None) -> OverlapScores:
        """reference: indexer.rs:239-277 (one request)."""
        return self.find_matches_batch([list(sequence)], early_exit, now)[0]

    def find_matches_device_arrays(self, hashes: np.ndarray, seq_offsets: np.ndarray, early_exit: bool = False):
        """Host arrays in, dense [n_req][MAX_WORKERS-in-use] scores out (slot order) plus matched depths."""
        L = _lib.lib()
        n_req = seq_offsets.size - 1
        nw = max([s for s in self._slot.values()], default=-1) + 1
        nw = max(nw, 1)
        if n_req == 0:
            return np.zeros((0, nw), dtype=np.uint32), np.zeros(0, dtype=np.int32)
        hashes = np.ascontiguousarray(hashes, dtype=np.uint64)
        seq_offsets = np.ascontiguousarray(seq_offsets, dtype=np.int64)
        scores = np.zeros((n_req, nw), dtype=np.uint32)
        depth = np.zeros(n_req, dtype=np.int32)
        _lib.check(L.kvdp_index_find_matches_host(self._h, hashes.ctypes.data if hashes.size else None,
                                                  seq_offsets.ctypes.data, n_req, int(early_exit), nw,
                                                  scores.ctypes.data, depth.ctypes.data))
        return scores, depth

    def scores_for_tokens(self, requests: Sequence, kv_block_size: int) -> np.ndarray:
        """[n_req][workers in use] overlap counts for token sequences: seed-1337 block hashing and the tree walk happen
        in one upload / launch sequence / download."""
        toks = [_as_u32(t) for t in requests]
        n_req = len(toks)
        nw = max(max([s for s in self._slot.values()], default=-1) + 1, 1)
        if n_req == 0:
            return np.zeros((0, nw), dtype=np.uint32)
        offs = np.zeros(n_req + 1, dtype=np.int64)
        offs[1:] = np.cumsum([t.size for t in toks])
        flat = np.ascontiguousarray(np.concatenate(toks)) if offs[-1] else np.zeros(1, dtype=np.uint32)
        scores = np.zeros((n_req, nw), dtype=np.uint32)
        _lib.check(_lib.lib().kvdp_index_find_matches_for_tokens_host(
            self._h, flat.ctypes.data, offs.ctypes.data, n_req, int(kv_block_size), 0, nw, scores.ctypes.data, None))
        return scores

    def _to_overlap(self, row: np.ndarray) -> OverlapScores:
        out = OverlapScores()
        for slot, c in enumerate(row):
            if c and self._worker_of_slot[slot] is not None:
                out.scores[self._worker_of_slot[slot]] = int(c)
        return out

    def stat(self, what: int, arg: int = 0) -> int:
        return int(_lib.lib().kvdp_index_stat(self._h, what, arg))

    def lookup_len(self, worker: WorkerId) -> int:
        s = self._slot_of(worker, False)
        return 0 if s is None else self.stat(2, s)

    def close(self) -> None:
        if self._h is not None and self._h.value:
            _lib.lib().kvdp_index_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class KvIndexer:
    """reference: KvIndexer (indexer.rs:498-667) without the tokio actor: same calls, synchronous."""

    def __init__(self, kv_block_size: int, max_nodes: int = 1 << 20, device: int = -1,
                 expiration_duration: Optional[float] = None, max_workers: int = MAX_WORKERS):
        self.kv_block_size = int(kv_block_size)
        self.tree = RadixTree(max_nodes, device, expiration_duration, max_workers)

    @classmethod
    def new_with_frequency(cls, expiration_duration: Optional[float], kv_block_size: int, **kw) -> "KvIndexer":
        """reference: indexer.rs:525-593."""
        return cls(kv_block_size, expiration_duration=expiration_duration, **kw)

    def block_size(self) -> int:
        return self.kv_block_size

    def apply_event(self, event: RouterEvent) -> None:
        self.tree.apply_event(event)

    def remove_worker(self, worker: WorkerId) -> None:
        self.tree.remove_worker(worker)

    def find_matches(self, sequence: Sequence[int], now: Optional[float] = None) -> OverlapScores:
        return self.tree.find_matches(sequence, False, now)

    def find_matches_for_request(self, tokens) -> OverlapScores:
        """reference: indexer.rs:639-651 — hash the request's full blocks (seed 1337, un-chained), then match."""
        return self.find_matches_for_requests([tokens])[0]

    def find_matches_for_requests(self, requests: Sequence) -> List[OverlapScores]:
        """Batched form: token lists in, one OverlapScores per request out; hashing and matching in ONE round trip to the
        GPU (kvdp_index_find_matches_for_tokens_host)."""
        if self.tree.expiration_duration is not None:          # the visit bookkeeping needs the block hashes on the host
            from .tokens import compute_block_hash_for_seq
            return self.tree.find_matches_batch([compute_block_hash_for_seq(t, self.kv_block_size) for t in requests])
        scores = self.tree.scores_for_tokens(requests, self.kv_block_size)
        return [self.tree._to_overlap(row) for row in scores]


class KvIndexerSharded:
    """reference: KvIndexerSharded (indexer.rs:677-884): several independent radix trees, every worker bound to the
    shard that had the fewest workers when its first event arrived; a lookup asks every shard and merges the answers.
    Here each shard is one device-resident prefix index holding up to `max_workers` (<= 512) workers, so the sharded
    indexer holds max_workers x num_shards. Like the reference, a worker's binding is dropped by remove_worker and chosen afresh afterwards."""

    def __init__(self, num_shards: int, kv_block_size: int, max_nodes: int = 1 << 20, device: int = -1,
                 expiration_duration: Optional[float] = None, max_workers: int = MAX_WORKERS):
        if num_shards < 1:
            raise KvError("num_shards must be at least 1")
        self.kv_block_size = int(kv_block_size)
        self.shards = [RadixTree(max_nodes, device, expiration_duration, max_workers) for _ in range(int(num_shards))]
        self.worker_assignments: Dict[WorkerId, int] = {}
        self.worker_counts = [0] * int(num_shards)

    def block_size(self) -> int:
        return self.kv_block_size

    def apply_event(self, event: RouterEvent) -> None:
        shard = self.worker_assignments.get(event.worker_id)
        if shard is None:
            shard = min(range(len(self.shards)), key=lambda i: self.worker_counts[i])   # first of the smallest
            self.worker_assignments[event.worker_id] = shard
            self.worker_counts[shard] += 1
        self.shards[shard].apply_event(event)

    def remove_worker(self, worker: WorkerId) -> None:
        shard = self.worker_assignments.pop(worker, None)
        if shard is not None:
            self.worker_counts[shard] -= 1
            self.shards[shard].remove_worker(worker)

    @staticmethod
    def _merge(parts: Sequence[OverlapScores]) -> OverlapScores:
        out = OverlapScores()
        for k, p in enumerate(parts):
            out.scores.update(p.scores)
            if k == 0:
                out.frequencies = list(p.frequencies)
            else:
                out.frequencies += [0] * (len(p.frequencies) - len(out.frequencies))
                for i, f in enumerate(p.frequencies):
                    out.frequencies[i] += f
        return out

    def find_matches_batch(self, sequences: Sequence[Sequence[int]], now: Optional[float] = None) -> List[OverlapScores]:
        per_shard = [t.find_matches_batch(sequences, False, now) for t in self.shards]
        return [self._merge([ps[i] for ps in per_shard]) for i in range(len(sequences))]

    def find_matches(self, sequence: Sequence[int], now: Optional[float] = None) -> OverlapScores:
        return self.find_matches_batch([list(sequence)], now)[0]

    def find_matches_for_request(self, tokens) -> OverlapScores:
        return self.find_matches_for_requests([tokens])[0]

    def find_matches_for_requests(self, requests: Sequence) -> List[OverlapScores]:
        if self.shards[0].expiration_duration is not None:
            from .tokens import compute_block_hash_for_seq
            return self.find_matches_batch([compute_block_hash_for_seq(t, self.kv_block_size) for t in requests])
        per_shard = [[t._to_overlap(row) for row in t.scores_for_tokens(requests, self.kv_block_size)]
                     for t in self.shards]
        return [self._merge([ps[i] for ps in per_shard]) for i in range(len(requests))]

    def shutdown(self) -> None:
        for t in self.shards:
            t.close()
