"""oracle/radix_oracle.py — TEST INFRASTRUCTURE ONLY: a CPU restatement of the KV router's prefix index.

Follows /root/reference/lib/llm/src/kv_router/indexer.rs:
    RadixBlock        :162-185   (children by LocalBlockHash, worker set)
    RadixTree         :187-387   (find_matches :239-277, apply_event :284-378, remove_worker :380-386)
    OverlapScores     :390-436
Pure-Python dicts and sets, object identity standing in for the reference's Rc<RefCell<..>> sharing (a block reachable
from several parents, or kept alive only by a worker's lookup table, behaves as it does there). The access-frequency
feature (expiration_duration, :251-267) is available through `find_matches_with_frequency`, with the clock passed in.

Parity pin: the reference's own unit test `test_radix_tree` (indexer.rs:949-1170) is replayed in
tests/test_oracle_radix.py, including its assertions on lookup-table and child counts.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple


class RadixBlock:
    __slots__ = ("children", "workers", "recent_uses")

    def __init__(self):
        self.children: Dict[int, "RadixBlock"] = {}
        self.workers: set = set()
        self.recent_uses: List[float] = []


class RadixTree:
    def __init__(self):
        self.root = RadixBlock()
        self.lookup: Dict[int, Dict[int, RadixBlock]] = {}

    # indexer.rs:239-277
    def find_matches(self, sequence: List[int], early_exit: bool) -> Dict[int, int]:
        scores: Dict[int, int] = {}
        current = self.root
        for h in sequence:
            block = current.children.get(h)
            if block is None:
                break
            for w in block.workers:
                scores[w] = scores.get(w, 0) + 1
            if early_exit and len(block.workers) == 1:
                break
            current = block
        return scores

    # indexer.rs:239-277 with expiration_duration = Some(expiration): returns (scores, frequencies)
    def find_matches_with_frequency(self, sequence: List[int], early_exit: bool, now: float, expiration: float):
        scores: Dict[int, int] = {}
        frequencies: List[int] = []
        current = self.root
        for h in sequence:
            block = current.children.get(h)
            if block is None:
                break
            for w in block.workers:
                scores[w] = scores.get(w, 0) + 1
            while block.recent_uses and now - block.recent_uses[0] > expiration:
                block.recent_uses.pop(0)
            if block.recent_uses:                       # add_frequency skips zeros (:429-436)
                frequencies.append(len(block.recent_uses))
            block.recent_uses.append(now)
            if early_exit and len(block.workers) == 1:
                break
            current = block
        return scores, frequencies

    # indexer.rs:284-378
    def apply_stored(self, worker_id: int, parent_hash: Optional[int], blocks: List[Tuple[int, int]]) -> bool:
        """blocks: [(block_hash (external), tokens_hash (local))]. Returns False when the parent is unknown."""
        worker_lookup = self.lookup.setdefault(worker_id, {})
        if parent_hash is None:
            current = self.root
        else:
            current = worker_lookup.get(parent_hash)
            if current is None:
                return False
        for block_hash, tokens_hash in blocks:
            block = current.children.get(tokens_hash)
            if block is None:
                block = worker_lookup.get(block_hash)
                if block is None:
                    block = RadixBlock()
                current.children[tokens_hash] = block
            block.workers.add(worker_id)
            worker_lookup[block_hash] = block
            current = block
        return True

    def apply_removed(self, worker_id: int, block_hashes: List[int]) -> None:
        worker_lookup = self.lookup.setdefault(worker_id, {})
        for h in block_hashes:
            entry = worker_lookup.get(h)
            if entry is None:
                continue
            entry.workers.discard(worker_id)
            if not entry.workers:
                entry.children.clear()
            del worker_lookup[h]

    # indexer.rs:380-386
    def remove_worker(self, worker_id: int) -> None:
        blocks = self.lookup.pop(worker_id, None)
        if blocks is not None:
            for block in blocks.values():
                block.workers.discard(worker_id)
