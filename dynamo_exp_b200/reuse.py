"""Host-side mirror of the KV block reuse pool and prefill planning:

    KvBlock               lib/llm/src/kv.rs:47-91
    AvailableBlocks       lib/llm/src/kv/reuse.rs:40-660  (match by sequence hash, priority / return-tick eviction order,
                          un-initialised set, update, reset, reset_all) — the actor/channel plumbing is replaced by
                          direct synchronous calls; `fence()` is therefore a no-op
    ReservedBlocks        lib/llm/src/kv/reserved.rs:15-160 (in-flight blocks shared by sequence hash; a block goes back
                          to the pool when its last holder lets go)
    KvStorageManager      lib/llm/src/kv/manager.rs:15-122 (prepare_prefill_sequence / prepare_prefill_offload)

This layer decides WHICH blocks move; it produces the block-id lists that the copy / relayout / transfer kernels
consume. Each KvBlock therefore carries `block_id`, its slot in the KV cache slabs (the reference leaves storage
binding as a TODO, kv.rs:54-63). Sequence hashes come from the GPU hashing path (dynamo_exp_b200.tokens) with the
in-crate twin's fixed seed 1337, as `crate::tokens` does for this module.
"""
from __future__ import annotations

from collections import deque
from dataclasses import dataclass, field
from typing import Deque, Dict, List, Optional, Sequence

from sortedcontainers import SortedDict

from .storage import KvError
from .tokens import PartialTokenBlock, TokenBlock, TokenBlockSequence, Tokens, XXH3_SEED

SequenceHash = int


def _empty_token_block() -> TokenBlock:
    return TokenBlock(Tokens(), 0, 0, 0, None)          # TokenBlock::default(): sequence_hash 0


@dataclass
class KvBlock:
    """reference: kv.rs:47-91."""
    token_block: TokenBlock = field(default_factory=_empty_token_block)
    priority: int = 0
    return_tick: int = 0
    block_id: int = -1                  # slot in the KV cache slabs (new: what the kernels index with)

    def update_token_block(self, token_block: TokenBlock) -> None:
        self.token_block = token_block

    def reset(self) -> None:
        self.token_block = _empty_token_block()
        self.priority = 0
        self.return_tick = 0


@dataclass
class UpdateBlock:                      # reuse.rs:596-600
    hash: SequenceHash
    priority: Optional[int] = None


class PoolItem:
    """A block on loan from the pool (`PoolItem<KvBlock>`); `drop()` (or garbage collection) returns it."""

    def __init__(self, block: KvBlock, pool: "AvailableBlocks"):
        self.block = block
        self._pool: Optional[AvailableBlocks] = pool

    def __getattr__(self, name):        # Deref to the KvBlock
        return getattr(self.block, name)

    def update_token_block(self, token_block: TokenBlock) -> None:
        self.block.update_token_block(token_block)

    def drop(self) -> None:
        if self._pool is not None:
            pool, self._pool = self._pool, None
            pool._handle_return(self.block)

    def into_shared(self) -> "PoolItem":
        return self

    def __del__(self):
        try:
            self.drop()
        except Exception:
            pass


class AvailableBlocks:
    """reference: AvailableBlocksState, reuse.rs:280-590."""

    def __init__(self):
        self.lookup_map: Dict[SequenceHash, KvBlock] = {}
        self.priority_set: SortedDict = SortedDict()        # (priority, return_tick) -> sequence_hash
        self.uninitialized_set: Deque[KvBlock] = deque()
        self.return_tick = 0
        self._total = 0
        self._available = 0

    # counters (reuse.rs:52-58)
    def total_blocks(self) -> int:
        return self._total

    def available_blocks(self) -> int:
        return self._available

    def is_active(self) -> bool:
        return True

    def fence(self) -> None:
        """The reference waits for its actor to drain; calls here are already synchronous."""

    # state machine
    def _insert(self, block: KvBlock) -> None:
        h = block.token_block.sequence_hash
        if h in self.lookup_map or h == 0:                  # one entry per hash; hash 0 = un-initialised
            self.uninitialized_set.append(block)
            return
        key = (block.priority, block.return_tick)
        assert key not in self.priority_set, "fatal error: multiple entries for the same key in priority set"
        self.priority_set[key] = h
        self.lookup_map[h] = block

    def _take_with_sequence_hash(self, h: SequenceHash) -> Optional[KvBlock]:
        block = self.lookup_map.pop(h, None)
        if block is not None:
            self.priority_set.pop((block.priority, block.return_tick), None)
        return block

    def insert(self, block: KvBlock) -> None:               # handle_insert, reuse.rs:505-517
        self._available += 1
        self._total += 1
        self.return_tick += 1
        block.return_tick = self.return_tick
        self._insert(block)

    def _handle_return(self, block: KvBlock) -> None:       # handle_return, reuse.rs:518-528
        self._available += 1
        self.return_tick += 1
        block.return_tick = self.return_tick
        self._insert(block)

    def match_blocks(self, hashes: Sequence[SequenceHash]) -> List[PoolItem]:
        """Take the blocks whose hashes match, in order, stopping at the first miss (reuse.rs:355-374)."""
        out: List[PoolItem] = []
        for h in hashes:
            block = self._take_with_sequence_hash(h)
            if block is None:
                break
            out.append(PoolItem(block, self))
        self._available -= len(out)
        return out

    def match_token_blocks(self, token_blocks: Sequence[TokenBlock]) -> List[PoolItem]:
        return self.match_blocks([b.sequence_hash for b in token_blocks])

    def _take(self) -> Optional[KvBlock]:                   # reuse.rs:400-420
        if self.uninitialized_set:
            return self.uninitialized_set.popleft()
        if self.priority_set:
            _, h = self.priority_set.popitem(0)
            block = self.lookup_map.pop(h, None)
            if block is None:
                raise RuntimeError("block from priority set not found in lookup map")
            return block
        return None

    def take_blocks(self, count: int) -> List[PoolItem]:
        out: List[PoolItem] = []
        for _ in range(count):
            block = self._take()
            if block is None:
                break
            out.append(PoolItem(block, self))
        self._available -= len(out)
        return out

    def update_single(self, update: UpdateBlock) -> None:
        self.update_multiple([update])

    def update_multiple(self, updates: Sequence[UpdateBlock]) -> None:   # reuse.rs:537-552
        for u in updates:
            block = self._take_with_sequence_hash(u.hash)
            if block is not None:
                if u.priority is not None:
                    block.priority = u.priority
                self._insert(block)

    def reset(self, sequence_hashes: Sequence[SequenceHash]) -> None:    # reuse.rs:554-561
        for h in sequence_hashes:
            block = self._take_with_sequence_hash(h)
            if block is not None:
                block.reset()
                self._insert(block)

    def reset_all(self) -> None:                                         # reuse.rs:563-573
        while self.priority_set:
            _, h = self.priority_set.popitem(0)
            block = self.lookup_map.pop(h, None)
            if block is None:
                raise RuntimeError("block from priority set not found in lookup map")
            block.reset()
            self._insert(block)


class _ReservedInner:
    def __init__(self, block: PoolItem, owner: "ReservedBlocks"):
        self.block = block
        self.owner = owner
        self.refs = 0


class ReservedBlock:
    """A shared in-flight block (reserved.rs:21-44). Clone with `clone()`, release with `drop()`."""

    def __init__(self, inner: _ReservedInner):
        self._inner: Optional[_ReservedInner] = inner
        inner.refs += 1

    def inflight_count(self) -> int:
        return self._inner.refs if self._inner else 0

    def clone(self) -> "ReservedBlock":
        return ReservedBlock(self._inner)

    def __getattr__(self, name):        # Deref to the SharedBlock / KvBlock
        return getattr(self._inner.block, name)

    def drop(self) -> None:
        inner, self._inner = self._inner, None
        if inner is None:
            return
        inner.refs -= 1
        if inner.refs == 0:             # Drop for ReservedBlockInner (reserved.rs:51-64) + PoolItem return
            h = inner.block.token_block.sequence_hash
            if inner.owner.blocks.get(h) is inner:
                del inner.owner.blocks[h]
            inner.block.drop()

    def __del__(self):
        try:
            self.drop()
        except Exception:
            pass


class ReservedBlocks:
    """reference: reserved.rs:66-150."""

    def __init__(self, block_size: int):
        self.block_size = block_size
        self.blocks: Dict[SequenceHash, _ReservedInner] = {}

    def match_sequence_hashes(self, sequence_hashes: Sequence[SequenceHash]) -> List[ReservedBlock]:
        out = []
        for h in sequence_hashes:
            inner = self.blocks.get(h)
            if inner is None or inner.refs == 0:
                break
            out.append(ReservedBlock(inner))
        return out

    def match_token_blocks(self, token_blocks: Sequence[TokenBlock]) -> List[ReservedBlock]:
        return self.match_sequence_hashes([b.sequence_hash for b in token_blocks])

    def register(self, block: PoolItem) -> ReservedBlock:
        h = block.token_block.sequence_hash
        if len(block.token_block.tokens) != self.block_size:
            raise KvError("Block size mismatch")
        existing = self.blocks.get(h)
        if existing is not None and existing.refs > 0:
            block.drop()                # the caller's duplicate goes back to the free pool (reserved.rs:131-140)
            return ReservedBlock(existing)
        inner = _ReservedInner(block.into_shared(), self)
        self.blocks[h] = inner
        return ReservedBlock(inner)


@dataclass
class PartialKvBlock:                   # manager.rs:124-128
    token_block: PartialTokenBlock
    kv_block: PoolItem


@dataclass
class PrefillMatched:                   # manager.rs:130-135
    inflight_blocks: List[ReservedBlock]
    remaining_blocks: List[TokenBlock]
    tail_block: PartialTokenBlock


@dataclass
class PrefillOffload:                   # manager.rs:137-142
    inflight_blocks: List[ReservedBlock]
    complete_prefill_blocks: List[PoolItem]
    tail_prefill_block: PartialKvBlock

    def block_ids(self) -> List[int]:
        """Cache slots of the whole request in sequence order — what the copy / transfer kernels take."""
        return ([b.block_id for b in self.inflight_blocks] + [b.block_id for b in self.complete_prefill_blocks]
                + [self.tail_prefill_block.kv_block.block_id])


class KvStorageManager:
    """reference: manager.rs:20-122 — reservation and priority reuse of KV blocks for one storage tier."""

    def __init__(self, block_size: int):
        self.available_blocks = AvailableBlocks()
        self.inflight_blocks = ReservedBlocks(block_size)
        self.block_size = block_size

    def prepare_prefill_sequence(self, tokens) -> PrefillMatched:
        seq = TokenBlockSequence(tokens if isinstance(tokens, Tokens) else Tokens(tokens), self.block_size, XXH3_SEED)
        blocks, tail_block = seq.into_parts()
        inflight = self.inflight_blocks.match_token_blocks(blocks)
        unmatched_hashes = [b.sequence_hash for b in blocks[len(inflight):]]
        for block in self.available_blocks.match_blocks(unmatched_hashes):
            inflight.append(self.inflight_blocks.register(block))
        return PrefillMatched(inflight, list(blocks[len(inflight):]), tail_block)

    def prepare_prefill_offload(self, matched: PrefillMatched) -> PrefillOffload:
        need = len(matched.remaining_blocks) + 1
        blocks_to_reuse = self.available_blocks.take_blocks(need)
        if len(blocks_to_reuse) != need:
            got = len(blocks_to_reuse)
            for b in blocks_to_reuse:
                b.drop()
            raise KvError(f"expected {need} blocks, got {got}")
        complete: List[PoolItem] = []
        for tb in matched.remaining_blocks:
            block = blocks_to_reuse.pop()
            block.update_token_block(tb)
            complete.append(block)
        assert len(blocks_to_reuse) == 1
        tail = PartialKvBlock(matched.tail_block, blocks_to_reuse.pop())
        return PrefillOffload(matched.inflight_blocks, complete, tail)
