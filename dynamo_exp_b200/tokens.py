"""Host-side mirror of the reference's token-block hashing API, computed on the GPU through libkvdp.so.

Mirrors `lib/tokens/src/lib.rs` (`compute_hash`, `Tokens`, `PartialTokenBlock`, `TokenBlock`, `TokenBlockSequence`),
the router indexer's `compute_block_hash_for_seq` (`lib/llm/src/kv_router/indexer.rs:64-135`) and, behind an explicit
flag, the in-crate twin (`lib/llm/src/tokens.rs`). Every hash is produced by the CUDA kernels in csrc/token_hash.cu;
nothing here hashes on the CPU.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Iterable, List, Optional, Tuple

import numpy as np

from . import _lib
from .storage import KvError

XXH3_SEED = 1337  # kv_router/indexer.rs:64

Token = int
SaltHash = int
BlockHash = int
SequenceHash = int


def _as_u32(tokens) -> np.ndarray:
    """`From<Vec<i32>>` casts `as u32` (lib/tokens/src/lib.rs:91-101): two's-complement reinterpretation."""
    a = np.asarray(tokens)
    if a.dtype == np.uint32:
        return np.ascontiguousarray(a)
    if a.size == 0:
        return np.zeros(0, dtype=np.uint32)
    if a.dtype.kind == "i" and a.dtype.itemsize == 4:
        return np.ascontiguousarray(a).view(np.uint32)
    return np.ascontiguousarray((a.astype(np.int64) & 0xFFFFFFFF).astype(np.uint32))


def compute_hash(data: bytes, seed: int) -> int:
    """xxh3_64_with_seed on the device (reference: lib/tokens/src/lib.rs:44-46)."""
    _lib.require_gpu()
    L = _lib.lib()
    n = len(data)
    d_in, d_out = C.c_void_p(), C.c_void_p()
    _lib.check(L.kvdp_device_alloc(C.byref(d_in), max(n, 1), -1))
    _lib.check(L.kvdp_device_alloc(C.byref(d_out), 8, -1))
    try:
        if n:
            buf = (C.c_char * n).from_buffer_copy(data)
            _lib.check(L.cuda_memcpy_sync(d_in, C.addressof(buf), n))
        _lib.check(L.kvdp_xxh3_64_batch(d_in, 1, n, seed & 0xFFFFFFFFFFFFFFFF, d_out, None))
        out = C.c_uint64(0)
        _lib.check(L.cuda_memcpy_sync(C.addressof(out), d_out, 8))
        return int(out.value)
    finally:
        L.kvdp_device_free(d_in)
        L.kvdp_device_free(d_out)


def hash_token_blocks(tokens: np.ndarray, offsets: np.ndarray, block_size: int, salt: int = 0, *, chained: bool = True,
                      twin_first_block: bool = False, device: int = -1) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Batched (CSR) hashing of many requests in one GPU pass.

    Returns (block_offsets[n_req+1], block_hash[], sequence_hash[]); `sequence_hash` is empty when `chained` is False.
    This is the v2 batched form of `TokenBlockSequence::split_tokens` (lib/tokens/src/lib.rs:369-404).
    """
    if block_size < 1:
        raise KvError("block_size must be at least 1")
    _lib.require_gpu()
    tok = _as_u32(tokens)
    off = np.ascontiguousarray(offsets, dtype=np.int64)
    if off.ndim != 1 or off.size < 1:
        raise KvError("offsets must hold n_req + 1 entries")
    n_req = off.size - 1
    if n_req and (np.any(np.diff(off) < 0) or off[0] < 0 or off[-1] > tok.size):
        raise KvError("offsets must be non-decreasing and within the token array")
    total = int(((off[1:] - off[:-1]) // block_size).sum()) if n_req else 0
    boff = np.zeros(n_req + 1, dtype=np.int64)
    bh = np.zeros(total, dtype=np.uint64)
    sh = np.zeros(total if chained else 0, dtype=np.uint64)
    flags = _lib.HASH_TWIN_FIRST if twin_first_block else 0
    _lib.check(_lib.lib().kvdp_hash_token_blocks_host(
        C.c_void_p(tok.ctypes.data if tok.size else 0), C.c_void_p(off.ctypes.data), n_req, int(block_size),
        salt & 0xFFFFFFFFFFFFFFFF, C.c_void_p(boff.ctypes.data), C.c_void_p(bh.ctypes.data if total else 0),
        C.c_void_p(sh.ctypes.data) if (chained and total) else None, flags, device))
    return boff, bh, sh


def compute_block_hash_for_seq(tokens, kv_block_size: int) -> List[int]:
    """Un-chained `LocalBlockHash` per full block, seed 1337 (reference: kv_router/indexer.rs:123-135)."""
    tok = _as_u32(tokens)
    _, bh, _ = hash_token_blocks(tok, np.array([0, tok.size], dtype=np.int64), kv_block_size, XXH3_SEED,
                                 chained=False)
    return [int(x) for x in bh]


def kv_event_create_stored_from_parts(event_id: int, token_ids, num_block_tokens, block_ids, kv_block_size: int,
                                      parent_hash: Optional[int] = None, lora_id: int = 0) -> dict:
    """The hashing half of the engine-facing `dynamo_kv_event_publish_stored` (lib/bindings/c/src/lib.rs:157-227,
    260-295): one stored-blocks event from parts. Blocks are taken in order until the first one whose token count is
    not `kv_block_size` (the reference stops the batch there, :191-210); each published block carries the engine's
    block id as `block_hash` and the seed-1337 un-chained token hash as `tokens_hash`. All kept blocks are hashed in ONE
    GPU batch instead of one call per block. (Publishing to NATS is out of scope; this returns the event payload.)"""
    tok = _as_u32(token_ids)
    counts = [int(x) for x in num_block_tokens]
    ids = [int(x) for x in block_ids]
    if len(counts) != len(ids):
        raise KvError("num_block_tokens and block_ids must have the same length")
    keep = 0
    for n in counts:
        if n != kv_block_size:
            break
        keep += 1
    if keep * kv_block_size > tok.size:
        raise KvError("token_ids is shorter than the blocks it is said to hold")
    hashes = compute_block_hash_for_seq(tok[:keep * kv_block_size], kv_block_size) if keep else []
    return {"event_id": int(event_id),
            "data": {"stored": {"parent_hash": None if parent_hash is None else int(parent_hash),
                                "blocks": [{"block_hash": ids[i], "tokens_hash": hashes[i]} for i in range(keep)]}}}


class Tokens:
    """reference: lib/tokens/src/lib.rs:48-128."""

    def __init__(self, tokens: Iterable[int] = ()):
        self._t = _as_u32(list(tokens) if not isinstance(tokens, np.ndarray) else tokens).copy()

    def __len__(self) -> int:
        return int(self._t.size)

    def __iter__(self):
        return iter(int(x) for x in self._t)

    def __getitem__(self, i):
        r = self._t[i]
        return [int(x) for x in r] if isinstance(r, np.ndarray) else int(r)

    def __eq__(self, other) -> bool:
        o = other._t if isinstance(other, Tokens) else _as_u32(list(other))
        return self._t.shape == o.shape and bool(np.all(self._t == o))

    def as_array(self) -> np.ndarray:
        return self._t

    def into_sequence(self, block_size: int, salt_hash: Optional[int] = None) -> "TokenBlockSequence":
        return TokenBlockSequence(self, block_size, salt_hash)

    def __repr__(self):
        return f"Tokens({self._t.tolist()})"


@dataclass
class TokenBlock:
    """reference: lib/tokens/src/lib.rs:217-262 (getters become attributes)."""
    tokens: Tokens
    salt_hash: int
    block_hash: int
    sequence_hash: int
    parent_sequence_hash: Optional[int]


class PartialTokenBlock:
    """reference: lib/tokens/src/lib.rs:139-187."""

    def __init__(self, tokens: Tokens, block_size: int, salt_hash: int, parent_sequence_hash: Optional[int],
                 twin: bool = False):
        self._tokens = tokens
        self.block_size = block_size
        self.salt_hash = salt_hash
        self.parent_sequence_hash = parent_sequence_hash
        self._twin = twin

    def push_token(self, token: int) -> Optional[TokenBlock]:
        if not self._twin:
            assert len(self._tokens) < self.block_size  # lib/tokens/src/lib.rs:153
        self._tokens._t = np.append(self._tokens._t, np.uint32(token & 0xFFFFFFFF))
        if len(self._tokens) != self.block_size:
            return None
        tokens, self._tokens = self._tokens, Tokens()
        # one block: block hash, then (if there is a parent) the 16-byte chain step, both on the device
        parent = self.parent_sequence_hash
        if self._twin:
            # lib/llm/src/tokens.rs:131-134 hashes [parent.unwrap_or(0), block_hash] even for the first block,
            # and never advances its own parent (the caller rebuilds the PartialTokenBlock)
            _, bh, _ = hash_token_blocks(tokens.as_array(), np.array([0, len(tokens)]), self.block_size,
                                         self.salt_hash, chained=False)
            seq = compute_chain_step(parent or 0, int(bh[0]), self.salt_hash)
            return TokenBlock(tokens, self.salt_hash, int(bh[0]), seq, parent)
        _, bh, _ = hash_token_blocks(tokens.as_array(), np.array([0, len(tokens)]), self.block_size, self.salt_hash,
                                     chained=False)
        block_hash = int(bh[0])
        seq = block_hash if parent is None else compute_chain_step(parent, block_hash, self.salt_hash)
        self.parent_sequence_hash = seq
        return TokenBlock(tokens, self.salt_hash, block_hash, seq, parent)

    def tokens(self) -> Tokens:
        return self._tokens

    def remaining_tokens(self) -> int:
        return self.block_size - len(self._tokens)

    def __len__(self) -> int:
        return len(self._tokens)


def compute_chain_step(parent_sequence_hash: int, block_hash: int, salt_hash: int) -> int:
    """hash(LE64(parent) || LE64(block_hash), seed = salt) (reference: lib/tokens/src/lib.rs:241-245)."""
    data = int(parent_sequence_hash).to_bytes(8, "little") + int(block_hash).to_bytes(8, "little")
    return compute_hash(data, salt_hash)


class TokenBlockSequence:
    """reference: lib/tokens/src/lib.rs:264-405."""

    def __init__(self, tokens, block_size: int, salt_hash: Optional[int] = None):
        if not isinstance(tokens, Tokens):
            tokens = Tokens(tokens)
        self._salt_hash = 0 if salt_hash is None else int(salt_hash)
        self._blocks, self._current = TokenBlockSequence.split_tokens(tokens, block_size, self._salt_hash)

    @staticmethod
    def split_tokens(tokens: Tokens, block_size: int, salt_hash: int) -> Tuple[List[TokenBlock], PartialTokenBlock]:
        t = tokens.as_array()
        _, bh, sh = hash_token_blocks(t, np.array([0, t.size], dtype=np.int64), block_size, salt_hash)
        blocks: List[TokenBlock] = []
        for b in range(bh.size):
            parent = int(sh[b - 1]) if b else None
            blocks.append(TokenBlock(Tokens(t[b * block_size:(b + 1) * block_size]), salt_hash, int(bh[b]),
                                     int(sh[b]), parent))
        rem = Tokens(t[bh.size * block_size:])
        current = PartialTokenBlock(rem, block_size, salt_hash, blocks[-1].sequence_hash if blocks else None)
        return blocks, current

    def push_token(self, token: int) -> Optional[TokenBlock]:
        block = self._current.push_token(token)
        if block is not None:
            self._blocks.append(block)
        return block

    def last(self) -> Optional[TokenBlock]:
        return self._blocks[-1] if self._blocks else None

    def blocks(self) -> List[TokenBlock]:
        return self._blocks

    def current_block(self) -> PartialTokenBlock:
        return self._current

    def into_parts(self) -> Tuple[List[TokenBlock], PartialTokenBlock]:
        return self._blocks, self._current

    def salt_hash(self) -> int:
        return self._salt_hash


# ---------------------------------------------------------------------------------------------- in-crate twin
class TwinTokens:
    """`lib/llm/src/tokens.rs` (seed hard-wired to 1337); exposed only so its quirks can be tested explicitly."""

    @staticmethod
    def compute_block_hash(tokens, block_size: int) -> List[int]:
        """reference: lib/llm/src/tokens.rs:109-114."""
        return compute_block_hash_for_seq(tokens, block_size)

    @staticmethod
    def split_tokens(tokens, block_size: int) -> Tuple[List[TokenBlock], PartialTokenBlock]:
        """reference: lib/llm/src/tokens.rs:209-243; like the original it fails on less than one full block."""
        t = _as_u32(tokens)
        if t.size // block_size < 1:
            raise IndexError("index out of bounds: the len is 0 but the index is 0")  # blocks[0] panics in the twin
        blocks, cur = TokenBlockSequence.split_tokens(Tokens(t), block_size, XXH3_SEED)
        return blocks, PartialTokenBlock(cur.tokens(), block_size, XXH3_SEED, cur.parent_sequence_hash, twin=True)

    @staticmethod
    def partial(block_size: int, parent_sequence_hash: Optional[int] = None) -> PartialTokenBlock:
        return PartialTokenBlock(Tokens(), block_size, XXH3_SEED, parent_sequence_hash, twin=True)
