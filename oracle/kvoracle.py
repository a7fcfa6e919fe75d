"""ctypes front-end of the CPU oracle (oracle/libkvoracle.so).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs.  Nothing under ``dynamo_exp_b200/`` may import this module.

The C sources carry the reference citations (xxh3_oracle.c, layout_oracle.c).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libkvoracle.so")
_REF_PATH = os.path.join(_HERE, "_ref", "libblock_copy_ref.so")

_u64p = C.POINTER(C.c_uint64)
_u32p = C.POINTER(C.c_uint32)
_i64p = C.POINTER(C.c_int64)
_i32p = C.POINTER(C.c_int)


def build(force: bool = False) -> None:
    """Compile the oracle (and, where /root/reference exists, oracle/_ref) with the committed Makefile."""
    if force or not os.path.exists(_LIB_PATH) or os.path.exists("/root/reference"):
        subprocess.run(["make", "-C", _HERE, "--no-print-directory"], check=True, capture_output=True)


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        L.kvo_xxh3_64.restype = C.c_uint64
        L.kvo_xxh3_64.argtypes = [C.c_void_p, C.c_size_t, C.c_uint64]
        L.kvo_split_tokens.restype = C.c_size_t
        L.kvo_split_tokens.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_uint64, C.c_void_p, C.c_void_p, C.c_int]
        L.kvo_hash_batch.restype = None
        L.kvo_hash_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_uint64, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_int]
        L.kvo_indexer_block_hashes.restype = C.c_size_t
        L.kvo_indexer_block_hashes.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p]
        L.kvo_hw_thread_count.restype = C.c_int
        L.kvo_copy_blocks_3d.restype = C.c_int
        L.kvo_copy_blocks_3d.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 6
        L.kvo_copy_blocks_layers.restype = C.c_int
        L.kvo_copy_blocks_layers.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p] + [C.c_int] * 7
        L.kvo_scatter_6d.restype = C.c_int
        L.kvo_scatter_6d.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_uint32] * 6
        L.kvo_rearrange.restype = C.c_int
        L.kvo_rearrange.argtypes = [C.c_void_p, C.c_void_p] + [C.c_size_t] * 6 + [C.c_int]
        L.kvo_relayout_contig.restype = C.c_int
        L.kvo_relayout_contig.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.kvo_reshard_push.restype = C.c_int
        L.kvo_reshard_push.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t,
                                       C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        _lib = L
    return _lib


def hw_threads() -> int:
    return int(lib().kvo_hw_thread_count())


def _ptr(a: np.ndarray) -> int:
    return a.ctypes.data


# ----------------------------------------------------------------------------------------------- hashing
def xxh3_64(data: bytes, seed: int = 0) -> int:
    buf = (C.c_char * max(len(data), 1)).from_buffer_copy(data if data else b"\0")
    return int(lib().kvo_xxh3_64(C.addressof(buf), len(data), seed & 0xFFFFFFFFFFFFFFFF))


def split_tokens(tokens, block_size: int, salt: int = 0, twin_first_block: bool = False):
    """(block_hash[], sequence_hash[]) of the full blocks of one request (lib/tokens split_tokens)."""
    tok = np.ascontiguousarray(np.asarray(tokens).astype(np.uint32, copy=False))
    nb = tok.size // block_size if block_size else 0
    bh = np.zeros(nb, dtype=np.uint64)
    sh = np.zeros(nb, dtype=np.uint64)
    n = lib().kvo_split_tokens(_ptr(tok), tok.size, block_size, salt & 0xFFFFFFFFFFFFFFFF, _ptr(bh), _ptr(sh),
                               int(twin_first_block))
    assert n == nb
    return bh, sh


def hash_batch(tokens: np.ndarray, offsets: np.ndarray, block_size: int, salt: int, threads: int = 1):
    """CSR batch. Returns (out_offsets[n_req+1], block_hash[], seq_hash[])."""
    tok = np.ascontiguousarray(tokens, dtype=np.uint32)
    off = np.ascontiguousarray(offsets, dtype=np.int64)
    n_req = off.size - 1
    nblk = (off[1:] - off[:-1]) // block_size
    out_off = np.zeros(n_req + 1, dtype=np.int64)
    np.cumsum(nblk, out=out_off[1:])
    total = int(out_off[-1])
    bh = np.zeros(total, dtype=np.uint64)
    sh = np.zeros(total, dtype=np.uint64)
    lib().kvo_hash_batch(_ptr(tok), _ptr(off), n_req, block_size, salt & 0xFFFFFFFFFFFFFFFF, _ptr(out_off), _ptr(bh),
                         _ptr(sh), threads)
    return out_off, bh, sh


def indexer_block_hashes(tokens, block_size: int) -> np.ndarray:
    tok = np.ascontiguousarray(np.asarray(tokens).astype(np.uint32, copy=False))
    nb = tok.size // block_size if block_size else 0
    out = np.zeros(nb, dtype=np.uint64)
    lib().kvo_indexer_block_hashes(_ptr(tok), tok.size, block_size, _ptr(out))
    return out


# ----------------------------------------------------------------------------------------------- layout
def _ids(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.int32)


def copy_blocks_3d(src: np.ndarray, dst: np.ndarray, src_ids, dst_ids, prefix_dim, src_blocks, dst_blocks,
                   suffix_dim, elem_size) -> int:
    s, d = _ids(src_ids), _ids(dst_ids)
    return int(lib().kvo_copy_blocks_3d(_ptr(src), _ptr(dst), _ptr(s), _ptr(d), s.size, prefix_dim, src_blocks,
                                        dst_blocks, suffix_dim, elem_size))


def copy_blocks_layers(src_layers, dst_layers, src_ids, dst_ids, prefix_dim, src_blocks, dst_blocks, suffix_dim,
                       elem_size, threads=1) -> int:
    sp = np.array([_ptr(a) for a in src_layers], dtype=np.uint64)
    dp = np.array([_ptr(a) for a in dst_layers], dtype=np.uint64)
    s, d = _ids(src_ids), _ids(dst_ids)
    return int(lib().kvo_copy_blocks_layers(_ptr(sp), _ptr(dp), len(src_layers), _ptr(s), _ptr(d), s.size, prefix_dim,
                                            src_blocks, dst_blocks, suffix_dim, elem_size, threads))


def scatter_6d(src: np.ndarray, dst: np.ndarray, dims6, elem_size, block_dim_index, src_block_dim, dst_block_dim,
               num_blocks) -> int:
    dims = np.ascontiguousarray(dims6, dtype=np.uint32)
    return int(lib().kvo_scatter_6d(_ptr(src), _ptr(dst), _ptr(dims), dims.size, elem_size, block_dim_index,
                                    src_block_dim, dst_block_dim, num_blocks))


def rearrange(t1: np.ndarray, t2: np.ndarray, d: int, direction: str) -> int:
    N, B, H, Cc = t1.shape
    return int(lib().kvo_rearrange(_ptr(t1), _ptr(t2), N, B, H, Cc, t1.itemsize, d, 1 if direction == "write" else 0))


def relayout_contig(layers, layout: int, layer_blocks: int, chunk_bytes: int, contig: np.ndarray, src_ids, dst_ids,
                    to_contig: bool, threads: int = 1) -> int:
    lp = np.array([_ptr(a) for a in layers], dtype=np.uint64)
    s, d = _ids(src_ids), _ids(dst_ids)
    return int(lib().kvo_relayout_contig(_ptr(lp), len(layers), layout, layer_blocks, chunk_bytes, _ptr(contig),
                                         _ptr(s), _ptr(d), s.size, int(to_contig), threads))


def reshard_push(src_layers, dst_layers_by_target, src_blocks, dst_blocks, bs, row_bytes, src_tp, dst_tp, src_rank,
                 src_ids, dst_ids) -> int:
    """dst_layers_by_target: list (targets) of lists (layers) of numpy arrays."""
    sp = np.array([_ptr(a) for a in src_layers], dtype=np.uint64)
    dp = np.array([_ptr(a) for tgt in dst_layers_by_target for a in tgt], dtype=np.uint64)
    s, d = _ids(src_ids), _ids(dst_ids)
    return int(lib().kvo_reshard_push(_ptr(sp), _ptr(dp), len(src_layers), src_blocks, dst_blocks, bs, row_bytes,
                                      src_tp, dst_tp, src_rank, _ptr(s), _ptr(d), s.size))


# ----------------------------------------------------------------------------------------------- oracle/_ref
def ref_lib():
    """The reference's own block_copy.cu built as a shared library (needs a GPU at run time). None if absent."""
    if not os.path.exists(_REF_PATH):
        return None
    return C.CDLL(_REF_PATH)
