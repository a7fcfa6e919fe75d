"""GPU: kvdp_hash_token_blocks_noalloc — the hashing entry point for callers that cannot allocate device memory
(for instance a router thread inside an engine that gave the whole GPU to its KV cache at start-up).

It must produce exactly what kvdp_hash_token_blocks produces with a caller-supplied workspace, call after call, even
though its scratch is one array shared by every call and still holds the previous call's partial sums. The batches
here hold tens of thousands of requests on purpose: only above 2 048 requests does the block-count scan go through the
scratch at all.
This file sorts last so that a run stopping at the first failure has already been through every other test.
"""
import ctypes as C

import numpy as np
import pytest

from tests.util import DevBuf

pytestmark = pytest.mark.gpu


def make_batch(n_req, seed, max_len):
    rng = np.random.Generator(np.random.PCG64(seed))
    lens = rng.integers(0, max_len + 1, size=n_req).astype(np.int64)
    lens[rng.integers(0, n_req, size=n_req // 50)] = 0          # empty requests
    lens[rng.integers(0, n_req, size=64)] = 64 * 40             # a few long ones
    offs = np.zeros(n_req + 1, dtype=np.int64)
    offs[1:] = np.cumsum(lens)
    toks = rng.integers(0, 128256, size=int(offs[-1]), dtype=np.uint32)
    return toks, offs


def run_both(gpu, toks, offs, block_size, salt, flags=0):
    """Hash one batch through the existing entry point (own workspace) and through the no-alloc one."""
    L = gpu.lib()
    n_req, total = offs.size - 1, int(offs[-1])
    nblk = int(((offs[1:] - offs[:-1]) // block_size).sum())
    d_tok, d_off = DevBuf(gpu, max(toks.nbytes, 16)), DevBuf(gpu, offs.nbytes)
    d_tok.upload(toks)
    d_off.upload(offs)
    ws_bytes = L.kvdp_hash_workspace_bytes(n_req)
    d_ws = DevBuf(gpu, ws_bytes)
    out = []
    for which in ("workspace", "noalloc"):
        d_boff, d_bh, d_sh = DevBuf(gpu, (n_req + 1) * 8), DevBuf(gpu, max(nblk, 1) * 8), DevBuf(gpu, max(nblk, 1) * 8)
        if which == "workspace":
            rc = L.kvdp_hash_token_blocks(d_tok.ptr, d_off.ptr, n_req, total, block_size, salt, d_boff.ptr, d_bh.ptr,
                                          d_sh.ptr, d_ws.ptr, ws_bytes, flags, None)
        else:
            rc = L.kvdp_hash_token_blocks_noalloc(d_tok.ptr, d_off.ptr, n_req, total, block_size, salt, d_boff.ptr,
                                                  d_bh.ptr, d_sh.ptr, flags, None)
        gpu.check(rc)
        out.append((d_boff.download(np.int64), d_bh.download(np.uint64)[:nblk], d_sh.download(np.uint64)[:nblk]))
    return out


def test_noalloc_matches_workspace_entry_twice_in_a_row(gpu):
    # call 1: 70 001 requests -> 69 scan partials land in the built-in scratch
    toks, offs = make_batch(70001, seed=1, max_len=200)
    (b0, h0, s0), (b1, h1, s1) = run_both(gpu, toks, offs, 64, 1337)
    assert b0[-1] > 0
    assert np.array_equal(b0, b1) and np.array_equal(h0, h1) and np.array_equal(s0, s1)
    # call 2, straight after: fewer requests and other data, so part of the scratch still holds call 1's sums
    toks2, offs2 = make_batch(66000, seed=2, max_len=150)
    (b0, h0, s0), (b1, h1, s1) = run_both(gpu, toks2, offs2, 64, 99)
    assert np.array_equal(b0, b1) and np.array_equal(h0, h1) and np.array_equal(s0, s1)
    # a generic-path block size and the twin flag go through the same scratch
    (b0, h0, s0), (b1, h1, s1) = run_both(gpu, toks2, offs2, 16, 7, flags=gpu.HASH_TWIN_FIRST)
    assert np.array_equal(b0, b1) and np.array_equal(h0, h1) and np.array_equal(s0, s1)


def test_noalloc_small_batch_and_known_answer(gpu):
    # small batches never touch the scratch (single-CTA scan); the reference's known answers must still come out
    toks = np.arange(1, 11, dtype=np.uint32)
    offs = np.array([0, 10], dtype=np.int64)
    (_, _, _), (boff, bh, sh) = run_both(gpu, toks, offs, 4, 1337)
    assert list(boff) == [0, 2]
    assert [int(x) for x in bh] == [14643705804678351452, 16777012769546811212]   # lib/tokens/src/lib.rs:539-545
    assert int(sh[1]) == 4945711292740353085


def test_noalloc_rejects_batches_beyond_its_limit(gpu):
    L = gpu.lib()
    d = DevBuf(gpu, 256)
    too_many = gpu.HASH_NOALLOC_MAX_REQUESTS + 1
    rc = L.kvdp_hash_token_blocks_noalloc(d.ptr, d.ptr, too_many, 0, 64, 0, d.ptr, d.ptr, d.ptr, 0, None)
    assert rc == gpu.KVDP_ERANGE
    msg = gpu.last_error()
    assert "at most" in msg and str(gpu.HASH_NOALLOC_MAX_REQUESTS) in msg
