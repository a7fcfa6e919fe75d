"""GPU: kvdp_hash_token_blocks_noalloc — the hashing entry point for callers that cannot allocate device memory
(for instance a router thread inside an engine that gave the whole GPU to its KV cache at start-up).

It must produce exactly what kvdp_hash_token_blocks produces with a caller-supplied workspace — and what the CPU oracle
produces — call after call, even though its scratch is one array shared by every call and still holds the previous
call's partial sums; and it must do so without taking a single byte of device memory. The batches here hold tens of
thousands of requests on purpose: only above 2 048 requests does the block-count scan go through the scratch at all.
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


def run_both(gpu, toks, offs, block_size, salt, flags=0, device=-1):
    """Hash one batch through the existing entry point (own workspace) and through the no-alloc one."""
    L = gpu.lib()
    n_req, total = offs.size - 1, int(offs[-1])
    nblk = int(((offs[1:] - offs[:-1]) // block_size).sum())
    d_tok, d_off = DevBuf(gpu, max(toks.nbytes, 16), device), DevBuf(gpu, offs.nbytes, device)
    d_tok.upload(toks)
    d_off.upload(offs)
    ws_bytes = L.kvdp_hash_workspace_bytes(n_req)
    d_ws = DevBuf(gpu, ws_bytes, device)
    out = []
    for which in ("workspace", "noalloc"):
        d_boff, d_bh, d_sh = (DevBuf(gpu, (n_req + 1) * 8, device), DevBuf(gpu, max(nblk, 1) * 8, device),
                              DevBuf(gpu, max(nblk, 1) * 8, device))
        if which == "workspace":
            rc = L.kvdp_hash_token_blocks(d_tok.ptr, d_off.ptr, n_req, total, block_size, salt, d_boff.ptr, d_bh.ptr,
                                          d_sh.ptr, d_ws.ptr, ws_bytes, flags, None)
        else:
            rc = L.kvdp_hash_token_blocks_noalloc(d_tok.ptr, d_off.ptr, n_req, total, block_size, salt, d_boff.ptr,
                                                  d_bh.ptr, d_sh.ptr, flags, None)
        gpu.check(rc)
        out.append((d_boff.download(np.int64), d_bh.download(np.uint64)[:nblk], d_sh.download(np.uint64)[:nblk]))
    return out


def same(a, b):
    return all(np.array_equal(x, y) for x, y in zip(a, b))


def test_noalloc_matches_workspace_entry_and_the_oracle_twice_in_a_row(gpu, oracle):
    # call 1: 70 001 requests -> 69 scan partials land in the built-in scratch
    toks, offs = make_batch(70001, seed=1, max_len=200)
    ws, na = run_both(gpu, toks, offs, 64, 1337)
    assert ws[0][-1] > 0
    assert same(ws, na)
    assert same(na, oracle.hash_batch(toks, offs, 64, 1337, oracle.hw_threads()))
    # call 2, straight after: fewer requests and other data, so part of the scratch still holds call 1's sums
    toks2, offs2 = make_batch(66000, seed=2, max_len=150)
    ws, na = run_both(gpu, toks2, offs2, 64, 99)
    assert same(ws, na)
    assert same(na, oracle.hash_batch(toks2, offs2, 64, 99, oracle.hw_threads()))
    # 24-token blocks (96 bytes: not one of the specialised sizes, so the generic block kernel runs) plus the twin flag
    ws, na = run_both(gpu, toks2, offs2, 24, 7, flags=gpu.HASH_TWIN_FIRST)
    assert same(ws, na)
    # 16-token blocks: the lane-parts kernel, against the oracle
    ws, na = run_both(gpu, toks2, offs2, 16, 7)
    assert same(ws, na)
    assert same(na, oracle.hash_batch(toks2, offs2, 16, 7, oracle.hw_threads()))


def test_noalloc_takes_no_device_memory(gpu):
    """The property the entry point exists for: free device memory, as the driver reports it, is the same before and
    after a call. The first call on a device is the set-up call (symbol lookup, a kernel attribute), so it is made
    before the measurement; inputs and outputs are allocated before it too."""
    import torch
    L = gpu.lib()
    toks, offs = make_batch(70001, seed=3, max_len=200)
    n_req, total = offs.size - 1, int(offs[-1])
    nblk = int(((offs[1:] - offs[:-1]) // 64).sum())
    d_tok, d_off = DevBuf(gpu, toks.nbytes), DevBuf(gpu, offs.nbytes)
    d_tok.upload(toks)
    d_off.upload(offs)
    d_boff, d_bh, d_sh = DevBuf(gpu, (n_req + 1) * 8), DevBuf(gpu, nblk * 8), DevBuf(gpu, nblk * 8)

    def call():
        gpu.check(L.kvdp_hash_token_blocks_noalloc(d_tok.ptr, d_off.ptr, n_req, total, 64, 1337, d_boff.ptr, d_bh.ptr,
                                                   d_sh.ptr, 0, None))
        gpu.check(L.kvdp_stream_synchronize(None))
    call()                                        # set-up call
    torch.cuda.mem_get_info(0)                    # let torch create whatever it creates on first use
    free_before, _ = torch.cuda.mem_get_info(0)
    call()
    call()
    free_after, _ = torch.cuda.mem_get_info(0)
    assert free_after == free_before, (free_before, free_after)
    # the allocating path, for contrast, is visible to the same probe (so the probe can see an allocation at all)
    extra = DevBuf(gpu, 64 << 20)
    free_with_extra, _ = torch.cuda.mem_get_info(0)
    assert free_with_extra < free_before
    extra.free()


def test_noalloc_on_a_second_gpu_resolves_its_own_scratch(gpu, oracle):
    """Every device has its own copy of the built-in scratch; its address is looked up with that device current."""
    import torch
    n = C.c_int(0)
    gpu.check(gpu.lib().kvdp_device_count(C.byref(n)))
    if n.value < 2:
        pytest.skip("one GPU on this box")
    toks, offs = make_batch(30011, seed=4, max_len=180)
    want = oracle.hash_batch(toks, offs, 64, 1337, oracle.hw_threads())
    torch.cuda.init()
    with torch.cuda.device(1):                   # the entry point works on the calling thread's current device
        for _ in range(2):
            ws, na = run_both(gpu, toks, offs, 64, 1337, device=1)
            assert same(ws, na) and same(na, want)
    ws, na = run_both(gpu, toks, offs, 64, 1337, device=0)      # and device 0 still has its own
    assert same(ws, na) and same(na, want)


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
