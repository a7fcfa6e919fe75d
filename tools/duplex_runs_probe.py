"""CopyStream both-ways step time against the number of host-side runs per direction (controlled id sets), and against
the size of the pieces the runs are issued in (kvdp_set_option("host_dma_piece_kib")). profiles/duplex_r02.txt."""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamo_exp_b200 import _lib as lib  # noqa: E402

L = lib.lib()
GB = 1e9
LAYERS, REQ, BPR, CHUNK = 32, 8, 64, 128 * 1024
pairs = REQ * BPR
nd, nh = pairs, 2 * pairs
layer_bytes, host_layer_bytes = 2 * nd * CHUNK, 2 * nh * CHUNK
payload = LAYERS * layer_bytes


def pinned_layers():
    out = []
    for _ in range(LAYERS):
        p = C.c_void_p()
        lib.check(L.cuda_malloc_host(C.byref(p), host_layer_bytes))
        C.memset(p.value, 1, host_layer_bytes)
        out.append(p.value)
    return out


h_in, h_out = pinned_layers(), pinned_layers()
d_in = [torch.zeros(layer_bytes, dtype=torch.uint8, device="cuda") for _ in range(LAYERS)]
d_out = [torch.ones(layer_bytes, dtype=torch.uint8, device="cuda") for _ in range(LAYERS)]
rng = np.random.Generator(np.random.PCG64(5))
cs_on, cs_off = C.c_void_p(), C.c_void_p()
assert L.copy_stream_create(C.byref(cs_on), LAYERS, pairs) == 0
assert L.copy_stream_create(C.byref(cs_off), LAYERS, pairs) == 0
suffix = CHUNK // 2


def slots_with_runs(r):
    """8 request slots out of 16 forming exactly r runs: r groups of 8/r adjacent slots, one free slot between groups."""
    g = REQ // r
    out, s = [], 0
    for _ in range(r):
        out += list(range(s, s + g))
        s += g + 1
    assert max(out) < nh // BPR, (r, out)
    return out


def hid(r):
    return np.concatenate([np.arange(s * BPR, (s + 1) * BPR) for s in rng.permutation(slots_with_runs(r))]).astype(np.int32)


def step(r_on, r_off, on=True, off=True):
    ids = [hid(r_on), rng.permutation(nd).astype(np.int32), rng.permutation(nd).astype(np.int32), hid(r_off)]
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    if on:
        lib.check(L.copy_stream_prepare_block_ids(cs_on, ids[0].ctypes.data, ids[1].ctypes.data, pairs))
    if off:
        lib.check(L.copy_stream_prepare_block_ids(cs_off, ids[2].ctypes.data, ids[3].ctypes.data, pairs))
    for l in range(LAYERS):
        if on:
            lib.check(L.copy_stream_memcpy(cs_on, h_in[l], d_in[l].data_ptr(), 2, suffix, 2, nh, nd))
        if off:
            lib.check(L.copy_stream_memcpy(cs_off, d_out[l].data_ptr(), h_out[l], 2, suffix, 2, nd, nh))
    t1 = time.perf_counter()
    if on:
        lib.check(L.copy_stream_sync(cs_on))
    if off:
        lib.check(L.copy_stream_sync(cs_off))
    t2 = time.perf_counter()
    return (t2 - t0) * 1e3, (t1 - t0) * 1e3


for _ in range(2):
    step(1, 1)
print("runs_on runs_off   step_ms(mean of 4)  issue_ms   GB/s")
for r_on, r_off in ((1, 1), (2, 2), (4, 4), (8, 8), (1, 8), (8, 1), (1, 1), (4, 4), (8, 8)):
    ts = [step(r_on, r_off) for _ in range(4)]
    m = float(np.mean([t[0] for t in ts]))
    print("%4d %7d   %8.2f  %14.2f   %6.1f   %s" % (r_on, r_off, m, float(np.mean([t[1] for t in ts])),
                                                 2 * payload / m / 1e6, [round(t[0], 1) for t in ts]), flush=True)
for kib in (0, 2048, 4096, 8192, 16384):
    lib.set_option("host_dma_piece_kib", kib)
    for r_on, r_off in ((8, 1), (1, 8), (1, 1), (8, 8)):
        ts = [step(r_on, r_off) for _ in range(3)]
        m = float(np.mean([t[0] for t in ts]))
        print("piece %5d KiB  runs %d/%d  %8.2f ms  %6.1f GB/s" % (kib, r_on, r_off, m, 2 * payload / m / 1e6), flush=True)
lib.set_option("host_dma_piece_kib", 8192)
print("one direction at a time")
for r in (1, 4, 8):
    a = float(np.mean([step(r, r, True, False)[0] for _ in range(3)]))
    b = float(np.mean([step(r, r, False, True)[0] for _ in range(3)]))
    print("runs %d: onboard alone %.2f ms (%.1f GB/s)   offload alone %.2f ms (%.1f GB/s)" %
          (r, a, payload / a / 1e6, b, payload / b / 1e6), flush=True)
