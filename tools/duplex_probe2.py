"""Second PCIe duplex probe: what separates the legacy CopyStream's both-ways figure (coalesced DMA path, 8 requests of
64 blocks per direction and step) from two large DMAs? The path's DMA pattern is rebuilt here with plain cudaMemcpyAsync
calls in steps of increasing fidelity (addresses -> staging-slot reuse -> the slot events), beside the path itself."""
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
slots_in = [torch.zeros(layer_bytes, dtype=torch.uint8, device="cuda") for _ in range(2)]
slots_out = [torch.ones(layer_bytes, dtype=torch.uint8, device="cuda") for _ in range(2)]
s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
m_in, m_out = torch.cuda.Stream(), torch.cuda.Stream()
rng = np.random.Generator(np.random.PCG64(5))


def request_slots(adjacent):
    if adjacent:
        return np.arange(REQ)
    return np.sort(rng.permutation(nh // BPR)[:REQ])


def dma(dst, src, n, stream):
    lib.check(L.cuda_memcpy_async(dst, src, n, C.c_void_p(stream.cuda_stream)))


def raw(label, runs_per_half, adjacent=False, two_slots=False, events=False, order="layer"):
    """`runs_per_half` DMAs per (layer, K|V) and direction; run size = half a layer / runs_per_half."""
    run_bytes = nd * CHUNK // runs_per_half
    best = 0.0
    for rep in range(4):
        slots_a = request_slots(adjacent) if runs_per_half == REQ else np.arange(runs_per_half)
        slots_b = request_slots(adjacent) if runs_per_half == REQ else np.arange(runs_per_half)
        filled = [[torch.cuda.Event() for _ in range(2)] for _ in range(2)]
        drained = [[torch.cuda.Event() for _ in range(2)] for _ in range(2)]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for l in range(LAYERS):
            slot = l & 1
            dev_a = slots_in[slot].data_ptr() if two_slots else d_in[l].data_ptr()
            dev_b = slots_out[slot].data_ptr() if two_slots else d_out[l].data_ptr()
            if events and l >= 2:
                s_in.wait_event(drained[0][slot])
                m_out.wait_event(drained[1][slot])
            if events:
                filled[1][slot].record(m_out)
                s_out.wait_event(filled[1][slot])
            ops = []
            for p in range(2):
                for r in range(runs_per_half):
                    ho_a = (p * nh + int(slots_a[r]) * (nd // runs_per_half)) * CHUNK
                    ho_b = (p * nh + int(slots_b[r]) * (nd // runs_per_half)) * CHUNK
                    do = (p * nd) * CHUNK + r * run_bytes
                    ops.append(((dev_a + do, h_in[l] + ho_a, run_bytes, s_in), (h_out[l] + ho_b, dev_b + do, run_bytes, s_out)))
            if order == "layer":
                for a, _ in ops:
                    dma(*a)
                for _, b in ops:
                    dma(*b)
            else:
                for a, b in ops:
                    dma(*a)
                    dma(*b)
            if events:
                filled[0][slot].record(s_in)
                m_in.wait_event(filled[0][slot])
                drained[0][slot].record(m_in)
                drained[1][slot].record(s_out)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if rep:
            best = max(best, 2 * payload / dt / GB)
    print("%-78s %6.1f GB/s" % (label, best), flush=True)


def path(label, adjacent, threads=False):
    import threading
    cs_on, cs_off = C.c_void_p(), C.c_void_p()
    assert L.copy_stream_create(C.byref(cs_on), LAYERS, pairs) == 0
    assert L.copy_stream_create(C.byref(cs_off), LAYERS, pairs) == 0
    suffix = CHUNK // 2
    best = 0.0
    for rep in range(4):
        def hid():
            return np.concatenate([np.arange(s * BPR, (s + 1) * BPR) for s in rng.permutation(request_slots(adjacent))]).astype(np.int32)
        ids = [hid(), rng.permutation(nd).astype(np.int32), rng.permutation(nd).astype(np.int32), hid()]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        lib.check(L.copy_stream_prepare_block_ids(cs_on, ids[0].ctypes.data, ids[1].ctypes.data, pairs))
        lib.check(L.copy_stream_prepare_block_ids(cs_off, ids[2].ctypes.data, ids[3].ctypes.data, pairs))

        def on():
            for l in range(LAYERS):
                lib.check(L.copy_stream_memcpy(cs_on, h_in[l], d_in[l].data_ptr(), 2, suffix, 2, nh, nd))

        def off():
            for l in range(LAYERS):
                lib.check(L.copy_stream_memcpy(cs_off, d_out[l].data_ptr(), h_out[l], 2, suffix, 2, nd, nh))
        if threads:
            ts = [threading.Thread(target=on), threading.Thread(target=off)]
            [t.start() for t in ts]
            [t.join() for t in ts]
        else:
            for l in range(LAYERS):
                lib.check(L.copy_stream_memcpy(cs_on, h_in[l], d_in[l].data_ptr(), 2, suffix, 2, nh, nd))
                lib.check(L.copy_stream_memcpy(cs_off, d_out[l].data_ptr(), h_out[l], 2, suffix, 2, nd, nh))
        lib.check(L.copy_stream_sync(cs_on))
        lib.check(L.copy_stream_sync(cs_off))
        dt = time.perf_counter() - t0
        if rep:
            best = max(best, 2 * payload / dt / GB)
    L.copy_stream_destroy(cs_on)
    L.copy_stream_destroy(cs_off)
    print("%-78s %6.1f GB/s" % (label, best), flush=True)


print("# %d layers x %d blocks x K|V x %d KiB per direction; host layer = 2 x %d blocks; best of 3" %
      (LAYERS, pairs, CHUNK // 1024, nh))
raw("raw: 1 DMA of 64 MiB per layer half", 1)
raw("raw: 2 DMAs of 32 MiB per layer half", 2)
raw("raw: 8 DMAs of 8 MiB per layer half, host runs adjacent", REQ, adjacent=True)
raw("raw: 8 DMAs of 8 MiB per layer half, host runs at random request slots", REQ)
raw("raw: same, directions interleaved DMA by DMA", REQ, order="dma")
raw("raw: same, device side = two staging slots", REQ, two_slots=True)
raw("raw: same, two staging slots + the slot events on two more streams", REQ, two_slots=True, events=True)
raw("raw: 1 DMA per layer half, two staging slots + the slot events", 1, two_slots=True, events=True)
path("CopyStream path: host runs adjacent (merge to 1 DMA per layer half)", True)
path("CopyStream path: host runs at random request slots (8 DMAs per layer half)", False)
path("CopyStream path: same, one host thread per direction", False, threads=True)


def fill(kind):
    for l in range(LAYERS):
        w = np.frombuffer((C.c_uint8 * host_layer_bytes).from_address(h_in[l]), dtype=np.uint8).view(np.uint64)
        if kind == "random":
            w[:] = np.arange(w.size, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15) + np.uint64(l)
            d_out[l].random_(0, 256)
        else:
            w[:] = 0
            d_out[l].zero_()
    torch.cuda.synchronize()


for kind in ("random", "zeros", "random"):
    fill(kind)
    print("# payload bytes now: %s" % kind)
    raw("raw: 1 DMA of 64 MiB per layer half", 1)
    raw("raw: 8 DMAs of 8 MiB per layer half, host runs at random request slots", REQ)
    path("CopyStream path: host runs adjacent (merge to 1 DMA per layer half)", True)
    path("CopyStream path: host runs at random request slots (8 DMAs per layer half)", False)
