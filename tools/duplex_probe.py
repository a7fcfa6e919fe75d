"""PCIe duplex probe: which mix of engines moves scattered blocks both ways at once fastest?
kernel = legacy copy_stream_memcpy on pinned memory (SMs read / write host memory), dma = one contiguous cudaMemcpyAsync."""
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
layers, nb, chunk = 8, 512, 128 * 1024
layer_bytes = 2 * nb * chunk


def pinned(n):
    out = []
    for _ in range(n):
        p = C.c_void_p()
        lib.check(L.cuda_malloc_host(C.byref(p), layer_bytes))
        C.memset(p.value, 1, layer_bytes)
        out.append(p.value)
    return out


h_in, h_out = pinned(layers), pinned(layers)
d_in = [torch.zeros(layer_bytes, dtype=torch.uint8, device="cuda") for _ in range(layers)]
d_out = [torch.ones(layer_bytes, dtype=torch.uint8, device="cuda") for _ in range(layers)]
cs_a, cs_b = C.c_void_p(), C.c_void_p()
assert L.copy_stream_create(C.byref(cs_a), layers, nb) == 0
assert L.copy_stream_create(C.byref(cs_b), layers, nb) == 0
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
rng = np.random.Generator(np.random.PCG64(1))
payload = layers * layer_bytes


def run(h2d, d2h, ctas=0):
    lib.set_option("host_ctas", ctas)
    best = 0
    for rep in range(3):
        ids = [rng.permutation(nb).astype(np.int32) for _ in range(4)]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if h2d == "kernel":
            lib.check(L.copy_stream_prepare_block_ids(cs_a, ids[0].ctypes.data, ids[1].ctypes.data, nb))
        if d2h == "kernel":
            lib.check(L.copy_stream_prepare_block_ids(cs_b, ids[2].ctypes.data, ids[3].ctypes.data, nb))
        for l in range(layers):
            if h2d == "kernel":
                lib.check(L.copy_stream_memcpy(cs_a, h_in[l], d_in[l].data_ptr(), 2, chunk // 2, 2, nb, nb))
            elif h2d == "dma":
                lib.check(L.cuda_memcpy_async(d_in[l].data_ptr(), h_in[l], layer_bytes, C.c_void_p(s1.cuda_stream)))
            if d2h == "kernel":
                lib.check(L.copy_stream_memcpy(cs_b, d_out[l].data_ptr(), h_out[l], 2, chunk // 2, 2, nb, nb))
            elif d2h == "dma":
                lib.check(L.cuda_memcpy_async(h_out[l], d_out[l].data_ptr(), layer_bytes, C.c_void_p(s2.cuda_stream)))
        if h2d == "kernel":
            lib.check(L.copy_stream_sync(cs_a))
        if d2h == "kernel":
            lib.check(L.copy_stream_sync(cs_b))
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        n = (h2d is not None) + (d2h is not None)
        if rep:
            best = max(best, n * payload / dt / GB)
    print("h2d=%-6s d2h=%-6s host_ctas=%3d  total %.1f GB/s" % (h2d, d2h, ctas, best), flush=True)


def run_chunked(chunk_mib):
    """Both directions with plain DMAs of `chunk_mib` MiB each: what the copy engines reach at that DMA size."""
    n = chunk_mib << 20
    best = 0
    for rep in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for l in range(layers):
            for o in range(0, layer_bytes, n):
                lib.check(L.cuda_memcpy_async(d_in[l].data_ptr() + o, h_in[l] + o, n, C.c_void_p(s1.cuda_stream)))
                lib.check(L.cuda_memcpy_async(h_out[l] + o, d_out[l].data_ptr() + o, n, C.c_void_p(s2.cuda_stream)))
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if rep:
            best = max(best, 2 * payload / dt / GB)
    print("dma both ways, %3d MiB per DMA: total %.1f GB/s" % (chunk_mib, best), flush=True)


for mib in (1, 2, 8, 32, 128):
    run_chunked(mib)
for combo in ((None, "kernel"), ("kernel", None), ("dma", None), (None, "dma"), ("dma", "dma"), ("kernel", "kernel"),
              ("kernel", "dma"), ("dma", "kernel")):
    run(*combo)
for ctas in (8, 16, 64, 148):
    run("kernel", "kernel", ctas)
lib.set_option("host_ctas", 0)
