"""Several GPUs of one box against the host at once: half of them upload (pinned host -> HBM), half download, plain DMAs.
What does the host sustain, and does the DMA size on either side matter? (full-trip leg of bench.py --gpus N)"""
import ctypes as C
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamo_exp_b200 import _lib as lib  # noqa: E402

L = lib.lib()
GB = 1e9
n = torch.cuda.device_count()
SIZE = 1 << 30
host, dev, streams, streams2 = [], [], [], []
for g in range(n):
    torch.cuda.set_device(g)
    p = C.c_void_p()
    lib.check(L.cuda_malloc_host(C.byref(p), SIZE))
    C.memset(p.value, g + 1, SIZE)
    host.append(p.value)
    dev.append(torch.zeros(SIZE, dtype=torch.uint8, device="cuda:%d" % g))
    streams.append(torch.cuda.Stream(device=g))
    streams2.append(torch.cuda.Stream(device=g))


def sync():
    for g in range(n):
        torch.cuda.synchronize(g)


def run(label, up, down, piece_up, piece_down, reps=4):
    best = 0.0
    for rep in range(3):
        sync()
        t0 = time.perf_counter()
        for r in range(reps):
            for o in range(0, SIZE, min(piece_up, piece_down)):
                for g in up:
                    if o % piece_up == 0:
                        torch.cuda.set_device(g)
                        lib.check(L.cuda_memcpy_async(dev[g].data_ptr() + o, host[g] + o, piece_up,
                                                      C.c_void_p(streams[g].cuda_stream)))
                for g in down:
                    if o % piece_down == 0:
                        torch.cuda.set_device(g)
                        lib.check(L.cuda_memcpy_async(host[g] + o, dev[g].data_ptr() + o, piece_down,
                                                      C.c_void_p(streams2[g].cuda_stream)))
        sync()
        dt = time.perf_counter() - t0
        best = max(best, (len(up) + len(down)) * reps * SIZE / dt / GB)
    print("%-64s %7.1f GB/s total, %5.1f per GPU" % (label, best, best / max(1, len(up) + len(down))), flush=True)


M = 1 << 20
print("# %d GPUs, 1 GiB per GPU and repetition, pinned host memory allocated by each GPU's context" % n)
every = list(range(n))
run("all upload, 1 GiB DMAs", every, [], SIZE, SIZE)
run("all download, 1 GiB DMAs", [], every, SIZE, SIZE)
if n >= 2:
    ups, downs = every[0::2], every[1::2]        # even ranks are the prefill side in bench_multi
    run("even upload + odd download, 1 GiB / 1 GiB", ups, downs, SIZE, SIZE)
    run("even upload + odd download, 8 MiB / 1 GiB", ups, downs, 8 * M, SIZE)
    run("even upload + odd download, 8 MiB / 8 MiB", ups, downs, 8 * M, 8 * M)
    run("even upload + odd download, 32 MiB / 32 MiB", ups, downs, 32 * M, 32 * M)
    run("first half upload + second half download, 8 MiB / 8 MiB", every[:n // 2], every[n // 2:], 8 * M, 8 * M)
run("every GPU both ways at once, 8 MiB / 8 MiB", every, every, 8 * M, 8 * M)
print(os.popen("nvidia-smi topo -m 2>&1 | head -14").read())
