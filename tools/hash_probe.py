"""Timing probe for the hashing path: isolates scan / block-hash / chain costs by varying the batch shape."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamo_exp_b200 import _lib as lib  # noqa: E402

L = lib.lib()
BS = 64


def run(lens, chained, label, iters=20, BS=BS):
    lens = np.asarray(lens, dtype=np.int64)
    n_req = lens.size
    offs = np.zeros(n_req + 1, dtype=np.int64)
    offs[1:] = np.cumsum(lens)
    total = int(offs[-1])
    nblk = int((lens // BS).sum())
    tok = torch.randint(0, 128256, (total,), dtype=torch.int32, device="cuda")
    d_off = torch.from_numpy(offs).cuda()
    d_boff = torch.empty(n_req + 1, dtype=torch.int64, device="cuda")
    d_bh = torch.empty(nblk, dtype=torch.int64, device="cuda")
    d_sh = torch.empty(nblk, dtype=torch.int64, device="cuda")
    ws_bytes = L.kvdp_hash_workspace_bytes(n_req)
    d_ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")

    def call():
        lib.check(L.kvdp_hash_token_blocks(tok.data_ptr(), d_off.data_ptr(), n_req, total, BS, 1337, d_boff.data_ptr(),
                                           d_bh.data_ptr(), d_sh.data_ptr() if chained else None, d_ws.data_ptr(),
                                           ws_bytes, 0, None))
    for _ in range(3):
        call()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        call()
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1e3 / iters
    print(f"{label:46s} n_req={n_req:7d} blocks={nblk:8d}  {us:8.1f} us/batch  {nblk / us * 1e-3:7.2f} G hashes/s  "
          f"{nblk * 272 / us * 1e-3:7.1f} GB/s")


isls = [512, 1024, 2048, 4096, 8192, 16384, 32768]
mix = [isls[i % 7] for i in range(7 * 2048)]
tot = sum(mix)
run(mix, False, "router mix, un-chained")
run(mix, True, "router mix, chained")
run([tot], False, "ONE request, same tokens, un-chained")
run([4096] * (tot // 4096), False, "all ISL 4k, un-chained")
run([4096] * (tot // 4096), True, "all ISL 4k, chained")
run([512] * (tot // 512), True, "all ISL 512, chained")
run([32768] * (tot // 32768), True, "all ISL 32k, chained")
run([64] * 70000, False, "70k one-block requests (3-kernel scan)")
run(mix * 4, False, "router mix x4, un-chained")
run(mix * 4, True, "router mix x4, chained")
big = mix * 4
run(big, True, "router mix x4, chained")
run(sorted(big), True, "router mix x4 SORTED by length, chained")
run(sorted(big), False, "router mix x4 SORTED by length, un-chained")

for bs in (16, 32, 64, 128, 256):
    run(mix, False, f"router mix, un-chained, block_size={bs}", BS=bs)
    run(mix, True, f"router mix, chained, block_size={bs}", BS=bs)
