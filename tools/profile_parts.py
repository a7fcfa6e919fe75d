"""Two launches of the short-block hash kernel (16-token blocks, router mix) for an ncu capture."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamo_exp_b200 import _lib as lib  # noqa: E402

L = lib.lib()
isls = [512, 1024, 2048, 4096, 8192, 16384, 32768]
lens = np.array([isls[i % 7] for i in range(7 * 2048)], dtype=np.int64)
offs = np.zeros(lens.size + 1, dtype=np.int64)
offs[1:] = np.cumsum(lens)
total = int(offs[-1])
for bs in (16, 32):
    nblk = int((lens // bs).sum())
    tok = torch.randint(0, 128256, (total,), dtype=torch.int32, device="cuda")
    d_off = torch.from_numpy(offs).cuda()
    d_boff = torch.empty(lens.size + 1, dtype=torch.int64, device="cuda")
    d_bh = torch.empty(nblk, dtype=torch.int64, device="cuda")
    ws = L.kvdp_hash_workspace_bytes(lens.size)
    d_ws = torch.empty(ws, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        lib.check(L.kvdp_hash_token_blocks(tok.data_ptr(), d_off.data_ptr(), lens.size, total, bs, 1337, d_boff.data_ptr(),
                                           d_bh.data_ptr(), None, d_ws.data_ptr(), ws, 0, None))
    torch.cuda.synchronize()
print("done")
