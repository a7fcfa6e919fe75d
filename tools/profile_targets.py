"""The three dominant kernels, two launches each, for one `ncu --set full -k regex:kvdp_` capture:
relayout (512 Llama-3-8B blocks, layer-separate -> contiguous), router-sweep hashing (57 344 requests, chained) and
token-granular transfer (32 768 tokens x 32 layers, 2 KiB rows)."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamo_exp_b200 import _lib as lib  # noqa: E402
from dynamo_exp_b200.transfer import token_rows_plan  # noqa: E402

L = lib.lib()
LAYERS, CHUNK = 32, 64 * 8 * 128 * 2
st = torch.cuda.Stream()
sp = C.c_void_p(st.cuda_stream)

# 1. relayout
nb = 1024
layers = [torch.empty(2 * nb * CHUNK, dtype=torch.uint8, device="cuda").random_(0, 256) for _ in range(LAYERS)]
contig = torch.empty(512 * LAYERS * 2 * CHUNK, dtype=torch.uint8, device="cuda")
ptrs = np.array([t.data_ptr() for t in layers], dtype=np.uint64)
plan = C.c_void_p()
lib.check(L.kvdp_plan_contig(C.byref(plan), ptrs.ctypes.data, LAYERS, 0, nb, CHUNK, C.c_void_p(contig.data_ptr()), 512, 1, 0))
g = torch.Generator(device="cpu").manual_seed(1)
sid = torch.randperm(nb, generator=g)[:512].to(torch.int32).cuda()
did = torch.randperm(512, generator=g).to(torch.int32).cuda()
for _ in range(2):
    lib.check(L.kvdp_plan_launch(plan, sid.data_ptr(), did.data_ptr(), 512, 0, LAYERS, sp))
torch.cuda.synchronize()
del layers, contig
torch.cuda.empty_cache()

# 2. hashing
isls = [512, 1024, 2048, 4096, 8192, 16384, 32768]
lens = np.array([isls[i % 7] for i in range(7 * 8192)], dtype=np.int64)
offs = np.zeros(lens.size + 1, dtype=np.int64)
offs[1:] = np.cumsum(lens)
total, nblk = int(offs[-1]), int((lens // 64).sum())
tok = torch.randint(0, 128256, (total,), dtype=torch.int32, device="cuda")
d_off = torch.from_numpy(offs).cuda()
d_boff = torch.empty(lens.size + 1, dtype=torch.int64, device="cuda")
d_bh = torch.empty(nblk, dtype=torch.int64, device="cuda")
d_sh = torch.empty(nblk, dtype=torch.int64, device="cuda")
ws = L.kvdp_hash_workspace_bytes(lens.size)
d_ws = torch.empty(ws, dtype=torch.uint8, device="cuda")
for _ in range(2):
    lib.check(L.kvdp_hash_token_blocks(tok.data_ptr(), d_off.data_ptr(), lens.size, total, 64, 1337, d_boff.data_ptr(),
                                       d_bh.data_ptr(), d_sh.data_ptr(), d_ws.data_ptr(), ws, 0, sp))
torch.cuda.synchronize()
del tok
torch.cuda.empty_cache()

# 3. token-granular transfer
row, slots, ntok = 2048, 1024 * 64, 8 * 4096
src = [torch.empty(2 * slots * row, dtype=torch.uint8, device="cuda").random_(0, 256) for _ in range(LAYERS)]
dst = [torch.zeros(2 * slots * row, dtype=torch.uint8, device="cuda") for _ in range(LAYERS)]
tplan = token_rows_plan([t.data_ptr() for t in src], [t.data_ptr() for t in dst], slots, slots, row, row, row)
s_ids = torch.randperm(slots, generator=g)[:ntok].to(torch.int32).cuda()
d_ids = torch.randperm(slots, generator=g)[:ntok].to(torch.int32).cuda()
for _ in range(2):
    lib.check(L.kvdp_plan_launch(tplan.handle, s_ids.data_ptr(), d_ids.data_ptr(), ntok, 0, LAYERS, sp))
torch.cuda.synchronize()
print("profile targets done")
