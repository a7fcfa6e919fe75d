"""Timing probe: TP head reshard on ONE GPU (HBM-bound), Llama-3-70B geometry, copy paths 1 and 2."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamo_exp_b200 import _lib as lib  # noqa: E402

L = lib.lib()
LAYERS, H, D, BS, E = 80, 8, 128, 64, 2
N = 256


def run(src_tp, dst_tp, path, tile=0, cps=0, stages=0):
    lib.set_option("copy_path", path)
    lib.set_option("tile_bytes", tile)
    lib.set_option("ctas_per_sm", cps)
    lib.set_option("stages", stages)
    hs, hd = H // src_tp, H // dst_tp
    src = [torch.empty(2 * N * BS * hs * D * E, dtype=torch.uint8, device="cuda").random_(0, 256) for _ in range(LAYERS)]
    ntgt = max(1, dst_tp // src_tp)
    dst = [[torch.zeros(2 * N * BS * hd * D * E, dtype=torch.uint8, device="cuda") for _ in range(LAYERS)]
           for _ in range(ntgt)]
    sp = np.array([t.data_ptr() for t in src], dtype=np.uint64)
    dp = np.array([t.data_ptr() for tg in dst for t in tg], dtype=np.uint64)
    plan = C.c_void_p()
    lib.check(L.kvdp_plan_reshard(C.byref(plan), sp.ctypes.data, dp.ctypes.data, LAYERS, N, N, BS, hs * D * E, src_tp,
                                  dst_tp, 0, 0))
    ids = torch.randperm(N, dtype=torch.int32, device="cuda")
    ids2 = torch.randperm(N, dtype=torch.int32, device="cuda")
    st = torch.cuda.Stream()
    for _ in range(2):
        lib.check(L.kvdp_plan_launch(plan, ids.data_ptr(), ids2.data_ptr(), N, 0, LAYERS, C.c_void_p(st.cuda_stream)))
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    it = 10
    for _ in range(it):
        lib.check(L.kvdp_plan_launch(plan, ids.data_ptr(), ids2.data_ptr(), N, 0, LAYERS, C.c_void_p(st.cuda_stream)))
    b.record(st)
    torch.cuda.synchronize()
    sec = a.elapsed_time(b) / 1e3 / it
    payload = LAYERS * 2 * N * BS * hs * D * E
    # spot check: block ids[0] layer 3 K, first token row
    i = 0
    s_id, d_id = int(ids[i]), int(ids2[i])
    srow = src[3].view(2, N, BS, hs * D * E)[0, s_id, 5]
    if dst_tp >= src_tp:
        for g in range(ntgt):
            drow = dst[g][3].view(2, N, BS, hd * D * E)[0, d_id, 5]
            assert torch.equal(drow, srow[g * hd * D * E:(g + 1) * hd * D * E])
    else:
        drow = dst[0][3].view(2, N, BS, hd * D * E)[0, d_id, 5]
        assert torch.equal(drow[:hs * D * E], srow)
    L.kvdp_plan_destroy(plan)
    print(f"TP{src_tp}->TP{dst_tp} path {path} tile {tile:6d} cps {cps} stages {stages}: {2 * payload / sec / 1e9:8.1f} GB/s r+w "
          f"(run {min(hs, hd) * D * E} B)", flush=True)


for tp in ((4, 2), (2, 4), (1, 8), (8, 1)):
    for path, tile, cps, stages in ((1, 0, 0, 0), (1, 16384, 8, 0), (2, 0, 0, 0), (2, 32768, 2, 3), (2, 16384, 4, 3),
                                    (2, 8192, 6, 4)):
        run(tp[0], tp[1], path, tile, cps, stages)
