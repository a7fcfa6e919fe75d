"""Timing probe for small-chunk copies (token-granular rows): copy paths x row sizes, one GPU, device to device."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamo_exp_b200 import _lib as lib  # noqa: E402
from dynamo_exp_b200.transfer import token_rows_plan  # noqa: E402

L = lib.lib()
LAYERS, SLOTS, NTOK = 32, 1024 * 64, 8 * 4096


def run(row, col, s_off, d_off, path, label):
    src = [torch.empty(2 * SLOTS * row, dtype=torch.uint8, device="cuda").random_(0, 256) for _ in range(LAYERS)]
    dst = [torch.zeros(2 * SLOTS * row, dtype=torch.uint8, device="cuda") for _ in range(LAYERS)]
    plan = token_rows_plan([t.data_ptr() for t in src], [t.data_ptr() for t in dst], SLOTS, SLOTS, row, row, col,
                           s_off, d_off)
    g = torch.Generator(device="cpu").manual_seed(9)
    s_ids = torch.randperm(SLOTS, generator=g)[:NTOK].to(torch.int32).cuda()
    d_ids = torch.randperm(SLOTS, generator=g)[:NTOK].to(torch.int32).cuda()
    lib.check(L.kvdp_set_option(b"copy_path", path))
    st = torch.cuda.Stream()

    def call():
        lib.check(L.kvdp_plan_launch(plan.handle, s_ids.data_ptr(), d_ids.data_ptr(), NTOK, 0, LAYERS,
                                     C.c_void_p(st.cuda_stream)))
    try:
        for _ in range(2):
            call()
    except lib.KvdpError as e:
        print(f"{label:40s} path={path}: {e}")
        return
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for _ in range(5):
        call()
    b.record(st)
    torch.cuda.synchronize()
    sec = a.elapsed_time(b) / 1e3 / 5
    payload = NTOK * col * 2 * LAYERS
    ok = all(torch.equal(src[l][(kv * SLOTS + int(s_ids[i])) * row + s_off:(kv * SLOTS + int(s_ids[i])) * row + s_off + col],
                         dst[l][(kv * SLOTS + int(d_ids[i])) * row + d_off:(kv * SLOTS + int(d_ids[i])) * row + d_off + col])
             for l in (0, LAYERS - 1) for kv in (0, 1) for i in (0, NTOK - 1))
    print(f"{label:40s} path={path}: {sec * 1e3:7.3f} ms  {2 * payload / sec / 1e9:7.1f} GB/s r+w  ok={ok}")
    plan.close()
    lib.check(L.kvdp_set_option(b"copy_path", 0))




for path in (0, 1, 3, 4):
    run(2048, 2048, 0, 0, path, "2 KiB rows, whole row")
for path in (0, 1, 4):
    run(2048, 512, 1024, 0, path, "2 KiB rows -> 512 B head slice")
for path in (0, 1, 3, 4):
    run(512, 512, 0, 0, path, "512 B rows, whole row")
for path in (0, 1, 3, 4):
    run(4096, 4096, 0, 0, path, "4 KiB rows, whole row")
for path in (0, 1):
    run(256, 256, 0, 0, path, "256 B rows, whole row")
for path in (0, 1):
    run(2048, 2048 - 256, 128, 128, path, "2 KiB rows, 1792 B slice at +128")
for path in (0, 1):
    run(2050, 2050, 0, 0, path, "2050 B rows (2-byte aligned)")
