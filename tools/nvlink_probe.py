"""NVLink probe, ONE process driving two GPUs (so that it can run under ncu, which must not wrap a multi-rank command).

  python tools/nvlink_probe.py launches   # a handful of push / pull launches of each copy path: the ncu target
  python tools/nvlink_probe.py sweep      # event-timed GB/s of the push over tile / CTAs-per-SM / store-hint settings

GPU 0 runs every kernel. push = GPU-0 kernel reading its own HBM and STORING into GPU 1 (bytes leave on NVLTX of GPU 0);
pull = GPU-0 kernel LOADING from GPU 1 into its own HBM (bytes arrive on NVLRX of GPU 0). Workloads: the Llama-3-8B
layer-separate -> contiguous relayout (128 KiB chunks, paths 3 and 1) and the Llama-3-70B TP4 -> TP2 reshard (512-byte
runs, path 2). Under ncu:
  ncu --metrics nvltx__bytes.sum,nvltx__bytes_data_user.sum,nvltx__bytes_data_protocol.sum,nvlrx__bytes.sum,\
nvlrx__bytes_data_user.sum,nvlrx__bytes_data_protocol.sum,gpu__time_duration.sum -k regex:kvdp_copy ...
"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamo_exp_b200 import CopyPlan, _lib as lib  # noqa: E402

L = lib.lib()
GB = 1e9
LAYERS, CHUNK = 32, 64 * 8 * 128 * 2
NB = 256                      # blocks per launch: 2 GiB payload


def alloc(nbytes, dev):
    p = C.c_void_p()
    lib.check(L.kvdp_device_alloc(C.byref(p), nbytes, dev))
    return p.value


def setup():
    lib.check(L.kvdp_peer_enable(0, 1))
    lib.check(L.kvdp_peer_enable(1, 0))
    torch.cuda.set_device(0)
    layers0 = [alloc(2 * NB * CHUNK, 0) for _ in range(LAYERS)]
    layers1 = [alloc(2 * NB * CHUNK, 1) for _ in range(LAYERS)]
    contig0 = alloc(NB * LAYERS * 2 * CHUNK, 0)
    contig1 = alloc(NB * LAYERS * 2 * CHUNK, 1)

    def contig_plan(src_layers, dst):
        ptrs = np.array(src_layers, dtype=np.uint64)
        h = C.c_void_p()
        lib.check(L.kvdp_plan_contig(C.byref(h), ptrs.ctypes.data, LAYERS, 0, NB, CHUNK, C.c_void_p(dst), NB, 1, 0))
        return h
    push = contig_plan(layers0, contig1)        # reads GPU 0, stores into GPU 1
    pull = contig_plan(layers1, contig0)        # loads from GPU 1, stores into GPU 0
    # 70B TP4 -> TP2 reshard of prefill rank 1 (2 heads = 512-byte runs into 4-head rows), 80 layers folded into 32
    hp, bs, d, e = 2, 64, 128, 2
    src70 = [alloc(2 * NB * bs * hp * d * e, 0) for _ in range(LAYERS)]
    dst70 = [alloc(2 * NB * bs * 2 * hp * d * e, 1) for _ in range(LAYERS)]
    reshard = CopyPlan.reshard(src70, [dst70], NB, NB, bs, hp * d * e, 4, 2, 1, 0)
    g = torch.Generator(device="cpu").manual_seed(1)
    sid = torch.randperm(NB, generator=g).to(torch.int32).cuda()
    did = torch.randperm(NB, generator=g).to(torch.int32).cuda()
    return push, pull, reshard, sid, did


def launch(plan, sid, did, stream):
    h = plan.handle if hasattr(plan, "handle") else plan
    lib.check(L.kvdp_plan_launch(h, sid.data_ptr(), did.data_ptr(), NB, 0, LAYERS, C.c_void_p(stream.cuda_stream)))


def timed(plan, sid, did, stream, reps=5):
    for _ in range(2):
        launch(plan, sid, did, stream)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(reps):
        launch(plan, sid, did, stream)
    b.record(stream)
    torch.cuda.synchronize()
    return a.elapsed_time(b) / 1e3 / reps


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "launches"
    push, pull, reshard, sid, did = setup()
    st = torch.cuda.Stream()
    payload = NB * LAYERS * 2 * CHUNK
    payload70 = NB * LAYERS * 2 * 64 * 2 * 128 * 2
    if mode == "nvml":
        # NVML's per-GPU NVLink byte counters (payload vs raw, i.e. with packet headers) around kernels AND around the copy
        # engine, which ncu cannot see. Cumulative KiB counters, read on GPU 0 before / after a burst of each.
        import pynvml
        pynvml.nvmlInit()
        h0 = pynvml.nvmlDeviceGetHandleByIndex(int(os.environ.get("CUDA_VISIBLE_DEVICES", "0").split(",")[0]))
        ids = [pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_TX,
               pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX, pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_RX]

        def read():
            vals = pynvml.nvmlDeviceGetFieldValues(h0, [(i, 0xFFFFFFFF) for i in ids])
            return [int(v.value.ullVal) * 1024 for v in vals]
        src = torch.empty(payload, dtype=torch.uint8, device="cuda:0")
        dst = torch.empty(payload, dtype=torch.uint8, device="cuda:1")
        print("burst                                   payload_B   data_tx_B    raw_tx_B  raw/data_tx   data_rx_B    raw_rx_B  raw/data_rx")
        for name, fn, nbytes in (
                ("push path 3 (bulk stores) x8", lambda: [launch(push, sid, did, st) for _ in range(8)], 8 * payload),
                ("pull path 3 (bulk loads) x8", lambda: [launch(pull, sid, did, st) for _ in range(8)], 8 * payload),
                ("cudaMemcpyAsync GPU0->GPU1 x8", lambda: [lib.check(L.cuda_memcpy_async(
                    dst.data_ptr(), src.data_ptr(), payload, C.c_void_p(st.cuda_stream))) for _ in range(8)], 8 * payload),
                ("cudaMemcpyAsync GPU1->GPU0 x8", lambda: [lib.check(L.cuda_memcpy_async(
                    src.data_ptr(), dst.data_ptr(), payload, C.c_void_p(st.cuda_stream))) for _ in range(8)], 8 * payload)):
            lib.set_option("copy_path", 3)
            torch.cuda.synchronize()
            import time
            time.sleep(0.3)
            b = read()
            fn()
            torch.cuda.synchronize()
            time.sleep(0.3)
            a = read()
            d = [x - y for x, y in zip(a, b)]
            print("%-36s %12d %11d %11d  %10.4f %11d %11d  %10.4f" %
                  (name, nbytes, d[0], d[1], d[1] / max(d[0], 1), d[2], d[3], d[3] / max(d[2], 1)), flush=True)
        lib.set_option("copy_path", 0)
        return
    if mode == "launches":
        for name, plan, path, bytes_ in (("push path 3 (bulk stores)", push, 3, payload), ("pull path 3 (bulk loads)", pull, 3, payload),
                                         ("push path 1 (16-byte stores)", push, 1, payload),
                                         ("pull path 1 (16-byte loads)", pull, 1, payload),
                                         ("push reshard path 2 (512-byte runs)", reshard, 0, payload70)):
            lib.set_option("copy_path", path)
            sec = timed(plan, sid, did, st, reps=2)
            print("%-40s %8.1f GB/s payload per direction" % (name, bytes_ / sec / GB), flush=True)
        lib.set_option("copy_path", 0)
        return
    # sweep of the push
    print("push sweep: path tile ctas/sm l2_hint vec32  GB/s")
    rows = []
    for path, tile, cps, hint, v32 in ([(3, t, c, h, 0) for t in (8192, 16384, 24576) for c in (1, 2, 4, 8) for h in (0, 2)] +
                                       [(1, t, c, 0, v) for t in (16384, 32768, 65536) for c in (2, 4, 8) for v in (0, 1)] +
                                       [(2, t, c, 0, 0) for t in (16384, 32768) for c in (1, 2, 4)]):
        try:
            for k, v in (("copy_path", path), ("tile_bytes", tile), ("ctas_per_sm", cps), ("l2_hint", hint), ("vec32", v32)):
                lib.set_option(k, v)
            sec = timed(push, sid, did, st, reps=3)
            rows.append((path, tile, cps, hint, v32, payload / sec / GB))
        except Exception as ex:  # noqa: BLE001
            print("failed", path, tile, cps, hint, v32, ex)
    for k in ("copy_path", "tile_bytes", "ctas_per_sm", "l2_hint", "vec32"):
        lib.set_option(k, 0)
    for r in sorted(rows, key=lambda r: -r[5]):
        print("%4d %6d %3d %2d %2d  %7.1f" % r)
    # copy-engine reference on the same buffers
    src = torch.empty(payload, dtype=torch.uint8, device="cuda:0")
    dst = torch.empty(payload, dtype=torch.uint8, device="cuda:1")
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(2):
        a.record()
        lib.check(L.cuda_memcpy_async(dst.data_ptr(), src.data_ptr(), payload, None))
        b.record()
        torch.cuda.synchronize()
    print("cudaMemcpyAsync GPU0 -> GPU1: %.1f GB/s" % (payload / (a.elapsed_time(b) / 1e3) / GB))


if __name__ == "__main__":
    main()
