"""bench_multi.py — the N>1 arm of bench.py: prefill GPUs push KV blocks into decode GPUs over NVLink.

Ranks [0, P) are prefill, ranks [P, N) decode (P = N // 2). Prefill rank r owns a layer-separate Llama-3-8B-shaped cache
and pushes each step's batch (R requests x 64 blocks) straight into the contiguous-per-block cache of decode rank
P + r % D: the relayout kernel's stores go to the peer-mapped destination (CUDA IPC), and the last CTA publishes a
completion flag in the decode GPU's memory. No NCCL on the data path; NCCL send/recv of the same payload is timed
beside it as the comparison the north star names, and a plain cudaMemcpyAsync to the peer as the link's practical ceiling.
"""
from __future__ import annotations

import ctypes as C
import json
import os

import numpy as np

from bench import BLOCK_BYTES, BLOCKS_PER_REQ, CHUNK, GB, L8B, ClockSampler, workload_config
from bench_single import Setup, measured_peaks

NVLINK_PEAK_MEASURED = 770.0   # GB/s per direction per GPU, peer copy (B200_PROFILING.md)
NVLINK_PEAK_NOMINAL = 900.0


class CudaArray:
    """Minimal __cuda_array_interface__ carrier so torch can view memory owned by libkvdp.so."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def run_multi(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist

    from dynamo_exp_b200 import _lib as lib
    L = lib.lib()
    P = world // 2
    D = world - P
    is_prefill = rank < P
    peer = (P + rank % D) if is_prefill else None
    senders = [r for r in range(P) if P + r % D == rank] if not is_prefill else []
    pairs = BLOCKS_PER_REQ * args.requests
    n = args.cache_blocks

    # ---- decode side: destination cache + flag, exported over CUDA IPC
    handles = {"cache": None, "flag": None}
    dst_ptr = C.c_void_p()
    flag_ptr = C.c_void_p()
    if not is_prefill:
        lib.check(L.kvdp_device_alloc(C.byref(dst_ptr), n * BLOCK_BYTES, local_rank))
        lib.check(L.kvdp_device_alloc(C.byref(flag_ptr), 256, local_rank))
        h1 = (C.c_uint8 * lib.IPC_HANDLE_BYTES)()
        h2 = (C.c_uint8 * lib.IPC_HANDLE_BYTES)()
        lib.check(L.kvdp_ipc_export(dst_ptr, h1))
        lib.check(L.kvdp_ipc_export(flag_ptr, h2))
        handles = {"cache": bytes(h1), "flag": bytes(h2)}
    all_handles = [None] * world
    dist.all_gather_object(all_handles, handles)

    setup = None
    peer_dst = C.c_void_p()
    peer_flag = C.c_void_p()
    if is_prefill:
        hb = (C.c_uint8 * lib.IPC_HANDLE_BYTES).from_buffer_copy(all_handles[peer]["cache"])
        fb = (C.c_uint8 * lib.IPC_HANDLE_BYTES).from_buffer_copy(all_handles[peer]["flag"])
        lib.check(L.kvdp_ipc_import(hb, local_rank, C.byref(peer_dst)))
        lib.check(L.kvdp_ipc_import(fb, local_rank, C.byref(peer_flag)))
        setup = Setup(args, torch, lib, device=local_rank, dst_contig_ptr=peer_dst.value, dst_blocks=n)
    dist.barrier()

    steps, warmup = args.steps, args.warmup
    my_stream = torch.cuda.Stream()

    def timed_region(run_step, total_steps):
        """Barrier + sync both sides, device events, MAX over ranks. Decode ranks wait for the last flag value."""
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st = setup.stream if is_prefill else my_stream
        a.record(st)
        run_step()
        b.record(st)
        torch.cuda.synchronize()
        dist.barrier()
        t = torch.tensor([a.elapsed_time(b) / 1e3], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- fused push: relayout kernel storing into the peer cache, in-kernel completion flag
    seq = {"v": 0}

    def push_steps(k0, count):
        def fn():
            if is_prefill:
                for k in range(count):
                    seq["v"] += 1
                    # senders that share a decode rank would need one flag each; with P == D every flag has one writer
                    setup.step(k0 + k, done_flag=peer_flag, done_value=seq["v"])
            else:
                seq["v"] += count
                lib.check(L.kvdp_stream_wait_flag(C.c_void_p(my_stream.cuda_stream), flag_ptr, seq["v"]))
        return fn

    timed_region(push_steps(0, warmup), warmup)
    clocks = ClockSampler(local_rank)
    clocks.start()
    sec = timed_region(push_steps(warmup, steps), steps)
    clk = clocks.stop()
    total_payload = P * pairs * BLOCK_BYTES * steps
    value = total_payload / sec / GB
    per_link = pairs * BLOCK_BYTES * steps / sec / GB * (max(1, P // D))

    # ---- correctness spot check: prefill sends sampled source chunks over NCCL; decode compares with its cache
    ok = torch.ones(1, dtype=torch.int32, device="cuda")
    last = warmup + steps - 1
    samples = 4
    if is_prefill:
        s_ids, d_ids = setup.ids(last)
        rng = np.random.Generator(np.random.PCG64(last))
        for _ in range(samples):
            i, l, kv = int(rng.integers(pairs)), int(rng.integers(L8B)), int(rng.integers(2))
            src = setup.src_layers[l][(kv * n + int(s_ids[i])) * CHUNK:(kv * n + int(s_ids[i]) + 1) * CHUNK]
            meta = torch.tensor([int(d_ids[i]), l, kv], dtype=torch.int64, device="cuda")
            dist.send(meta, peer)
            dist.send(src.contiguous(), peer)
    else:
        cache = torch.as_tensor(CudaArray(dst_ptr.value, n * BLOCK_BYTES), device="cuda")
        for s in senders:
            for _ in range(samples):
                meta = torch.empty(3, dtype=torch.int64, device="cuda")
                buf = torch.empty(CHUNK, dtype=torch.uint8, device="cuda")
                dist.recv(meta, s)
                dist.recv(buf, s)
                d, l, kv = [int(x) for x in meta.tolist()]
                off = d * BLOCK_BYTES + (l * 2 + kv) * CHUNK
                if not torch.equal(buf, cache[off:off + CHUNK]):
                    ok.zero_()
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if int(ok.item()) != 1:
        raise RuntimeError("peer push result differs from its source")

    # ---- comparators: (i) local relayout into HBM staging + NCCL send/recv, (ii) plain peer memcpy of the payload
    nsteps_c = max(2, min(steps, 10))
    staging = torch.empty(pairs * BLOCK_BYTES, dtype=torch.uint8, device="cuda")
    nccl_plan = C.c_void_p()
    if is_prefill:
        lib.check(L.kvdp_plan_contig(C.byref(nccl_plan), setup.src_ptrs.ctypes.data, L8B, 0, n, CHUNK,
                                     C.c_void_p(staging.data_ptr()), pairs, 1, local_rank))
    iota = np.arange(pairs, dtype=np.int32)

    def nccl_steps():
        cur = torch.cuda.current_stream()
        for k in range(nsteps_c):
            if is_prefill:
                s_ids, _ = setup.ids(k)
                lib.check(L.kvdp_xfer_run(setup.xfer, nccl_plan, s_ids.ctypes.data, iota.ctypes.data, pairs, 0, L8B, 0,
                                          None, 0))
                ev = torch.cuda.Event()
                ev.record(setup.stream)
                cur.wait_event(ev)
                dist.send(staging, peer)
                ev2 = torch.cuda.Event()
                ev2.record(cur)
                setup.stream.wait_event(ev2)
            else:
                for s in senders:
                    dist.recv(staging, s)

    def timed_default_stream(fn):
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        dist.barrier()
        t = torch.tensor([a.elapsed_time(b) / 1e3], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    nccl_value = memcpy_value = None
    if not args.no_comparators:
        nccl_steps()  # warm-up (NCCL connection setup)
        t_nccl = timed_default_stream(nccl_steps)
        nccl_value = P * pairs * BLOCK_BYTES * nsteps_c / t_nccl / GB

    def memcpy_steps():
        if is_prefill:
            for _ in range(nsteps_c):
                lib.check(L.cuda_memcpy_async(peer_dst, C.c_void_p(staging.data_ptr()), pairs * BLOCK_BYTES,
                                              C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    if not args.no_comparators:
        memcpy_steps()
        t_cpy = timed_default_stream(memcpy_steps)
        memcpy_value = P * pairs * BLOCK_BYTES * nsteps_c / t_cpy / GB

    if rank == 0:
        peak, peak_src = measured_peaks()
        cfg = workload_config(args)
        cfg["pairs"] = ["%d->%d" % (r, P + r % D) for r in range(P)]
        line = {
            "metric": "kv_transfer_GBps", "value": value, "unit": "GB/s", "n_gpus": world, "steps": steps,
            "warmup": warmup, "ms_per_step": sec / steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": cfg,
            "blocks_per_s": P * pairs * steps / sec,
            "roofline": {"bound": "nvlink", "achieved": per_link, "peak": NVLINK_PEAK_MEASURED, "unit": "GB/s",
                         "frac": per_link / NVLINK_PEAK_MEASURED, "traffic": None,
                         "kernel": "kvdp_copy_tma_in_kernel / kvdp_copy_vec_kernel (peer stores)",
                         "peak_source": "measured peer copy per direction per GPU (B200_PROFILING.md); nominal 900",
                         "frac_of_nominal_900": per_link / NVLINK_PEAK_NOMINAL,
                         "note": "achieved = wire bytes per direction on the busiest port / time"},
            "comparators": {"nccl_send_recv_GBps": nccl_value, "cuda_memcpy_peer_GBps": memcpy_value,
                            "steps": nsteps_c},
            "options": {k: lib.get_option(k) for k in ("copy_path", "tile_bytes", "stages", "ctas_per_sm")},
            "gpu_launches": steps, "clocks": clk,
            "e2e": {"value": None, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                    "note": "see the N=1 line: the host-buffer path is PCIe-bound per GPU and does not cross NVLink"},
        }
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()
