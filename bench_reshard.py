#!/usr/bin/env python
"""bench_reshard.py — BASELINE config 3: Llama-3-70B-shaped KV cache, TP-P prefill -> TP-D decode head resharding
across P + D GPUs, through the transfer connector (dynamo_exp_b200/connector.py).

`reshard_leg()` is what `bench.py --gpus N` runs after its 1P->1D legs (N=8: TP4->TP2 on GPUs 0-5 — the north-star case,
two GPUs idle — and TP4->TP4 on all eight; N=4: TP2->TP2 and TP2->TP1 + a spare; N=2: TP1->TP1). Standalone:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 6 --master-addr 127.0.0.1 --master-port 29520 \
        bench_reshard.py --prefill-tp 4 --decode-tp 2 --steps 20 --warmup 3

Ranks [0, P) are the prefill engine's TP ranks (caches [2, N, 64, 8/P, 128] x 80 layers, bf16), ranks [P, P+D) the
decode engine's, the rest idle (they only join the barriers). Two modes are timed:
  push : every prefill rank write_blocks() its head slice into the decode caches (kernel stores over NVLink); decode
         ranks stream-wait on the notification counters the senders bump behind their data;
  pull : every decode rank read_blocks() from the prefill caches (kernel loads over NVLink).
Roofline: NVLink per direction on the busiest port — decode ingress = (P/D) x per-prefill-rank egress.
Timing: barrier + device sync on both sides, CUDA events on the connector's stream, MAX over ranks.

Verification (every byte, not samples): layer l of a canonical TP1 cache is a closed-form function of (l, element index),
so every GPU can regenerate it; prefill rank p holds its head slice, and after each mode every decode rank compares ALL
layers of ALL blocks of the last step with the head slice it should now hold. One extra un-timed push step checks the
ordering of the completion signal across GPUs: the decode side enqueues, on one stream, a wait on the senders' counters
followed immediately by the comparison kernels, with no other synchronisation between the two GPUs.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GB = 1e9
L70, H70, D70, BS, ELEM = 80, 8, 128, 64, 2
NVLINK_NOMINAL = 900.0


class CudaArray:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def canonical_layer(torch, layer: int, n_blocks: int):
    """Layer `layer` of the canonical TP1 cache as int16 [2, n_blocks, 64, 8, 128]: integer mixing of the element index,
    identical on every GPU (no RNG state involved)."""
    n = 2 * n_blocks * BS * H70 * D70
    x = torch.arange(n, dtype=torch.int64, device="cuda")
    x = x * 0x9E3779B97F4A7C1 + (((layer + 1) * 0x2545F4914F6CDD1D) & 0x7FFFFFFFFFFFFFFF)   # wraps mod 2^64
    x = x ^ (x >> 29)
    x = x * (0x94D049BB133111EB - (1 << 64))
    x = x ^ (x >> 32)
    return x.to(torch.int16).view(2, n_blocks, BS, H70, D70)


def reshard_leg(torch, dist, lib, rank, local_rank, world, P, D, steps, warmup, requests=4, cache_blocks=512,
                layers=L70, clock_sampler=None, flag_layers=8):
    """One TP-P -> TP-D configuration on ranks [0, P+D) of `world`. Returns the result dict on rank 0, None elsewhere."""
    from dynamo_exp_b200.connector import DynamoKvdpConnector, tp_targets
    L = lib.lib()
    assert world >= P + D
    active = rank < P + D
    is_prefill = rank < P
    is_decode = P <= rank < P + D
    tp, tp_rank = (P, rank) if is_prefill else (D, rank - P)
    heads = H70 // tp
    n, pairs = cache_blocks, 64 * requests
    row = heads * D70 * ELEM
    layer_bytes = 2 * n * BS * row
    hp, hd = H70 // P, H70 // D

    bufs, conn = [], None
    if active:
        for l in range(layers):
            p = C.c_void_p()
            lib.check(L.kvdp_device_alloc(C.byref(p), layer_bytes, local_rank))
            if is_prefill:
                canon = canonical_layer(torch, l, n)
                mine = torch.as_tensor(CudaArray(p.value, layer_bytes), device="cuda").view(torch.int16)
                mine.view(2, n, BS, heads, D70).copy_(canon[:, :, :, tp_rank * hp:(tp_rank + 1) * hp, :])
                del canon
            bufs.append(p.value)
        conn = DynamoKvdpConnector("prefill" if is_prefill else "decode", tp_rank, tp, device=local_rank,
                                   max_blocks_per_transfer=pairs)
        conn.register_kv_caches(bufs, shape=(2, n, BS, heads, D70), element_size=ELEM)
    torch.cuda.synchronize()
    blobs = [None] * world
    dist.all_gather_object(blobs, conn.get_agent_metadata() if active else None)
    pre_meta, dec_meta = blobs[:P], blobs[P:P + D]
    if is_prefill:
        conn.add_remote_agent("decode", dec_meta, D, None, n)
    elif is_decode:
        conn.add_remote_agent("prefill", pre_meta, P, None, n)
    dist.barrier()

    rng = np.random.Generator(np.random.PCG64(99))      # same ids on every rank
    perm_s, perm_d = rng.permutation(n).astype(np.int32), rng.permutation(n).astype(np.int32)

    def ids(step):
        w = max(1, n // pairs)
        o = (step % w) * pairs
        return perm_s[o:o + pairs], perm_d[o:o + pairs]

    stream = torch.cuda.ExternalStream(conn._engine.stream) if active else None
    wait_stream = torch.cuda.Stream()
    sent = {"n": 0}
    senders = [p for p in range(P) if is_decode and tp_rank in tp_targets(P, D, p)]
    head_ptrs = []

    def wait_for_senders(count_total):
        if not head_ptrs:
            for p in senders:
                ptr = conn.notif_head_ptr("prefill:%d" % p)
                assert ptr, "sender prefill:%d never claimed a lane here" % p
                head_ptrs.append(ptr)
        for ptr in head_ptrs:
            lib.check(L.kvdp_stream_wait_flag(C.c_void_p(wait_stream.cuda_stream), C.c_void_p(ptr), count_total))

    def timed(mode, k0, count):
        runner = is_prefill if mode == "push" else is_decode
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st = stream if (runner and active) else wait_stream
        a.record(st)
        if mode == "push":
            sent["n"] += count
            if is_prefill:
                for k in range(count):
                    s, d = ids(k0 + k)
                    conn.write_blocks(s, s, d, "decode", "r%d" % (k0 + k))
            elif is_decode:
                wait_for_senders(sent["n"])
        elif is_decode:
            for k in range(count):
                s, d = ids(k0 + k)
                conn.read_blocks(d, d, s, "prefill")     # pull prefill block s into decode block d
        b.record(st)
        torch.cuda.synchronize()
        if is_decode and mode == "push":
            conn.get_new_notifs()                          # drain the rings like a polling worker
        dist.barrier()
        t = torch.tensor([a.elapsed_time(b) / 1e3], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def my_layer(l):
        return torch.as_tensor(CudaArray(bufs[l], layer_bytes), device="cuda").view(torch.int16).view(2, n, BS, heads, D70)

    def verify_all(step, which):
        """Every byte of every layer of the blocks moved in `step`, on every decode rank."""
        bad = torch.zeros(1, dtype=torch.int64, device="cuda")
        checked = 0
        if is_decode:
            s_ids, d_ids = ids(step)
            s_t = torch.from_numpy(s_ids.astype(np.int64)).cuda()
            d_t = torch.from_numpy(d_ids.astype(np.int64)).cuda()
            for l in range(layers):
                want = canonical_layer(torch, l, n)[:, s_t][:, :, :, tp_rank * hd:(tp_rank + 1) * hd, :]
                got = my_layer(l)[:, d_t]
                bad += (want != got).sum()
                checked += want.numel() * ELEM
        tot = torch.tensor([checked], dtype=torch.int64, device="cuda")
        dist.all_reduce(bad, op=dist.ReduceOp.SUM)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        if int(bad.item()) != 0:
            raise RuntimeError("reshard %s TP%d->TP%d: %d elements differ from the canonical head slice" %
                               (which, P, D, int(bad.item())))
        return int(tot.item())

    def zero_decode():
        if is_decode:
            for p in bufs:
                lib.check(L.kvdp_memset(C.c_void_p(p), 0, layer_bytes))

    k_steps, k_warm = max(2, min(steps, 10)), max(1, min(warmup, 2))
    results, clocks = {}, None
    for mode in ("push", "pull"):
        zero_decode()
        timed(mode, 0, k_warm)
        if clock_sampler is not None and rank == 0:
            clock_sampler.start()
        sec = timed(mode, k_warm, k_steps)
        if clock_sampler is not None and rank == 0:
            c = clock_sampler.stop()
            clocks = clocks or c
            clock_sampler = type(clock_sampler)(clock_sampler.idx)
        verified = verify_all(k_warm + k_steps - 1, mode)
        per_prefill = pairs * layers * 2 * BS * hp * D70 * ELEM
        total = per_prefill * P * k_steps
        ingress, egress = total / D / sec / GB, total / P / sec / GB
        results[mode] = {"aggregate_GBps": total / sec / GB, "busiest_direction_GBps": max(ingress, egress),
                         "decode_ingress_GBps": ingress, "prefill_egress_GBps": egress,
                         "frac_of_nominal_900": max(ingress, egress) / NVLINK_NOMINAL,
                         "blocks_per_s": pairs * k_steps / sec, "ms_per_step": sec / k_steps * 1e3,
                         "verified_bytes_last_step": verified}

    # ---- flag-ordering step: wait on the senders' counters, then read the data at once on the same stream
    zero_decode()
    torch.cuda.synchronize()
    step = 10_000
    pick = sorted({int(x) for x in np.linspace(0, layers - 1, min(flag_layers, layers))})
    want, bad = [], torch.zeros(1, dtype=torch.int64, device="cuda")
    if is_decode:
        s_ids, d_ids = ids(step)
        s_t = torch.from_numpy(s_ids.astype(np.int64)).cuda()
        d_t = torch.from_numpy(d_ids.astype(np.int64)).cuda()
        want = [canonical_layer(torch, l, n)[:, s_t][:, :, :, tp_rank * hd:(tp_rank + 1) * hd, :].contiguous()
                for l in pick]
    torch.cuda.synchronize()
    dist.barrier()
    sent["n"] += 1
    if is_prefill:
        s, d = ids(step)
        conn.write_blocks(s, s, d, "decode", "flag-order")
    elif is_decode:
        wait_for_senders(sent["n"])
        with torch.cuda.stream(wait_stream):
            for l, w in zip(pick, want):
                bad += (w != my_layer(l)[:, d_t]).sum()
    torch.cuda.synchronize()
    dist.all_reduce(bad, op=dist.ReduceOp.SUM)
    if int(bad.item()) != 0:
        raise RuntimeError("TP%d->TP%d: data read right after the completion signal was incomplete (%d elements)" %
                           (P, D, int(bad.item())))
    if is_decode:
        conn.get_new_notifs()
    del want

    if active:
        conn.sync()
    dist.barrier()
    if active:
        conn.shutdown()                                    # closes this rank's mappings of its peers' caches
    dist.barrier()                                         # ... before any owner frees them
    for p in bufs:
        L.kvdp_device_free(C.c_void_p(p))
    dist.barrier()
    if rank != 0:
        return None
    return {"config": "TP%d->TP%d" % (P, D), "gpus_used": P + D, "gpus_idle": world - P - D,
            "workload": "llama3-70b-shape KV (%d layers, 8 KV heads x 128, 64-token blocks, bf16), TP%d prefill -> TP%d "
                        "decode head reshard, %d blocks/step per prefill rank" % (layers, P, D, pairs),
            "run_bytes": min(hp, hd) * D70 * ELEM, "cache_blocks": n, "steps": k_steps, "warmup": k_warm,
            "modes": results, "clocks": clocks,
            "verification": "all bytes: every layer of every block of the last step of each mode, on every decode rank, "
                            "against the canonical TP1 head slice regenerated on that GPU",
            "flag_order_check": "passed: %d layers read on the decode stream directly behind kvdp_stream_wait_flag on "
                                "the senders' counters, no other cross-GPU synchronisation" % len(pick),
            "api": "DynamoKvdpConnector.write_blocks / read_blocks (kvdp_plan_reshard + kvdp_xfer_run over CUDA-IPC "
                   "peer mappings)"}


def configs_for(world: int):
    """(prefill TP, decode TP) configurations bench.py runs at a given GPU count."""
    return {2: [(1, 1)], 3: [(2, 1)], 4: [(2, 2), (2, 1)], 6: [(4, 2)], 8: [(4, 2), (4, 4)]}.get(world, [])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--prefill-tp", type=int, default=4)
    ap.add_argument("--decode-tp", type=int, default=2)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--requests", type=int, default=4, help="ISL-4k requests (64 blocks) per step")
    ap.add_argument("--cache-blocks", type=int, default=512)
    ap.add_argument("--layers", type=int, default=L70)
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    from bench import ClockSampler
    from dynamo_exp_b200 import _lib as lib, build
    build.build()
    world = int(os.environ["WORLD_SIZE"])
    rank = int(os.environ["RANK"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    res = reshard_leg(torch, dist, lib, rank, local_rank, world, args.prefill_tp, args.decode_tp, args.steps,
                      args.warmup, args.requests, args.cache_blocks, args.layers, ClockSampler(local_rank))
    if rank == 0:
        print(json.dumps({"metric": "kv_transfer_GBps", "unit": "GB/s", "n_gpus": world, **res}), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
