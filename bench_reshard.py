#!/usr/bin/env python
"""bench_reshard.py — BASELINE config 3: Llama-3-70B-shaped KV cache, TP-P prefill -> TP-D decode head resharding
across P + D GPUs (default TP4 -> TP2 on 6 B200), through the transfer connector (dynamo_exp_b200/connector.py).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 6 --master-addr 127.0.0.1 --master-port 29520 \
        bench_reshard.py --prefill-tp 4 --decode-tp 2 --steps 20 --warmup 3

Ranks [0, P) are the prefill engine's TP ranks (caches [2, N, 64, 8/P, 128] x 80 layers, bf16), ranks [P, P+D) the
decode engine's. Two modes are timed:
  push : every prefill rank write_blocks() its head slice into the decode caches (kernel stores over NVLink); decode
         ranks stream-wait on their mailbox heads;
  pull : every decode rank read_blocks() from the prefill caches (kernel loads over NVLink).
Roofline: NVLink per direction on the busiest port — decode ingress = (P/D) x per-prefill-rank egress.
Timing: barrier + device sync on both sides, CUDA events on the connector's stream, MAX over ranks. One JSON line.
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


class CudaArray:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


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

    from dynamo_exp_b200 import _lib as lib, build
    from dynamo_exp_b200.connector import DynamoKvdpConnector, tp_targets
    build.build()
    world = int(os.environ["WORLD_SIZE"])
    rank = int(os.environ["RANK"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    P, D = args.prefill_tp, args.decode_tp
    assert world == P + D, "launch with --nproc-per-node = prefill_tp + decode_tp"
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    L = lib.lib()
    is_prefill = rank < P
    tp, tp_rank = (P, rank) if is_prefill else (D, rank - P)
    heads = H70 // tp
    n = args.cache_blocks
    pairs = 64 * args.requests
    row = heads * D70 * ELEM
    layer_bytes = 2 * n * BS * row

    # caches (cudaMalloc through libkvdp so that they are IPC-exportable), prefill side filled with random bytes
    layers = []
    g = torch.Generator(device="cuda")
    g.manual_seed(7 + rank)
    for _ in range(args.layers):
        p = C.c_void_p()
        lib.check(L.kvdp_device_alloc(C.byref(p), layer_bytes, local_rank))
        if is_prefill:
            torch.as_tensor(CudaArray(p.value, layer_bytes), device="cuda").random_(0, 256, generator=g)
        layers.append(p.value)
    conn = DynamoKvdpConnector("prefill" if is_prefill else "decode", tp_rank, tp, device=local_rank,
                               max_blocks_per_transfer=pairs)
    conn.register_kv_caches(layers, shape=(2, n, BS, heads, D70), element_size=ELEM)
    blobs = [None] * world
    dist.all_gather_object(blobs, conn.get_agent_metadata())
    pre_meta, dec_meta = blobs[:P], blobs[P:]
    if is_prefill:
        conn.add_remote_agent("decode", dec_meta, D, None, n)
    else:
        conn.add_remote_agent("prefill", pre_meta, P, None, n)
    dist.barrier()

    rng = np.random.Generator(np.random.PCG64(99))      # same ids on every rank
    perm_s, perm_d = rng.permutation(n).astype(np.int32), rng.permutation(n).astype(np.int32)

    def ids(step):
        w = max(1, n // pairs)
        o = (step % w) * pairs
        return perm_s[o:o + pairs], perm_d[o:o + pairs]

    stream = torch.cuda.ExternalStream(conn._engine.stream)
    wait_stream = torch.cuda.Stream()
    sent = {"n": 0}

    def timed(mode, k0, count):
        runner = is_prefill if mode == "push" else not is_prefill
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st = stream if runner else wait_stream
        a.record(st)
        if mode == "push":
            sent["n"] += count
            if is_prefill:
                for k in range(count):
                    s, d = ids(k0 + k)
                    conn.write_blocks(s, s, d, "decode", "r%d" % (k0 + k))
            else:
                senders = [p for p in range(P) if tp_rank in tp_targets(P, D, p)]
                for p in senders:   # one mailbox lane per sending prefill rank; its head counts delivered messages
                    lib.check(L.kvdp_stream_wait_flag(C.c_void_p(wait_stream.cuda_stream),
                                                      C.c_void_p(conn.notif_head_ptr("prefill:%d" % p)), sent["n"]))
        else:
            if not is_prefill:
                for k in range(count):
                    s, d = ids(k0 + k)
                    conn.read_blocks(d, d, s, "prefill")     # pull prefill block s into decode block d
        b.record(st)
        torch.cuda.synchronize()
        dist.barrier()
        t = torch.tensor([a.elapsed_time(b) / 1e3], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def verify(step):
        """Decode rank q, layer l: its heads of block d equal the matching head slice of the prefill ranks' block s."""
        s_ids, d_ids = ids(step)
        ok = torch.ones(1, dtype=torch.int32, device="cuda")
        hp, hd = H70 // P, H70 // D
        for q in range(D):
            for p in [p for p in range(P) if q in tp_targets(P, D, p)]:
                l, i, kv = 3 % args.layers, 5, 1
                if rank == p:
                    t = torch.as_tensor(CudaArray(layers[l], layer_bytes), device="cuda").view(2, n, BS, hp * D70 * ELEM)
                    dist.send(t[kv, int(s_ids[i])].contiguous(), P + q)
                if rank == P + q:
                    buf = torch.empty(BS, hp * D70 * ELEM, dtype=torch.uint8, device="cuda")
                    dist.recv(buf, p)
                    t = torch.as_tensor(CudaArray(layers[l], layer_bytes), device="cuda").view(2, n, BS, hd * D70 * ELEM)
                    if P >= D:
                        j = p % (P // D)
                        mine = t[kv, int(d_ids[i])][:, j * hp * D70 * ELEM:(j + 1) * hp * D70 * ELEM]
                        want = buf
                    else:
                        m = D // P
                        gq = q % m
                        mine = t[kv, int(d_ids[i])]
                        want = buf[:, gq * hd * D70 * ELEM:(gq + 1) * hd * D70 * ELEM]
                    if not torch.equal(mine, want):
                        ok.zero_()
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) != 1:
            raise RuntimeError("reshard result differs from its source")

    results = {}
    for mode in ("push", "pull"):
        if not is_prefill:
            for p in layers:
                lib.check(L.kvdp_memset(C.c_void_p(p), 0, layer_bytes))
        timed(mode, 0, args.warmup)
        sec = timed(mode, args.warmup, args.steps)
        verify(args.warmup + args.steps - 1)
        per_prefill = pairs * args.layers * 2 * BS * (H70 // P) * D70 * ELEM
        total = per_prefill * P * args.steps
        ingress = total / D / sec / GB
        egress = total / P / sec / GB
        results[mode] = {"aggregate_GBps": total / sec / GB, "busiest_direction_GBps": max(ingress, egress),
                         "decode_ingress_GBps": ingress, "prefill_egress_GBps": egress,
                         "frac_of_nominal_900": max(ingress, egress) / 900.0,
                         "frac_of_measured_770": max(ingress, egress) / 770.0,
                         "blocks_per_s": pairs * args.steps / sec, "ms_per_step": sec / args.steps * 1e3}
    if rank == 0:
        print(json.dumps({
            "metric": "kv_transfer_GBps", "unit": "GB/s", "n_gpus": world,
            "config": {"workload": "llama3-70b-shape KV (%d layers, 8 KV heads x 128, 64-token blocks, bf16), "
                                   "TP%d prefill -> TP%d decode head reshard, %d blocks/step/rank" %
                                   (args.layers, P, D, pairs),
                       "run_bytes": min(H70 // P, H70 // D) * D70 * ELEM, "cache_blocks": n},
            "steps": args.steps, "warmup": args.warmup, "modes": results}), flush=True)
    conn.shutdown()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
