#!/usr/bin/env python
"""bench.py — KV-cache data-plane benchmark (contract: see the task statement / DESIGN.md §Measurement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path over one batch of synthetic KV blocks:
  N = 1 : BASELINE config 2 on one GPU — Llama-3-8B / DeepSeek-R1-Distill-Llama-8B shape (32 layers, 8 KV heads x 128,
          64-token blocks, bf16), a batch of R requests of ISL 4k (64 blocks each): layer-separate -> contiguous-per-block
          relayout, ONE kernel launch per step (HBM-bound).
  N > 1 : the same batch per prefill GPU, pushed from inside the relayout kernel into the contiguous cache of a decode
          GPU through NVLink peer-mapped stores (prefill ranks = first half, decode ranks = second half).
`value` is payload GB/s delivered (1 GB = 1e9 B) over the whole job, inputs resident in HBM. `e2e` is the same unit
measured through the reference-facing legacy C ABI (CopyStream: prepare_block_ids / memcpy per layer / sync) with the
source blocks in pinned HOST memory, i.e. including the host->device copy, plus a device->host read of the result.
Token-block hashing (hashes/s) is timed in the same run and reported under "hash".
`--impl reference` times the CPU restatement of the reference's host block-copy loops (oracle/) on the host cores.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GB = 1e9
# Llama-3-8B shape (BASELINE.json configs[1]); 64-token blocks, bf16
L8B, H8B, D8B, BS, ELEM = 32, 8, 128, 64, 2
CHUNK = BS * H8B * D8B * ELEM          # 128 KiB per (layer, K|V, block)
BLOCK_BYTES = L8B * 2 * CHUNK          # 8 MiB per block over all layers
BLOCKS_PER_REQ = 64                    # ISL 4k


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--requests", type=int, default=8, help="requests (of 64 blocks) per step")
    ap.add_argument("--cache-blocks", type=int, default=1024, help="blocks in the source / destination caches")
    ap.add_argument("--pairing", choices=["split", "interleaved"], default="interleaved",
                    help="N>1: which GPUs prefill: the even ones, each pushing to its odd neighbour (default: GPUs that\n"
                         "share a PCIe switch are then not both reading host memory in the e2e leg), or the first half")
    ap.add_argument("--cpu-affinity", type=int, default=0, help="N>1: 1 = bind each rank to the CPUs NVML names for its GPU")
    ap.add_argument("--copy-path", type=int, default=0)
    ap.add_argument("--tile-bytes", type=int, default=0)
    ap.add_argument("--stages", type=int, default=0)
    ap.add_argument("--ctas-per-sm", type=int, default=0)
    ap.add_argument("--host-path", type=int, default=0)
    ap.add_argument("--host-ctas", type=int, default=0)
    ap.add_argument("--l2-hint", type=int, default=0)
    ap.add_argument("--vec32", type=int, default=0, help="copy path 1: 256-bit loads/stores")
    ap.add_argument("--no-hash", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--mode", default="both", choices=["both", "push", "pull"],
                    help="N>1: push = prefill-side kernel stores into the decode GPU; pull = decode-side kernel loads "
                         "from the prefill GPU; both = measure both, headline the faster")
    ap.add_argument("--no-comparators", action="store_true", help="N>1: skip the NCCL / memcpy-peer comparison legs")
    ap.add_argument("--e2e-path", default="dma", choices=["dma", "kernel"],
                    help="N>1 e2e leg, prefill side: dma = legacy CopyStream (copy engines into HBM staging + push kernel), "
                         "kernel = one zero-copy kernel reading host memory")
    ap.add_argument("--e2e-links", default="auto", choices=["auto", "both", "prefill"],
                    help="N>1 e2e leg: auto = time both settings, headline the faster; both = the prefill side's host blocks live in memory both workers of a pair map and "
                         "pin, the decode GPU fetches half of each step over its own PCIe link; prefill = the prefill GPU's "
                         "link carries everything")
    ap.add_argument("--no-reshard", action="store_true", help="N>1: skip the Llama-3-70B TP reshard legs (config 3)")
    ap.add_argument("--sweep", action="store_true", help="print a table of kernel variants instead of the JSON line")
    return ap.parse_args()


def env_info(torch=None, device_index: int = 0) -> dict:
    """Where the numbers came from: GPU, driver / CUDA / library versions, host cores and the synthetic-data seeds
    (SURVEY D.3)."""
    info = {"host_cores": os.cpu_count(), "seeds": {"tokens": "0xD15A66", "kv_payload": 42, "block_ids": 1234}}
    try:
        from dynamo_exp_b200 import _lib
        info["libkvdp_version"] = int(_lib.lib().kvdp_version())
    except Exception:
        pass
    if torch is not None:
        try:
            p = torch.cuda.get_device_properties(device_index)
            info.update({"gpu": p.name, "sms": p.multi_processor_count, "hbm_GiB": round(p.total_memory / 2 ** 30, 1),
                         "torch": torch.__version__, "cuda_runtime": torch.version.cuda})
        except Exception:
            pass
    try:
        import pynvml
        pynvml.nvmlInit()
        v = pynvml.nvmlSystemGetDriverVersion()
        info["driver"] = v.decode() if isinstance(v, bytes) else v
    except Exception:
        pass
    return info


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock, power and throttle reasons through NVML (in-process, every 5 ms) while the timed region runs;
    falls back to an `nvidia-smi -lms` subprocess when NVML cannot be loaded."""

    def __init__(self, device_index: int):
        self.idx = device_index
        self.samples = []          # (sm_mhz, power_w, reasons_bitmask)
        self.max_mhz = None
        self._stop = threading.Event()
        self.thread = None
        self.nvml = None
        self.handle = None
        self.proc = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            idx = self.idx
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            if vis:
                try:
                    idx = int(vis.split(",")[self.idx])
                except Exception:
                    pass
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None
            return self._start_smi()
        self.thread = threading.Thread(target=self._loop, daemon=True)
        self.thread.start()

    def _loop(self):
        n = self.nvml
        while not self._stop.is_set():
            try:
                mhz = float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
                try:
                    pw = n.nvmlDeviceGetPowerUsage(self.handle) / 1000.0
                except Exception:
                    pw = None
                try:
                    rs = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
                except Exception:
                    rs = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
                self.samples.append((mhz, pw, rs))
            except Exception:
                pass
            time.sleep(0.005)

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def _start_smi(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-i", str(self.idx), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.nvml is not None:
            self._stop.set()
            self.thread.join(timeout=1)
            n = self.nvml
            names = (("hw_slowdown", n.nvmlClocksEventReasonHwSlowdown),
                     ("hw_thermal_slowdown", n.nvmlClocksEventReasonHwThermalSlowdown),
                     ("sw_thermal_slowdown", n.nvmlClocksEventReasonSwThermalSlowdown),
                     ("sw_power_cap", n.nvmlClocksEventReasonSwPowerCap),
                     ("hw_power_brake", n.nvmlClocksEventReasonHwPowerBrakeSlowdown))
            reasons = sorted({name for _, _, rs in self.samples for name, bit in names if rs & bit})
            sm = [s[0] for s in self.samples]
            pw = [s[1] for s in self.samples if s[1] is not None]
            return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_mhz,
                    "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": reasons,
                    "source": "nvml, 5 ms period, during the timed region"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock sampling unavailable"], "samples": 0}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=3)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons, pw = [], [], set(), []
        for ln in out.splitlines():
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                pw.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons),
                "source": "nvidia-smi -lms 100"}


# ------------------------------------------------------------------------------------------------ CPU legs (oracle)
class CpuRelayout:
    """The reference's host block-copy loop restated in C (oracle/layout_oracle.c): layer-separate -> contiguous-per-block
    relayout of Llama-3-8B blocks by memcpy on the host cores. `cache_blocks` blocks live in each cache; a step moves
    `blocks` of them, a different window every step, exactly like the GPU arm's step."""

    def __init__(self, blocks: int, cache_blocks: int):
        from oracle import kvoracle as ko
        ko.lib()
        self.ko, self.blocks, self.n = ko, blocks, max(cache_blocks, blocks)
        self.layers = [np.empty(2 * self.n * CHUNK, dtype=np.uint8) for _ in range(L8B)]
        for i, a in enumerate(self.layers):
            a[:] = (i * 7 + 1) & 0xFF  # touch every page
        self.contig = np.zeros(self.n * BLOCK_BYTES, dtype=np.uint8)
        rng = np.random.Generator(np.random.PCG64(1234))
        self.sperm, self.dperm = rng.permutation(self.n).astype(np.int32), rng.permutation(self.n).astype(np.int32)

    def step(self, k: int, threads: int) -> float:
        w = max(1, self.n // self.blocks)
        o = (k % w) * self.blocks
        t0 = time.perf_counter()
        self.ko.relayout_contig(self.layers, 0, self.n, CHUNK, self.contig, self.sperm[o:o + self.blocks],
                                self.dperm[o:o + self.blocks], True, threads)
        return time.perf_counter() - t0


def cpu_relayout_leg(threads: int, min_seconds: float, blocks: int = BLOCKS_PER_REQ, cache_blocks: int = 0, leg=None):
    """GB/s of payload and seconds per step of the CPU relayout, repeated for at least `min_seconds`."""
    leg = leg or CpuRelayout(blocks, cache_blocks or blocks)
    leg.step(0, threads)  # warm-up
    reps, spent = 0, 0.0
    while True:
        spent += leg.step(reps + 1, threads)
        reps += 1
        if spent >= min_seconds:
            break
    return leg.blocks * BLOCK_BYTES * reps / spent / GB, spent / reps


def cpu_hash_leg(threads: int, min_seconds: float):
    from oracle import kvoracle as ko
    from tests.util import tokens_for
    isls = [512, 1024, 2048, 4096, 8192, 16384, 32768]
    reqs = [isls[i % 7] for i in range(7 * 16)]
    offs = np.zeros(len(reqs) + 1, dtype=np.int64)
    offs[1:] = np.cumsum(reqs)
    toks = np.concatenate([tokens_for(r, n) for r, n in enumerate(reqs)])
    nblk = int(offs[-1]) // BS
    ko.hash_batch(toks, offs, BS, 1337, threads)
    reps, t0 = 0, time.perf_counter()
    while True:
        ko.hash_batch(toks, offs, BS, 1337, threads)
        reps += 1
        dt = time.perf_counter() - t0
        if dt >= min_seconds:
            break
    return nblk * reps / dt


def run_reference(args):
    """--impl reference: the CPU implementation of the path on the host cores, same metric / unit / config / step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import kvoracle as ko
    threads = ko.hw_threads()
    blocks = BLOCKS_PER_REQ * args.requests
    # same step as the GPU arm: `blocks` blocks out of caches of `cache_blocks`; the cache is capped at 2 steps' worth
    # so that the host allocation stays at 2 x 8 GiB
    leg = CpuRelayout(blocks, min(args.cache_blocks, 2 * blocks))
    for k in range(args.warmup):
        leg.step(k, threads)
    secs = []
    t0 = time.perf_counter()
    for k in range(args.steps):
        secs.append(leg.step(args.warmup + k, threads))
        if time.perf_counter() - t0 > 150:
            break
    step_s = float(np.mean(secs))
    value = blocks * BLOCK_BYTES / step_s / GB
    del leg
    hps = cpu_hash_leg(threads, 2.0)
    line = {
        "impl": "reference", "metric": "kv_transfer_GBps", "value": value, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": len(secs), "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": "GB/s", "cores": threads, "kind": "port",
                         "sample": "%d Llama-3-8B blocks (%d MiB payload) per step, layer-separate -> contiguous relayout "
                                   "by memcpy (oracle/layout_oracle.c), all host threads; host-resident caches of %d "
                                   "blocks" % (blocks, blocks * BLOCK_BYTES >> 20, min(args.cache_blocks, 2 * blocks))},
        "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "blocks_per_s": value * GB / BLOCK_BYTES,
        "hash": {"hashes_per_s": hps, "cores": threads, "kind": "port"},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args):
    return {
        "workload": "llama3-8b-shape KV blocks (32 layers, 8 KV heads x 128, 64-token blocks, bf16 = 8 MiB/block), "
                    "ISL 4k requests (64 blocks), layer-separate -> contiguous-per-block"
                    + (" + NVLink peer push prefill->decode" if args.gpus > 1 else " relayout on one GPU"),
        "requests_per_step": args.requests,
        "blocks_per_step": BLOCKS_PER_REQ * args.requests,
        "cache_blocks": args.cache_blocks,
        "l2_policy": "inputs larger than L2: every step reads/writes a different %d MiB window of 8 GiB caches"
                     % (BLOCKS_PER_REQ * args.requests * BLOCK_BYTES >> 20),
        "parallelism": "1 gpu" if args.gpus == 1 else "%dP:%dD" % (args.gpus // 2, args.gpus - args.gpus // 2),
    }


# ------------------------------------------------------------------------------------------------ our arm
def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist

    from dynamo_exp_b200 import _lib, build
    build.build()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node %d for --gpus %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local_rank)
    _lib.require_gpu()
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    for name, val in (("copy_path", args.copy_path), ("tile_bytes", args.tile_bytes), ("stages", args.stages),
                      ("ctas_per_sm", args.ctas_per_sm), ("host_path", args.host_path), ("host_ctas", args.host_ctas),
                      ("l2_hint", args.l2_hint), ("vec32", args.vec32)):
        _lib.set_option(name, val)

    if world > 1:
        from bench_multi import run_multi
        return run_multi(args, rank, local_rank, world)

    from bench_single import run_single, run_sweep
    if args.sweep:
        return run_sweep(args)
    return run_single(args)


if __name__ == "__main__":
    main()
