"""bench_single.py — the N=1 arm of bench.py (one B200): relayout kernel, e2e through the legacy C ABI, hashing."""
from __future__ import annotations

import ctypes as C
import json
import os
import time

import numpy as np

from bench import (BLOCK_BYTES, BLOCKS_PER_REQ, BS, CHUNK, ELEM, GB, H8B, D8B, L8B, ROOT, ClockSampler, env_info,
                   cpu_hash_leg, cpu_relayout_leg, workload_config)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kernel_key: str):
    """Per-launch DRAM bytes of a kernel from the committed ncu capture — only while the capture still describes the
    source this run was built from: profiles/ncu_traffic.json stores, next to each number, a hash of the kernel's source
    files (tools/ncu_summary.py writes both); a mismatch means the kernel changed after the capture and yields None."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        import importlib.util
        spec = importlib.util.spec_from_file_location("kvdp_ncu_summary", os.path.join(ROOT, "tools", "ncu_summary.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        entry = json.load(open(p)).get(kernel_key)
        if not isinstance(entry, dict) or entry.get("source_hash") != mod.source_hash(kernel_key):
            return None
        return entry["dram_bytes_per_launch"]
    except Exception:
        return None


def ncu_traffic_note(kernel_key: str) -> str:
    t = ncu_traffic(kernel_key)
    if t is None:
        return "no ncu capture matches the current source of %s (profiles/ncu_traffic.json is keyed by a source hash)" % kernel_key
    return "dram__bytes_read + dram__bytes_write per launch from the ncu --set full capture of this source revision"


class Setup:
    """8 GiB layer-separate source cache + 8 GiB contiguous destination cache on cuda:0 (torch owns the memory)."""

    def __init__(self, args, torch, lib, device=0, dst_contig_ptr=None, dst_blocks=None, src_ptrs=None):
        self.torch, self.lib, self.args = torch, lib, args
        n = args.cache_blocks
        self.n = n
        self.src_layers = []
        if src_ptrs is None:
            g = torch.Generator(device="cuda")
            g.manual_seed(42 + device)
            for l in range(L8B):
                t = torch.empty(2 * n * CHUNK, dtype=torch.uint8, device="cuda")
                t.random_(0, 256, generator=g)
                self.src_layers.append(t)
            self.src_ptrs = np.array([t.data_ptr() for t in self.src_layers], dtype=np.uint64)
        else:
            self.src_ptrs = np.asarray(src_ptrs, dtype=np.uint64)  # e.g. peer-mapped layers (pull mode)
        if dst_contig_ptr is None:
            self.dst = torch.zeros(n * BLOCK_BYTES, dtype=torch.uint8, device="cuda")
            dst_contig_ptr, dst_blocks = self.dst.data_ptr(), n
        self.dst_ptr, self.dst_blocks = dst_contig_ptr, dst_blocks
        self.plan = C.c_void_p()
        lib.check(lib.lib().kvdp_plan_contig(C.byref(self.plan), self.src_ptrs.ctypes.data, L8B, 0, n, CHUNK,
                                             C.c_void_p(dst_contig_ptr), dst_blocks, 1, device))
        self.pairs = BLOCKS_PER_REQ * args.requests
        assert self.pairs <= n and self.pairs <= dst_blocks
        self.xfer = C.c_void_p()
        lib.check(lib.lib().kvdp_xfer_create(C.byref(self.xfer), self.pairs, device))
        self.stream_ptr = lib.lib().kvdp_xfer_stream(self.xfer)
        self.stream = torch.cuda.ExternalStream(self.stream_ptr)
        rng = np.random.Generator(np.random.PCG64(1234 + device))
        self.sperm = rng.permutation(n).astype(np.int32)
        self.dperm = rng.permutation(dst_blocks).astype(np.int32)

    def ids(self, step):
        """A different window of the caches every step, so no step finds its data in L2."""
        w = max(1, self.n // self.pairs)
        o = (step % w) * self.pairs
        wd = max(1, self.dst_blocks // self.pairs)
        od = (step % wd) * self.pairs
        return self.sperm[o:o + self.pairs], self.dperm[od:od + self.pairs]

    def step(self, step, done_flag=None, done_value=0):
        s, d = self.ids(step)
        self.lib.check(self.lib.lib().kvdp_xfer_run(self.xfer, self.plan, s.ctypes.data, d.ctypes.data, self.pairs, 0,
                                                    L8B, 0, done_flag, done_value))

    def verify_local(self, step, samples=6):
        """Spot check without the oracle: destination chunks equal the source chunks they were mapped from."""
        torch = self.torch
        s, d = self.ids(step)
        rng = np.random.Generator(np.random.PCG64(step))
        for _ in range(samples):
            i, l, kv = int(rng.integers(self.pairs)), int(rng.integers(L8B)), int(rng.integers(2))
            src = self.src_layers[l][(kv * self.n + int(s[i])) * CHUNK:(kv * self.n + int(s[i]) + 1) * CHUNK]
            off = int(d[i]) * BLOCK_BYTES + (l * 2 + kv) * CHUNK
            if not torch.equal(src, self.dst[off:off + CHUNK]):
                raise RuntimeError("relayout result differs from its source (pair %d layer %d kv %d)" % (i, l, kv))


def time_steps(torch, setup, steps, warmup, barrier=None, clocks=None):
    for w in range(warmup):
        setup.step(w)
    torch.cuda.synchronize()
    if barrier:
        barrier()
    if clocks is not None:
        clocks.start()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    start.record(setup.stream)
    for k in range(steps):
        setup.step(warmup + k)
    end.record(setup.stream)
    torch.cuda.synchronize()
    if barrier:
        barrier()
    return start.elapsed_time(end) / 1e3  # seconds


def time_kernel_only(torch, lib, setup, launches, warmup=3):
    """Average duration of the relayout kernel alone (ids already on the device): the roofline numerator's clock."""
    s_all = torch.from_numpy(np.concatenate([setup.ids(k)[0] for k in range(launches + warmup)])).cuda()
    d_all = torch.from_numpy(np.concatenate([setup.ids(k)[1] for k in range(launches + warmup)])).cuda()
    p = setup.pairs
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(launches)]
    torch.cuda.synchronize()
    for k in range(launches + warmup):
        if k >= warmup:
            evs[k - warmup][0].record(setup.stream)
        lib.check(lib.lib().kvdp_plan_launch(setup.plan, s_all.data_ptr() + 4 * k * p, d_all.data_ptr() + 4 * k * p, p,
                                             0, L8B, setup.stream_ptr))
        if k >= warmup:
            evs[k - warmup][1].record(setup.stream)
    torch.cuda.synchronize()
    ms = [a.elapsed_time(b) for a, b in evs]
    return float(np.mean(ms)) / 1e3, float(np.min(ms)) / 1e3


def e2e_legacy(torch, lib, args, steps, warmup):
    """Reference-facing path, a full trip with HOST buffers through the legacy CopyStream C ABI
    (copy_stream_prepare_block_ids + copy_stream_memcpy per layer + copy_stream_sync, kv/layer.rs:785-979): every step
    ONBOARDS one batch of blocks (pinned host layer-separate cache -> HBM cache) and OFFLOADS another batch of the same
    size (HBM cache -> pinned host cache), each through its own CopyStream, as the reference's pool of CopyStreams
    allows; the two directions share the PCIe link in opposite directions. value = payload of both batches / wall time.
    Each direction is also timed alone against its own link ceiling. Block ids: the host caches hold twice the batch;
    every request's 64 blocks sit next to each other on the host side at a random request slot (how a host tier stores
    an offloaded request), the device-side ids are a random permutation. A second, fully scattered id set (random host
    blocks, no runs) is timed beside it."""
    L = lib.lib()
    pairs = BLOCKS_PER_REQ * args.requests
    nd, nh = pairs, 2 * pairs
    layer_bytes, host_layer_bytes = 2 * nd * CHUNK, 2 * nh * CHUNK

    def pinned_layers(fill):
        arrs, ptrs = [], []
        for l in range(L8B):
            p = C.c_void_p()
            lib.check(L.cuda_malloc_host(C.byref(p), host_layer_bytes))
            a = np.frombuffer((C.c_uint8 * host_layer_bytes).from_address(p.value), dtype=np.uint8)
            if fill:
                w = a.view(np.uint64)
                w[:] = np.arange(w.size, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15) + np.uint64(l)
            else:
                a[:] = 0
            arrs.append(a)
            ptrs.append(p.value)
        return arrs, ptrs
    host_in, host_in_p = pinned_layers(True)            # onboard source
    host_out, host_out_p = pinned_layers(False)         # offload destination
    dev_in = [torch.zeros(layer_bytes, dtype=torch.uint8, device="cuda") for _ in range(L8B)]
    g = torch.Generator(device="cuda")
    g.manual_seed(4242)
    dev_out = [torch.empty(layer_bytes, dtype=torch.uint8, device="cuda").random_(0, 256, generator=g)
               for _ in range(L8B)]
    cs_on, cs_off = C.c_void_p(), C.c_void_p()
    assert L.copy_stream_create(C.byref(cs_on), L8B, pairs) == 0
    assert L.copy_stream_create(C.byref(cs_off), L8B, pairs) == 0
    rng = np.random.Generator(np.random.PCG64(77))
    suffix = BS * H8B * D8B

    def host_ids(scattered):
        if scattered:
            return rng.permutation(nh)[:pairs].astype(np.int32)
        slots = rng.permutation(nh // BLOCKS_PER_REQ)[:args.requests]
        return np.concatenate([np.arange(s * BLOCKS_PER_REQ, (s + 1) * BLOCKS_PER_REQ) for s in slots]).astype(np.int32)

    def one_step(onboard=True, offload=True, scattered=False):
        # (onboard src = host, dst = device; offload src = device, dst = host)
        ids = [host_ids(scattered), rng.permutation(nd).astype(np.int32), rng.permutation(nd).astype(np.int32),
               host_ids(scattered)]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if onboard:
            lib.check(L.copy_stream_prepare_block_ids(cs_on, ids[0].ctypes.data, ids[1].ctypes.data, pairs))
        if offload:
            lib.check(L.copy_stream_prepare_block_ids(cs_off, ids[2].ctypes.data, ids[3].ctypes.data, pairs))
        for l in range(L8B):
            if onboard:
                lib.check(L.copy_stream_memcpy(cs_on, host_in_p[l], dev_in[l].data_ptr(), 2, suffix, ELEM, nh, nd))
            if offload:
                lib.check(L.copy_stream_memcpy(cs_off, dev_out[l].data_ptr(), host_out_p[l], 2, suffix, ELEM, nd, nh))
        if onboard:
            lib.check(L.copy_stream_sync(cs_on))
        if offload:
            lib.check(L.copy_stream_sync(cs_off))
        return time.perf_counter() - t0, ids

    t_total, timed, ids, step_ms = 0.0, 0, None, []
    for k in range(steps + warmup):
        dt, ids = one_step()
        if k >= warmup:
            t_total += dt
            timed += 1
            step_ms.append(round(dt * 1e3, 3))
    # every byte of the last step, both directions
    for l in range(L8B):
        want = torch.from_numpy(host_in[l]).view(2, nh, CHUNK)[:, torch.from_numpy(ids[0].astype(np.int64))].cuda()
        if not torch.equal(dev_in[l].view(2, nd, CHUNK)[:, torch.from_numpy(ids[1].astype(np.int64)).cuda()], want):
            raise RuntimeError("e2e onboard result differs from its source (layer %d)" % l)
        got = torch.from_numpy(host_out[l]).view(2, nh, CHUNK)[:, torch.from_numpy(ids[3].astype(np.int64))].cuda()
        if not torch.equal(dev_out[l].view(2, nd, CHUNK)[:, torch.from_numpy(ids[2].astype(np.int64)).cuda()], got):
            raise RuntimeError("e2e offload result differs from its source (layer %d)" % l)
    one_step(True, False)
    t_on = min(one_step(True, False)[0] for _ in range(3))
    one_step(False, True)
    t_off = min(one_step(False, True)[0] for _ in range(3))
    one_step(scattered=True)
    t_sc = min(one_step(scattered=True)[0] for _ in range(2))
    t_sc_on = min(one_step(True, False, True)[0] for _ in range(2))
    t_sc_off = min(one_step(False, True, True)[0] for _ in range(2))
    # comparator: the same call sequence with the reference's data movement behind it, one cudaMemcpyAsync per
    # (K|V, pair) chunk (block_copy.cu:246-305), which libkvdp keeps as host_path = 1
    dma_loop = None
    if lib.get_option("host_path") == 0:
        lib.set_option("host_path", 1)
        dma_loop = pairs * BLOCK_BYTES / min(one_step(True, False, True)[0] for _ in range(2)) / GB
        lib.set_option("host_path", 0)
    # practical ceilings of the host link on this box: contiguous 1 GiB pinned DMAs, each way alone and both at once
    n_c = 1 << 30
    big_h = C.c_void_p()
    lib.check(L.cuda_malloc_host(C.byref(big_h), 2 * n_c))
    big_d = torch.empty(2 * n_c, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def dma(h2d, d2h):
        best = 0.0
        for _ in range(3):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            if h2d:
                lib.check(L.cuda_memcpy_async(big_d.data_ptr(), big_h.value, n_c, C.c_void_p(s1.cuda_stream)))
            if d2h:
                lib.check(L.cuda_memcpy_async(big_h.value + n_c, big_d.data_ptr() + n_c, n_c, C.c_void_p(s2.cuda_stream)))
            torch.cuda.synchronize()
            best = max(best, (int(h2d) + int(d2h)) * n_c / (time.perf_counter() - t0) / GB)
        return best
    ceil_h2d, ceil_d2h, ceil_both = dma(True, False), dma(False, True), dma(True, True)
    L.cuda_free_host(big_h)
    L.copy_stream_destroy(cs_on)
    L.copy_stream_destroy(cs_off)
    for p in host_in_p + host_out_p:
        L.cuda_free_host(C.c_void_p(p))
    sec = t_total / timed
    payload = pairs * BLOCK_BYTES
    return {"value": 2 * payload / sec / GB, "unit": "GB/s", "h2d_bytes_per_step": payload, "d2h_bytes_per_step": payload,
            "ms_per_step": sec * 1e3, "steps": timed, "step_ms": step_ms,
            "step": "onboard %d blocks (pinned host -> HBM) + offload %d blocks (HBM -> pinned host), concurrently, "
                    "each through its own legacy CopyStream" % (pairs, pairs),
            "duplex_link_ceiling_GBps": ceil_both, "frac_of_duplex_link_ceiling": 2 * payload / sec / GB / ceil_both,
            "onboard": {"GBps": payload / t_on / GB, "h2d_link_ceiling_GBps": ceil_h2d,
                        "frac_of_link_ceiling": payload / t_on / GB / ceil_h2d},
            "offload": {"GBps": payload / t_off / GB, "d2h_link_ceiling_GBps": ceil_d2h,
                        "frac_of_link_ceiling": payload / t_off / GB / ceil_d2h},
            "h2d_link_ceiling_GBps": ceil_h2d, "frac_of_link_ceiling": payload / t_on / GB / ceil_h2d,
            "host_dma_piece_kib": lib.get_option("host_dma_piece_kib"),
            "host_path": {0: "auto: DMAs over each run of consecutive host blocks (pieces of one size) through HBM staging + one kernel per layer "
                             "when the host-side runs average >= 1 MiB, else SMs read / write pinned memory",
                          1: "per-chunk cudaMemcpyAsync loop", 2: "TMA kernel on pinned memory",
                          3: "SMs read / write pinned memory"}[lib.get_option("host_path")],
            "scattered_host_ids": {"both_GBps": 2 * payload / t_sc / GB, "onboard_GBps": payload / t_sc_on / GB,
                                   "offload_GBps": payload / t_sc_off / GB,
                                   "note": "512 random host blocks out of 1024 (no runs): zero-copy kernels; SM accesses "
                                           "to host memory are 128-byte PCIe requests (0.92 of the link, "
                                           "profiles/sweep_r02_host_paths.txt)"},
            "verification": "all bytes of the last step, both directions",
            "per_chunk_dma_loop_GBps": dma_loop,
            "per_chunk_dma_loop_note": "onboard alone with the reference's data movement behind the same ABI calls (one "
                                       "cudaMemcpyAsync per chunk, 4096 per 64-block request): kvdp_set_option(host_path, 1)",
            "api": "legacy C ABI: copy_stream_prepare_block_ids + 32 x copy_stream_memcpy + copy_stream_sync per "
                   "direction, timed on the host clock around the blocking calls"}


def host_tier_bench(torch, lib):
    """BASELINE config 4: offload / onboard of Llama-3-8B blocks between HBM (layer-separate) and a pinned
    contiguous-per-block host slab, mixed ISL 1k..32k (16..512 blocks per request), through kvdp_host_tier_*."""
    L = lib.lib()
    isl_blocks = [16, 32, 64, 128, 256, 512]          # ISL 1k, 2k, 4k, 8k, 16k, 32k = 1008 blocks = 7.9 GiB
    nb = sum(isl_blocks)
    dev_layers = [torch.empty(2 * nb * CHUNK, dtype=torch.uint8, device="cuda").random_(0, 256) for _ in range(L8B)]
    back_layers = [torch.zeros(2 * nb * CHUNK, dtype=torch.uint8, device="cuda") for _ in range(L8B)]
    host = C.c_void_p()
    lib.check(L.cuda_malloc_host(C.byref(host), nb * BLOCK_BYTES))
    ptrs = np.array([t.data_ptr() for t in dev_layers], dtype=np.uint64)
    ptrs_b = np.array([t.data_ptr() for t in back_layers], dtype=np.uint64)
    tier, tier_b = C.c_void_p(), C.c_void_p()
    lib.check(L.kvdp_host_tier_create(C.byref(tier), ptrs.ctypes.data, L8B, 0, nb, CHUNK, host, nb, 32, 0))
    lib.check(L.kvdp_host_tier_create(C.byref(tier_b), ptrs_b.ctypes.data, L8B, 0, nb, CHUNK, host, nb, 32, 0))
    rng = np.random.Generator(np.random.PCG64(4))
    dev_ids = rng.permutation(nb).astype(np.int32)      # scattered on the device
    back_ids = rng.permutation(nb).astype(np.int32)
    host_ids = np.arange(nb, dtype=np.int32)            # each request lands in a consecutive host range
    reqs, o = [], 0
    for n in isl_blocks:
        reqs.append((o, n))
        o += n

    def run(t, fn, a_ids, b_ids):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for o, n in reqs:
            lib.check(fn(t, a_ids[o:o + n].ctypes.data, b_ids[o:o + n].ctypes.data, n))
        lib.check(L.kvdp_host_tier_sync(t))
        return time.perf_counter() - t0
    run(tier, L.kvdp_host_tier_offload, dev_ids, host_ids)                      # warm-up
    t_off = min(run(tier, L.kvdp_host_tier_offload, dev_ids, host_ids) for _ in range(3))
    run(tier_b, L.kvdp_host_tier_onboard, host_ids, back_ids)
    t_on = min(run(tier_b, L.kvdp_host_tier_onboard, host_ids, back_ids) for _ in range(3))
    # round trip identity on sampled blocks
    for i in (0, 17, nb - 1):
        for l in (0, 13, L8B - 1):
            for kv in (0, 1):
                a = dev_layers[l][(kv * nb + int(dev_ids[i])) * CHUNK:(kv * nb + int(dev_ids[i]) + 1) * CHUNK]
                b = back_layers[l][(kv * nb + int(back_ids[i])) * CHUNK:(kv * nb + int(back_ids[i]) + 1) * CHUNK]
                if not torch.equal(a, b):
                    raise RuntimeError("host tier round trip mismatch")
    # link ceilings: one contiguous 1 GiB DMA each way
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ceil = {}
    for name, dst, src in (("d2h", host.value, dev_layers[0].data_ptr()), ("h2d", dev_layers[0].data_ptr(), host.value)):
        best = 0.0
        nbytes = min(1 << 30, 2 * nb * CHUNK)
        for _ in range(3):
            ev[0].record()
            lib.check(L.cuda_memcpy_async(dst, src, nbytes, None))
            ev[1].record()
            torch.cuda.synchronize()
            best = max(best, nbytes / (ev[0].elapsed_time(ev[1]) / 1e3) / GB)
        ceil[name] = best
    L.kvdp_host_tier_destroy(tier)
    L.kvdp_host_tier_destroy(tier_b)
    L.cuda_free_host(host)
    payload = nb * BLOCK_BYTES
    return {"workload": "Llama-3-8B blocks, requests of %s blocks (ISL 1k..32k mix), %.1f GiB per direction"
                        % (isl_blocks, nb * BLOCK_BYTES / 2 ** 30),
            "offload_GBps": payload / t_off / GB, "onboard_GBps": payload / t_on / GB,
            "d2h_link_ceiling_GBps": ceil["d2h"], "h2d_link_ceiling_GBps": ceil["h2d"],
            "offload_frac_of_ceiling": payload / t_off / GB / ceil["d2h"],
            "onboard_frac_of_ceiling": payload / t_on / GB / ceil["h2d"],
            "blocks_per_s_offload": nb / t_off, "blocks_per_s_onboard": nb / t_on,
            "staging": "2 x 32 blocks (2 x 256 MiB HBM), one DMA per request range, gather/scatter kernel per slot"}


def token_transfer_bench(torch, lib, steps=10):
    """token-granular transfer, the fused form of the NCCL connector's per-layer gather +
    send + reshape_and_cache. Llama-3-8B shape, 8 requests x 4096 tokens scattered over 64 Ki slots, every layer, K and
    V, 2 KiB rows (8 heads x 128 x bf16) in ONE launch; device to device on one GPU (HBM-bound: read + write)."""
    from dynamo_exp_b200.transfer import token_rows_plan
    L = lib.lib()
    row, slots, ntok = 8 * 128 * 2, 1024 * 64, 8 * 4096
    src = [torch.empty(2 * slots * row, dtype=torch.uint8, device="cuda").random_(0, 256) for _ in range(L8B)]
    dst = [torch.zeros(2 * slots * row, dtype=torch.uint8, device="cuda") for _ in range(L8B)]
    plan = token_rows_plan([t.data_ptr() for t in src], [t.data_ptr() for t in dst], slots, slots, row, row, row)
    g = torch.Generator(device="cpu").manual_seed(9)
    s_ids = torch.randperm(slots, generator=g)[:ntok].to(torch.int32).cuda()
    d_ids = torch.randperm(slots, generator=g)[:ntok].to(torch.int32).cuda()
    st = torch.cuda.Stream()

    def run():
        lib.check(L.kvdp_plan_launch(plan.handle, s_ids.data_ptr(), d_ids.data_ptr(), ntok, 0, L8B,
                                     C.c_void_p(st.cuda_stream)))
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for _ in range(steps):
        run()
    b.record(st)
    torch.cuda.synchronize()
    sec = a.elapsed_time(b) / 1e3 / steps
    for layer in (0, L8B - 1):          # spot check: K and V rows of a few tokens
        for kv in (0, 1):
            for i in (0, ntok // 2, ntok - 1):
                so = (kv * slots + int(s_ids[i])) * row
                do = (kv * slots + int(d_ids[i])) * row
                if not torch.equal(src[layer][so:so + row], dst[layer][do:do + row]):
                    raise RuntimeError("token transfer mismatch")
    payload = ntok * row * 2 * L8B
    plan.close()
    return {"payload_GBps": payload / sec / GB, "hbm_GBps_read_plus_write": 2 * payload / sec / GB,
            "tokens_per_s": ntok / sec, "ms_per_launch": sec * 1e3, "tokens": ntok, "run_bytes": row,
            "runs_per_launch": ntok * 2 * L8B}


def event_publish_bench(lib, n_events=20000, blocks_per_event=4, bs=64):
    """the engine-facing event ABI. One engine thread publishes stored events (4 full 64-token blocks
    each, chained parents) as fast as it can; the publisher thread hashes whatever queued up per GPU batch and applies
    the events to a prefix index. Wall clock from the first publish to the flush returning (host path: no device
    timer applies)."""
    L = lib.lib()
    rng = np.random.Generator(np.random.PCG64(7))
    toks = rng.integers(0, 128256, size=(n_events, blocks_per_event * bs), dtype=np.uint32)
    ix = C.c_void_p()
    lib.check(L.kvdp_index_create(C.byref(ix), 1 << 18, 64, 0))
    if L.dynamo_llm_init(b"bench", b"worker", 1, bs) != 0:
        raise RuntimeError("dynamo_llm_init failed")
    lib.check(L.kvdp_event_sink_index(ix, 0))
    counts = (C.c_size_t * blocks_per_event)(*([bs] * blocks_per_event))
    ids = (C.c_uint64 * blocks_per_event)()
    u32p = C.POINTER(C.c_uint32)
    b0 = [int(L.kvdp_event_stat(i)) for i in range(4)]
    t0 = time.perf_counter()
    for e in range(n_events):
        for j in range(blocks_per_event):
            ids[j] = e * blocks_per_event + j + 1
        if L.dynamo_kv_event_publish_stored(e, toks[e].ctypes.data_as(u32p), counts, ids, blocks_per_event, None, 0):
            raise RuntimeError("publish failed")
    t_pub = time.perf_counter() - t0
    lib.check(L.kvdp_event_flush())
    t_all = time.perf_counter() - t0
    b1 = [int(L.kvdp_event_stat(i)) for i in range(4)]
    nodes = int(L.kvdp_index_stat(ix, 0, 0))
    lib.check(L.kvdp_event_sink_index(None, 0))
    L.dynamo_llm_shutdown()
    L.kvdp_index_destroy(ix)
    assert b1[1] - b0[1] == n_events and nodes >= n_events
    return {"events_per_s": n_events / t_all, "blocks_per_s": n_events * blocks_per_event / t_all,
            "publish_call_us": t_pub / n_events * 1e6, "events": n_events, "blocks_per_event": blocks_per_event,
            "hash_batches": b1[2] - b0[2], "avg_blocks_per_batch": (b1[3] - b0[3]) / max(1, b1[2] - b0[2]),
            "note": "dynamo_kv_event_publish_stored from one Python thread (ctypes call overhead included); sink = "
                    "prefix index; batches form by themselves while the previous one is on the GPU"}


def router_index_bench(torch, lib, steps=10):
    """the router's prefix index. 8 workers hold prefixes of 256 synthetic documents; 57 344 requests
    (ISL 512..32k -> 8..512 block hashes each) share a document prefix of random length and then diverge. One launch
    scores the whole batch (RadixTree::find_matches, indexer.rs:239-277)."""
    L = lib.lib()
    rng = np.random.Generator(np.random.PCG64(2024))
    n_workers, n_docs, doc_blocks = 8, 256, 512

    def doc_hash(d, j):
        x = (np.uint64(d) << np.uint64(32)) | np.asarray(j, dtype=np.uint64)
        x = (x + np.uint64(0x9E3779B97F4A7C15)) * np.uint64(0xBF58476D1CE4E5B9)
        return x ^ (x >> np.uint64(31))
    ix = C.c_void_p()
    lib.check(L.kvdp_index_create(C.byref(ix), 1 << 20, 64, 0))
    t0 = time.perf_counter()
    events = 0
    with np.errstate(over="ignore"):
        for w in range(n_workers):
            for d in rng.permutation(n_docs)[:n_docs // 2]:
                depth = int(rng.integers(16, doc_blocks + 1))
                th = doc_hash(d, np.arange(depth))
                bh = th ^ np.uint64(w + 1)              # engine-side ids: distinct per worker
                lib.check(L.kvdp_index_apply_stored(ix, w, 0, 0, bh.ctypes.data, th.ctypes.data, depth))
                events += 1
    lib.check(L.kvdp_index_flush(ix))
    build_s = time.perf_counter() - t0
    isl_blocks = [8, 16, 32, 64, 128, 256, 512]
    n_req = 7 * 8192
    lens = np.array([isl_blocks[i % 7] for i in range(n_req)], dtype=np.int64)
    offs = np.zeros(n_req + 1, dtype=np.int64)
    offs[1:] = np.cumsum(lens)
    total = int(offs[-1])
    seqs = np.empty(total, dtype=np.uint64)
    docs = rng.integers(0, n_docs, size=n_req)
    share = (rng.random(n_req) * (lens + 1)).astype(np.int64)      # blocks shared with the document
    with np.errstate(over="ignore"):
        for r in range(n_req):
            n, k = int(lens[r]), int(share[r])
            seqs[offs[r]:offs[r] + k] = doc_hash(docs[r], np.arange(k))
            seqs[offs[r] + k:offs[r] + n] = doc_hash(1 << 20 | r, np.arange(n - k))
    d_seq = torch.from_numpy(seqs.view(np.int64)).cuda()
    d_off = torch.from_numpy(offs).cuda()
    d_scores = torch.zeros(n_req * n_workers, dtype=torch.int32, device="cuda")
    d_depth = torch.zeros(n_req, dtype=torch.int32, device="cuda")
    st = torch.cuda.Stream()     # an explicit stream: NULL would select the index's own stream

    def run():
        lib.check(L.kvdp_index_find_matches(ix, d_seq.data_ptr(), d_off.data_ptr(), n_req, 0, n_workers,
                                            d_scores.data_ptr(), d_depth.data_ptr(), C.c_void_p(st.cuda_stream)))
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for _ in range(steps):
        run()
    b.record(st)
    torch.cuda.synchronize()
    sec = a.elapsed_time(b) / 1e3 / steps
    depth = d_depth.cpu().numpy()
    probes = int(depth.sum()) + int((depth < lens).sum())          # matched steps + the miss that ends a walk
    nodes = int(L.kvdp_index_stat(ix, 0, 0))
    L.kvdp_index_destroy(ix)
    return {"requests_per_s": n_req / sec, "ms_per_batch": sec * 1e3, "requests": n_req, "workers": n_workers,
            "tree_nodes": nodes, "store_events": events, "tree_build_s": build_s,
            "avg_matched_blocks": float(depth.mean()), "probes_per_s": probes / sec,
            "bytes_per_probe": 32, "random_access_GBps": probes * 32 / sec / GB,
            "note": "one thread per request; each step = ONE dependent 32-byte random read (edge record carrying the "
                    "child's worker set); latency-bound on the longest walk (512 blocks), not bandwidth-bound"}


def hash_bench(torch, lib, steps, warmup, peak):
    """Router-sweep shaped batch (ISL 512..32k round-robin, 64-token blocks) resident in HBM; chained + block hashes."""
    from tests.util import mix64
    L = lib.lib()
    isls = [512, 1024, 2048, 4096, 8192, 16384, 32768]
    n_req = 7 * 8192          # 57 344 requests, 8.3 M blocks, 2.1 GB of tokens per batch (>> L2)
    lens = np.array([isls[i % 7] for i in range(n_req)], dtype=np.int64)
    offs = np.zeros(n_req + 1, dtype=np.int64)
    offs[1:] = np.cumsum(lens)
    total = int(offs[-1])
    nblk = total // BS
    # tokens generated on the device (torch RNG) — 0.5 GiB, larger than L2
    g = torch.Generator(device="cuda")
    g.manual_seed(0xD15A66)
    tok = torch.randint(0, 128256, (total,), dtype=torch.int32, device="cuda", generator=g)
    d_off = torch.from_numpy(offs).cuda()
    d_boff = torch.empty(n_req + 1, dtype=torch.int64, device="cuda")
    d_bh = torch.empty(nblk, dtype=torch.int64, device="cuda")
    d_sh = torch.empty(nblk, dtype=torch.int64, device="cuda")
    ws_bytes = L.kvdp_hash_workspace_bytes(n_req)
    d_ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    # Two streams, alternating batches, each with its own outputs: the serial chain kernel of batch k (latency-bound,
    # a few hundred warps) overlaps the HBM-bound block-hash kernel of batch k+1, as a router streaming 1M requests
    # through in batches would run it.
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    outs = [(d_boff, d_bh, d_sh, d_ws)]
    outs.append((torch.empty_like(d_boff), torch.empty_like(d_bh), torch.empty_like(d_sh), torch.empty_like(d_ws)))

    def run(salt, k=0):
        st = streams[k & 1]
        boff, bh, sh, ws = outs[k & 1]
        lib.check(L.kvdp_hash_token_blocks(tok.data_ptr(), d_off.data_ptr(), n_req, total, BS, salt, boff.data_ptr(),
                                           bh.data_ptr(), sh.data_ptr(), ws.data_ptr(), ws_bytes, 0,
                                           C.c_void_p(st.cuda_stream)))
    torch.cuda.synchronize()
    for k in range(warmup):
        run(1337, k)
    torch.cuda.synchronize()
    # (1) latency of one batch on one stream
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(streams[0])
    for k in range(steps):
        run(1337 + k, 0)
    b.record(streams[0])
    torch.cuda.synchronize()
    sec_serial = a.elapsed_time(b) / 1e3 / steps
    # spot parity inside the timed configuration, against an implementation that is neither ours nor the oracle
    try:
        import xxhash
        run(1337, 0)
        torch.cuda.synchronize()
        t_host = tok[:4 * BS].cpu().numpy().astype(np.uint32)
        got_bh = outs[0][1][:4].cpu().numpy().view(np.uint64)
        got_sh = outs[0][2][:4].cpu().numpy().view(np.uint64)
        parent = None
        for bi in range(4):
            want = xxhash.xxh3_64_intdigest(t_host[bi * BS:(bi + 1) * BS].tobytes(), seed=1337)
            seq = want if parent is None else xxhash.xxh3_64_intdigest(
                int(parent).to_bytes(8, "little") + int(want).to_bytes(8, "little"), seed=1337)
            if int(got_bh[bi]) != want or int(got_sh[bi]) != seq:
                raise RuntimeError("hash bench: GPU hashes differ from xxhash")
            parent = seq
        spot = "first 4 blocks of request 0 equal python-xxhash (block and chained)"
    except ImportError:
        spot = "python xxhash not importable: no in-bench spot check"
    # (2) throughput with batches alternating over the two streams
    start = torch.cuda.Event(enable_timing=True)
    ends = [torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)]
    start.record(streams[0])
    streams[1].wait_event(start)
    n_pipe = 2 * steps
    for k in range(n_pipe):
        run(1337 + k, k)
    ends[0].record(streams[0])
    ends[1].record(streams[1])
    torch.cuda.synchronize()
    sec = max(start.elapsed_time(ends[0]), start.elapsed_time(ends[1])) / 1e3 / n_pipe
    # (3) router form: un-chained seed-1337 LocalBlockHash only (kv_router/indexer.rs:123-135), one stream
    def run_unchained(salt):
        st = streams[0]
        boff, bh, sh, ws = outs[0]
        lib.check(L.kvdp_hash_token_blocks(tok.data_ptr(), d_off.data_ptr(), n_req, total, BS, salt, boff.data_ptr(),
                                           bh.data_ptr(), None, ws.data_ptr(), ws_bytes, 0,
                                           C.c_void_p(st.cuda_stream)))
    run_unchained(1337)
    torch.cuda.synchronize()
    a.record(streams[0])
    for k in range(steps):
        run_unchained(1337)
    b.record(streams[0])
    torch.cuda.synchronize()
    sec_un = a.elapsed_time(b) / 1e3 / steps
    out = {"hashes_per_s": nblk / sec_serial, "tokens_per_s": total / sec_serial, "ms_per_batch": sec_serial * 1e3,
           "requests": n_req, "blocks": nblk, "launches_per_batch": 5, "batches_timed": steps,
           "requests_timed": n_req * steps, "spot_check": spot,
           "semantics": "lib/tokens: block_hash + chained sequence_hash per full 64-token block, one stream",
           "two_stream_pipeline": {"hashes_per_s": nblk / sec, "ms_per_batch": sec * 1e3},
           "unchained_router_form": {"hashes_per_s": nblk / sec_un, "ms_per_batch": sec_un * 1e3,
                                     "semantics": "kv_router LocalBlockHash (seed 1337), scan + block-hash kernels"},
           # dominant kernel = the block-hash kernel; measured through the un-chained call (scan kernels included,
           # so this is a lower bound for the kernel itself); 264 B/block = 256 read + 8 written
           "roofline": {"bound": "hbm", "achieved": nblk * 264 / sec_un / GB, "peak": peak, "unit": "GB/s",
                        "frac": nblk * 264 / sec_un / GB / peak, "kernel": "kvdp_hash_blocks_tma<4,8>",
                        "algorithmic_bytes_per_launch": nblk * 264,
                        "traffic": ncu_traffic("kvdp_hash_blocks_tma"),
                        "traffic_note": ncu_traffic_note("kvdp_hash_blocks_tma")},
           "chained_hbm_frac": nblk * 272 / sec_serial / GB / peak}
    # e2e: pinned host tokens -> device -> hashes back on the host, through kvdp_hash_token_blocks_host
    n_e = 7 * 1024           # 7 168 requests, 266 MB of tokens up, 17 MB of hashes down per call
    lens_e = lens[:n_e]
    offs_e = np.zeros(n_e + 1, dtype=np.int64)
    offs_e[1:] = np.cumsum(lens_e)
    tot_e = int(offs_e[-1])
    h_tok = torch.randint(0, 128256, (tot_e,), dtype=torch.int32).pin_memory()
    nb_e = tot_e // BS
    h_boff = torch.empty(n_e + 1, dtype=torch.int64).pin_memory()
    h_bh = torch.empty(nb_e, dtype=torch.int64).pin_memory()
    h_sh = torch.empty(nb_e, dtype=torch.int64).pin_memory()

    def run_host():
        lib.check(L.kvdp_hash_token_blocks_host(h_tok.data_ptr(), offs_e.ctypes.data, n_e, BS, 1337, h_boff.data_ptr(),
                                                h_bh.data_ptr(), h_sh.data_ptr(), 0, -1))
    run_host()
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        run_host()
    dt = (time.perf_counter() - t0) / reps
    out["e2e"] = {"hashes_per_s": nb_e / dt, "h2d_bytes_per_batch": tot_e * 4 + (n_e + 1) * 8,
                  "d2h_bytes_per_batch": nb_e * 16 + (n_e + 1) * 8, "ms_per_batch": dt * 1e3}
    return out


def run_single(args):
    import torch

    from dynamo_exp_b200 import _lib as lib
    peak, peak_src = measured_peaks()
    setup = Setup(args, torch, lib)
    clocks = ClockSampler(0)
    sec = time_steps(torch, setup, args.steps, args.warmup, clocks=clocks)
    clk = clocks.stop()
    setup.verify_local(args.warmup + args.steps - 1)
    payload = setup.pairs * BLOCK_BYTES
    value = payload * args.steps / sec / GB
    k_avg, k_min = time_kernel_only(torch, lib, setup, max(args.steps, 10))
    path = lib.get_option("copy_path")
    kname = {0: "kvdp_copy_tma_tma_kernel", 1: "kvdp_copy_vec_kernel", 2: "kvdp_copy_tma_in_kernel",
             3: "kvdp_copy_tma_tma_kernel"}[path]
    line = {
        "metric": "kv_transfer_GBps", "value": value, "unit": "GB/s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": workload_config(args),
        "blocks_per_s": setup.pairs * args.steps / sec,
        "roofline": {"bound": "hbm", "achieved": 2 * payload / k_avg / GB, "peak": peak, "unit": "GB/s",
                     "frac": 2 * payload / k_avg / GB / peak, "traffic": ncu_traffic(kname),
                     "traffic_note": ncu_traffic_note(kname), "kernel": kname,
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": 2 * payload,
                     "avg_launch_ms": k_avg * 1e3, "best_launch_ms": k_min * 1e3,
                     "frac_of_nominal_8TBps": 2 * payload / k_avg / GB / 8000.0},
        "gpu_launches": args.steps,
        "clocks": clk,
        "env": env_info(torch, 0),
    }
    del setup
    torch.cuda.empty_cache()
    if not args.no_e2e:
        line["e2e"] = e2e_legacy(torch, lib, args, min(args.steps, 8), max(4, min(args.warmup, 8)))
    if not args.no_e2e:
        line["host_tier"] = host_tier_bench(torch, lib)
        torch.cuda.empty_cache()
    if not args.no_hash:
        line["hash"] = hash_bench(torch, lib, 18, 3, peak)   # 18 x 57 344 = 1.03 M requests per timed loop
        line["hashes_per_s"] = line["hash"]["hashes_per_s"]
        line["router_index"] = router_index_bench(torch, lib)
        line["event_publish"] = event_publish_bench(lib)
        line["token_transfer"] = token_transfer_bench(torch, lib)
        torch.cuda.empty_cache()
    if not args.no_cpu:
        from oracle import kvoracle as ko
        threads = ko.hw_threads()
        from bench import CpuRelayout
        blocks = BLOCKS_PER_REQ * args.requests
        leg = CpuRelayout(blocks, min(args.cache_blocks, 2 * blocks))
        v_all, _ = cpu_relayout_leg(threads, 6.0, leg=leg)
        v_one, _ = cpu_relayout_leg(1, 3.0, leg=leg)
        del leg
        line["cpu_baseline"] = {
            "value": v_all, "unit": "GB/s", "cores": threads, "kind": "port",
            "sample": "the bench step itself (%d Llama-3-8B blocks = %d MiB payload, layer-separate -> contiguous) done by "
                      "the C restatement of the reference's host copy loop (oracle/layout_oracle.c) on host-resident "
                      "caches, repeated for >= 6 s" % (blocks, blocks * BLOCK_BYTES >> 20),
            "single_thread_value": v_one,
            "hash_hashes_per_s": cpu_hash_leg(threads, 3.0), "hash_hashes_per_s_single_thread": cpu_hash_leg(1, 2.0)}
    print(json.dumps(line), flush=True)


def run_sweep(args):
    """Exploration table (not the bench line): every copy path x tile x CTAs/SM on the N=1 workload."""
    import torch

    from dynamo_exp_b200 import _lib as lib
    setup = Setup(args, torch, lib)
    payload = setup.pairs * BLOCK_BYTES
    rows = []
    combos = []
    for path in (1, 2, 3):
        for tile in (8192, 16384, 32768):
            for cps in ((2, 4, 8) if path == 1 else (1, 2, 3, 4, 6)):
                for stages in ((0,) if path != 2 else (3, 4, 6)):
                    combos.append((path, tile, cps, stages))
    for path, tile, cps, stages in combos:
        try:
            lib.set_option("copy_path", path)
            lib.set_option("tile_bytes", tile)
            lib.set_option("ctas_per_sm", cps)
            lib.set_option("stages", stages)
            k_avg, k_min = time_kernel_only(torch, lib, setup, 6, warmup=2)
            setup.verify_local(7, samples=2)
            rows.append((path, tile, cps, stages, 2 * payload / k_avg / GB, 2 * payload / k_min / GB))
        except Exception as e:  # noqa: BLE001
            rows.append((path, tile, cps, stages, -1.0, -1.0))
            print("sweep point failed:", path, tile, cps, stages, e, flush=True)
            break
    print("path tile ctas/sm stages  avg_GB/s(r+w)  best_GB/s(r+w)")
    for r in sorted(rows, key=lambda r: -r[4]):
        print("%4d %6d %4d %4d  %10.1f  %10.1f" % r)
