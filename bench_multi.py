"""bench_multi.py — the N>1 arm of bench.py: KV blocks move from prefill GPUs to decode GPUs over NVLink.

Ranks [0, P) are prefill, ranks [P, N) decode (P = N // 2); prefill rank r is paired with decode rank P + r % D.
Each step moves one batch (R requests x 64 Llama-3-8B-shaped blocks) from the prefill GPU's layer-separate cache into
the decode GPU's contiguous-per-block cache with ONE launch of the relayout kernel, addressing the other GPU's memory
through CUDA-IPC peer mappings, and the kernel's last CTA publishes a completion flag in the other GPU's memory:
  push : the kernel runs on the prefill GPU and STORES into the decode cache  (the reference's NIXL WRITE, write_blocks)
  pull : the kernel runs on the decode GPU and LOADS from the prefill cache   (the reference's NIXL READ, read_blocks)
No NCCL on the data path. NCCL send/recv of the same payload (the comparison the north star names) and a plain
cudaMemcpyAsync to the peer (the link's practical ceiling) are timed beside it.
"""
from __future__ import annotations

import ctypes as C
import json
import os

import numpy as np

from bench import BLOCK_BYTES, BLOCKS_PER_REQ, CHUNK, GB, L8B, ClockSampler, env_info, workload_config
from bench_single import Setup

NVLINK_PEAK_FALLBACK = 770.0   # GB/s per direction per GPU, peer copy (B200_PROFILING.md); used only when the in-run
                               # cudaMemcpyAsync-to-peer comparator is skipped (--no-comparators)
NVLINK_PEAK_NOMINAL = 900.0


def synth_words(torch, seed: int, nbytes: int):
    """`nbytes` (a multiple of 8) of synthetic payload as int64 words: a closed-form mix of (seed, word index), so the
    GPU that receives a transfer can regenerate what its peer holds and compare every byte."""
    x = torch.arange(nbytes // 8, dtype=torch.int64, device="cuda")
    x = x * 0x9E3779B97F4A7C1 + (((seed + 1) * 0x2545F4914F6CDD1D) & 0x7FFFFFFFFFFFFFFF)      # wraps mod 2^64
    x = x ^ (x >> 29)
    x = x * (0x94D049BB133111EB - (1 << 64))
    return x ^ (x >> 32)


class CudaArray:
    """Minimal __cuda_array_interface__ carrier so torch can view memory owned by libkvdp.so."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def run_multi(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist

    from dynamo_exp_b200 import _lib as lib
    L = lib.lib()
    from dynamo_exp_b200.connector import split_roles
    prefill_ranks, decode_ranks, role_pairs = split_roles(world, interleaved=args.pairing == "interleaved")
    if args.cpu_affinity:
        try:                                  # run (and first-touch pinned memory) on the CPUs next to this GPU
            import pynvml
            pynvml.nvmlInit()
            pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local_rank))
        except Exception:
            pass
    P, D = len(prefill_ranks), len(decode_ranks)
    is_prefill = rank in prefill_ranks
    partner = dict(role_pairs)[rank] if is_prefill else [p for p, d in role_pairs if d == rank][0]
    pairs = BLOCKS_PER_REQ * args.requests
    n = args.cache_blocks
    steps, warmup = args.steps, args.warmup

    def export(ptr):
        h = (C.c_uint8 * lib.IPC_HANDLE_BYTES)()
        lib.check(L.kvdp_ipc_export(ptr, h))
        return bytes(h)

    def imp(hbytes):
        hb = (C.c_uint8 * lib.IPC_HANDLE_BYTES).from_buffer_copy(hbytes)
        p = C.c_void_p()
        lib.check(L.kvdp_ipc_import(hb, local_rank, C.byref(p)))
        return p

    # ---- memory the other side must reach: allocated by libkvdp (cudaMalloc) and exported over CUDA IPC
    flag_ptr = C.c_void_p()
    lib.check(L.kvdp_device_alloc(C.byref(flag_ptr), 256, local_rank))
    handles = {"flag": export(flag_ptr)}
    own_layers, own_cache = [], C.c_void_p()
    def layer_seed(prefill_rank, l):
        return 1000 * (prefill_rank + 1) + l

    if is_prefill:
        for l in range(L8B):
            p = C.c_void_p()
            lib.check(L.kvdp_device_alloc(C.byref(p), 2 * n * CHUNK, local_rank))
            torch.as_tensor(CudaArray(p.value, 2 * n * CHUNK), device="cuda").view(torch.int64).copy_(
                synth_words(torch, layer_seed(rank, l), 2 * n * CHUNK))
            own_layers.append(p)
        handles["layers"] = [export(p) for p in own_layers]
    else:
        lib.check(L.kvdp_device_alloc(C.byref(own_cache), n * BLOCK_BYTES, local_rank))
        handles["cache"] = export(own_cache)
    all_handles = [None] * world
    dist.all_gather_object(all_handles, handles)
    peer_flag = imp(all_handles[partner]["flag"])
    if is_prefill:
        peer_cache = imp(all_handles[partner]["cache"])
        setup = Setup(args, torch, lib, device=local_rank, dst_contig_ptr=peer_cache.value, dst_blocks=n,
                      src_ptrs=[p.value for p in own_layers])           # push plan
    else:
        peer_layers = [imp(h).value for h in all_handles[partner]["layers"]]
        setup = Setup(args, torch, lib, device=local_rank, dst_contig_ptr=own_cache.value, dst_blocks=n,
                      src_ptrs=peer_layers)                             # pull plan
    wait_stream = torch.cuda.Stream()
    dist.barrier()
    seq = {"v": 0}

    def timed(mode, k0, count, su=None):
        """Barrier + device sync on both sides, CUDA events on the launching stream, MAX over ranks. The passive side
        stream-waits on the completion flag the kernel's last CTA writes into its memory."""
        su = su or setup
        runner = is_prefill if mode == "push" else not is_prefill
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st = su.stream if runner else wait_stream
        a.record(st)
        if runner:
            for k in range(count):
                seq["v"] += 1
                su.step(k0 + k, done_flag=peer_flag, done_value=seq["v"])
        else:
            seq["v"] += count
            lib.check(L.kvdp_stream_wait_flag(C.c_void_p(wait_stream.cuda_stream), flag_ptr, seq["v"]))
        b.record(st)
        torch.cuda.synchronize()
        dist.barrier()
        t = torch.tensor([a.elapsed_time(b) / 1e3], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def verify(mode, step):
        """Every byte of the step's blocks: the decode rank regenerates its partner's layers from their seeds and compares
        all 32 layers x K|V of every (source id -> destination id) pair with what the transfer left in its cache."""
        bad = torch.zeros(1, dtype=torch.int64, device="cuda")
        if not is_prefill:
            rng_p = np.random.Generator(np.random.PCG64(1234 + (rank if mode == "pull" else partner)))
            sperm = rng_p.permutation(n).astype(np.int32)
            dperm = rng_p.permutation(n).astype(np.int32)
            w = max(1, n // pairs)
            o = (step % w) * pairs
            s_t = torch.from_numpy(sperm[o:o + pairs].astype(np.int64)).cuda()
            d_t = torch.from_numpy(dperm[o:o + pairs].astype(np.int64)).cuda()
            cache = torch.as_tensor(CudaArray(own_cache.value, n * BLOCK_BYTES), device="cuda").view(torch.int64)
            cache = cache.view(n, L8B, 2, CHUNK // 8)
            for l in range(L8B):
                src = synth_words(torch, layer_seed(partner, l), 2 * n * CHUNK).view(2, n, CHUNK // 8)
                bad += (src[:, s_t].transpose(0, 1) != cache[d_t, l]).sum()
        dist.all_reduce(bad, op=dist.ReduceOp.SUM)
        if int(bad.item()) != 0:
            raise RuntimeError("%s: %d 8-byte words differ from their source" % (mode, int(bad.item())))
        return P * pairs * BLOCK_BYTES

    modes = ["push", "pull"] if args.mode == "both" else [args.mode]
    results, clk = {}, None
    for mode in modes:
        if not is_prefill:
            lib.check(L.kvdp_memset(own_cache, 0, n * BLOCK_BYTES))  # each mode must fill the cache itself
        timed(mode, 0, warmup)
        clocks = ClockSampler(local_rank)
        clocks.start()
        sec = timed(mode, warmup, steps)
        c = clocks.stop()
        clk = clk or c
        verified = verify(mode, warmup + steps - 1)
        results[mode] = {"value": P * pairs * BLOCK_BYTES * steps / sec / GB, "sec": sec,
                         "per_direction": pairs * BLOCK_BYTES * steps / sec / GB * max(1, P // D),
                         "verified_bytes_last_step": verified}

    # ---- e2e: a full trip with HOST buffers on both ends. Source blocks sit in pinned host memory of the prefill side;
    # ONE kernel per step reads them over PCIe and stores them into the decode GPU's cache over NVLink (no bounce through
    # prefill HBM) and publishes the step's completion flag there; the decode side's stream waits on that flag and copies
    # the step's blocks out to ITS pinned host memory. Steps pipeline: step k+1 crosses NVLink while step k goes down.
    e2e = None
    if not args.no_e2e:
        import copy
        nh = 128                                      # 1 GiB of pinned blocks per prefill rank
        e_args = copy.copy(args)
        e_args.cache_blocks, e_args.requests = nh, nh // BLOCKS_PER_REQ
        host_ptrs, e_setup, host_out = [], None, C.c_void_p()
        e_path = getattr(args, "e2e_path", "dma")
        nhc = 2 * nh                                  # host cache: twice the step's blocks
        dst_layer_bytes = 2 * nh * CHUNK              # the first 1 GiB of the decode cache, viewed as 32 layer slabs
        cs = C.c_void_p()
        # --e2e-links both (default): the prefill side's host blocks live in memory BOTH workers of the pair map and pin
        # (shared_host.SharedHostRegion); the prefill GPU relays its share of a step's requests over NVLink and the decode
        # GPU fetches the rest over its OWN PCIe link - two links instead of one for the same onboard.
        # "auto" times the onboard leg both ways and headlines the faster (on the 8-GPU boxes four uploading GPUs already take
        # what the host delivers; eight take slightly less).
        links_flag = getattr(args, "e2e_links", "auto")
        links = 2 if (links_flag in ("both", "auto") and e_path == "dma") else 1
        from dynamo_exp_b200.shared_host import SharedHostRegion, split_for_links
        n_req_e = nh // BLOCKS_PER_REQ
        use = {"links": links}                        # links in use by the leg being timed
        host_layer_bytes = 2 * nhc * CHUNK
        region, handles_e, links_fallback = None, [None] * world, None
        if links == 2:
            # every rank must agree: if one cannot create / open / register the shared region (another pid namespace, a
            # locked-memory limit), all go back to one link per pair with cuda_malloc_host memory and the line says so
            err = None
            if is_prefill:
                try:
                    if os.environ.get("KVDP_BENCH_NO_SHARED_HOST"):      # exercises the fallback below
                        raise RuntimeError("disabled by KVDP_BENCH_NO_SHARED_HOST")
                    region = SharedHostRegion.create(L8B * host_layer_bytes)
                except Exception as e:  # noqa: BLE001
                    err = "create: %s" % e
            dist.all_gather_object(handles_e, region.handle() if region is not None else None)
            if not is_prefill:
                try:
                    if handles_e[partner] is None:
                        raise RuntimeError("the prefill side has no region")
                    region = SharedHostRegion.open(handles_e[partner])
                except Exception as e:  # noqa: BLE001
                    err = "open: %s" % e
            errs = [None] * world
            dist.all_gather_object(errs, err)
            if any(e_ is not None for e_ in errs):
                links_fallback = "; ".join("rank %d %s" % (i_, e_) for i_, e_ in enumerate(errs) if e_)
                if region is not None:
                    region.close()
                    region = None
                links = 1
                use["links"] = 1

        def step_ids(pair_seed, k):
            """The step's (host ids, decode-cache ids), the same on both workers of a pair."""
            r = np.random.Generator(np.random.PCG64([4321, pair_seed, k]))
            slots_ = r.permutation(nhc // BLOCKS_PER_REQ)[:n_req_e]
            sid = np.concatenate([np.arange(s_ * BLOCKS_PER_REQ, (s_ + 1) * BLOCKS_PER_REQ) for s_ in slots_]).astype(np.int32)
            return sid, r.permutation(nh).astype(np.int32)
        if is_prefill:
            for l in range(L8B):
                if region is not None:
                    ptr_l = region.ptr + l * host_layer_bytes
                else:
                    p = C.c_void_p()
                    lib.check(L.cuda_malloc_host(C.byref(p), host_layer_bytes))
                    ptr_l = p.value
                a = np.frombuffer((C.c_uint8 * host_layer_bytes).from_address(ptr_l), dtype=np.uint8)
                a[:] = (l * 11 + rank) & 0xFF
                host_ptrs.append(ptr_l)
            if e_path == "kernel":
                # ONE zero-copy kernel per step: SMs read host memory and store into the peer (contiguous-per-block)
                e_args.cache_blocks = nhc
                e_setup = Setup(e_args, torch, lib, device=local_rank, dst_contig_ptr=peer_cache.value, dst_blocks=nh,
                                src_ptrs=[p for p in host_ptrs])
            else:
                assert L.copy_stream_create(C.byref(cs), L8B, nh) == 0
        else:
            lib.check(L.cuda_malloc_host(C.byref(host_out), nh * BLOCK_BYTES))
            C.memset(host_out.value, 0, nh * BLOCK_BYTES)
            if links == 2:
                host_ptrs = [region.ptr + l * host_layer_bytes for l in range(L8B)]
                assert L.copy_stream_create(C.byref(cs), L8B, nh) == 0

        def link_step(k, cache_base):
            """This worker's share of step k through the legacy CopyStream: requests of 64 consecutive host blocks at random
            request slots -> copy engines into HBM staging, one kernel per layer storing into the decode cache (the peer
            mapping over NVLink on the prefill side, local memory on the decode side)."""
            share = split_for_links(n_req_e, use["links"])    # [(first, last request) per link]; link 0 = the prefill GPU
            my_share = share[0] if is_prefill else share[use["links"] - 1]
            b0, b1 = my_share[0] * BLOCKS_PER_REQ, my_share[1] * BLOCKS_PER_REQ
            sid, did = step_ids(min(rank, partner), k)
            sid, did = sid[b0:b1].copy(), did[b0:b1].copy()
            if b1 > b0:
                lib.check(L.copy_stream_prepare_block_ids(cs, sid.ctypes.data, did.ctypes.data, b1 - b0))
                for l in range(L8B):
                    lib.check(L.copy_stream_memcpy(cs, host_ptrs[l], cache_base + l * dst_layer_bytes, 2, CHUNK // 2, 2,
                                                   nhc, nh))

        def prefill_step(k, value):
            if e_path == "kernel":
                e_setup.step(k, done_flag=peer_flag, done_value=value)
                return
            link_step(k, peer_cache.value)
            lib.check(L.copy_stream_sync(cs))
            lib.check(L.kvdp_xfer_signal(setup.xfer, peer_flag, value))

        def timed_e2e(k0, count, download="full"):
            torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            st = ((e_setup.stream if e_path == "kernel" else setup.stream) if is_prefill else wait_stream)
            a.record(st)
            for k in range(count):
                seq["v"] += 1
                if is_prefill:
                    prefill_step(k0 + k, seq["v"])
                else:
                    if use["links"] == 2:               # this GPU's own link carries its share of the step
                        link_step(k0 + k, own_cache.value)
                        lib.check(L.copy_stream_sync(cs))
                    lib.check(L.kvdp_stream_wait_flag(C.c_void_p(wait_stream.cuda_stream), flag_ptr, seq["v"]))
                    # "full": the step's blocks go on to this side's pinned host memory; "probe": 8 bytes of the step's
                    # result are read back (the onboard leg's device -> host read)
                    lib.check(L.cuda_memcpy_async(host_out, own_cache, nh * BLOCK_BYTES if download == "full" else 8,
                                                  C.c_void_p(wait_stream.cuda_stream)))
            b.record(st)
            torch.cuda.synchronize()
            dist.barrier()
            t = torch.tensor([a.elapsed_time(b) / 1e3], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        k_e = max(2, min(steps, 20))     # 1 GiB steps: enough of them that filling and draining the two-stage pipe is small
        if not is_prefill:
            lib.check(L.kvdp_memset(own_cache, 0, nh * BLOCK_BYTES))
        bad = torch.zeros(1, dtype=torch.int64, device="cuda")

        def check_bytes(flat):
            got = flat.view(nh, L8B, 2 * CHUNK) if e_path == "kernel" else flat.view(L8B, 1, 2 * nh * CHUNK).transpose(0, 1)
            return sum(int((got[:, l] != ((l * 11 + partner) & 0xFF)).sum()) for l in range(L8B))
        # onboard leg with every link setting asked for; each run starts from a zeroed decode cache and is checked byte by byte
        sec_by_links, k_next = {}, 0
        for n_links in ([2, 1] if (links == 2 and links_flag == "auto") else [links]):
            use["links"] = n_links
            if not is_prefill:
                lib.check(L.kvdp_memset(own_cache, 0, nh * BLOCK_BYTES))
            timed_e2e(k_next, 1, download="probe")
            sec_by_links[n_links] = timed_e2e(k_next + 1, k_e, download="probe")
            k_next += 1 + k_e
            if not is_prefill:      # what the onboard leg left in the decode GPU's cache
                bad += check_bytes(torch.as_tensor(CudaArray(own_cache.value, nh * BLOCK_BYTES), device="cuda"))
        best_links = min(sec_by_links, key=sec_by_links.get)
        sec_on = sec_by_links[best_links]
        use["links"] = best_links
        if not is_prefill:
            lib.check(L.kvdp_memset(own_cache, 0, nh * BLOCK_BYTES))
        timed_e2e(k_next, 1)
        sec_full = timed_e2e(k_next + 1, k_e)
        if not is_prefill:      # what arrived in host memory: layer l of every block is the partner's fill byte of layer l
            bad += check_bytes(torch.from_numpy(np.frombuffer((C.c_uint8 * (nh * BLOCK_BYTES)).from_address(host_out.value),
                                                              dtype=np.uint8)))
        dist.all_reduce(bad, op=dist.ReduceOp.SUM)
        if int(bad.item()) != 0:
            raise RuntimeError("e2e: %d bytes on the decode side differ from the prefill-side host source" % int(bad.item()))
        links = best_links
        share = split_for_links(n_req_e, links)
        prefill_api = ("legacy CopyStream on the prefill side (copy_stream_prepare_block_ids + 32 x copy_stream_memcpy from "
                       "pinned host layers into the PEER-mapped decode cache: copy engines host -> HBM staging, one kernel "
                       "per layer staging -> peer over NVLink; copy_stream_sync + kvdp_xfer_signal of the completion flag)"
                       if e_path != "kernel" else
                       "kvdp_plan_contig over pinned host layer pointers (prefill side) -> peer-mapped decode cache, "
                       "kvdp_xfer_run with in-kernel completion flag, one zero-copy launch per step")
        if links == 2:
            prefill_api += ("; links=both: the host blocks sit in memory both workers of the pair map and pin "
                            "(SharedHostRegion: memfd + kvdp_host_register), the prefill GPU carries requests %s of each step "
                            "and the decode GPU fetches requests %s through its own CopyStream over its own PCIe link"
                            % (list(range(*share[0])), list(range(*share[1]))))
        e2e = {"value": P * nh * BLOCK_BYTES * k_e / sec_on / GB, "unit": "GB/s",
               "leg": "onboard: blocks held in the prefill side's pinned host memory are delivered into the decode GPU's "
                      "cache (PCIe on the prefill GPU, NVLink to the decode GPU%s); the decode side waits on the step's "
                      "completion flag and reads 8 bytes of the result back to its host"
                      % ("; with links=both the decode GPU fetches half of each step's requests itself over its own "
                         "PCIe link" if links == 2 else ""),
               "links": "both" if links == 2 else "prefill",
               "links_fallback": links_fallback,
               "onboard_GBps_by_links": {("both" if n_ == 2 else "prefill"): P * nh * BLOCK_BYTES * k_e / t_ / GB
                                         for n_, t_ in sec_by_links.items()},
               "links_note": "--e2e-links auto times the onboard leg with the decode GPU's own link helping (both) and "
                             "without (prefill) and headlines the faster; the full trip uses that setting",
               "h2d_bytes_per_step": P * nh * BLOCK_BYTES, "d2h_bytes_per_step": 8 * D, "steps": k_e,
               "verification": "all bytes of the step's region of every decode cache, and of the decode-side host "
                               "buffers in the full trip",
               "path": e_path,
               "full_trip_host_to_host": {
                   "value": P * nh * BLOCK_BYTES * k_e / sec_full / GB, "unit": "GB/s", "steps": k_e,
                   "h2d_bytes_per_step": P * nh * BLOCK_BYTES, "d2h_bytes_per_step": P * nh * BLOCK_BYTES,
                   "note": "the same steps with the decode side copying every step's blocks on to ITS pinned host memory "
                           "(kvdp_stream_wait_flag + cuda_memcpy_async), pipelined with the next step. No deployment "
                           "moves KV host -> GPU -> NVLink -> GPU -> host; with several pairs it is bounded by the host, "
                           "whose memory is read by every prefill GPU and written by every decode GPU at once. Every "
                           "step carries the same payload into the same cache region, so step k+1's stores overlapping "
                           "step k's download are benign here"},
               "api": prefill_api + "; decode side: kvdp_stream_wait_flag; PCIe-bound per GPU whose link is used"}
        if cs.value:
            L.copy_stream_destroy(cs)
        if region is not None:
            dist.barrier()
            region.close()
        else:
            for p in host_ptrs:
                L.cuda_free_host(C.c_void_p(p))
        if host_out.value:
            L.cuda_free_host(host_out)

    # ---- token-granular transfer over NVLink: the same memory viewed per token slot, 2 KiB rows, all
    # layers and K|V in one launch; the decode cache region is treated as 32 layer slabs of [2, n*64 slots, 2 KiB].
    token = None
    if not args.no_comparators:
        from dynamo_exp_b200.transfer import token_rows_plan
        row, slots, ntok = 2048, n * 64, 8 * 4096
        layer_bytes = 2 * n * CHUNK
        if is_prefill:
            t_src, t_dst = [p.value for p in own_layers], [peer_cache.value + l * layer_bytes for l in range(L8B)]
        else:
            t_src, t_dst = peer_layers, [own_cache.value + l * layer_bytes for l in range(L8B)]
        t_plan = token_rows_plan(t_src, t_dst, slots, slots, row, row, row, device=local_rank)
        rng_t = np.random.Generator(np.random.PCG64(99 + partner if is_prefill else 99 + rank))
        ts_ids = rng_t.permutation(slots)[:ntok].astype(np.int32)
        td_ids = rng_t.permutation(slots)[:ntok].astype(np.int32)

        def token_timed(mode, count):
            runner = is_prefill if mode == "push" else not is_prefill
            torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            st = setup.stream if runner else wait_stream
            a.record(st)
            if runner:
                for _ in range(count):
                    seq["v"] += 1
                    lib.check(L.kvdp_xfer_run(setup.xfer, t_plan.handle, ts_ids.ctypes.data, td_ids.ctypes.data, ntok, 0,
                                              L8B, 0, peer_flag, seq["v"]))
            else:
                seq["v"] += count
                lib.check(L.kvdp_stream_wait_flag(C.c_void_p(wait_stream.cuda_stream), flag_ptr, seq["v"]))
            b.record(st)
            torch.cuda.synchronize()
            dist.barrier()
            t = torch.tensor([a.elapsed_time(b) / 1e3], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        def token_verify(mode):
            """Every transferred row of every layer, K and V, against the regenerated source."""
            bad = torch.zeros(1, dtype=torch.int64, device="cuda")
            if not is_prefill:
                rng_v = np.random.Generator(np.random.PCG64(99 + rank))
                s_t = torch.from_numpy(rng_v.permutation(slots)[:ntok].astype(np.int64)).cuda()
                d_t = torch.from_numpy(rng_v.permutation(slots)[:ntok].astype(np.int64)).cuda()
                cache = torch.as_tensor(CudaArray(own_cache.value, n * BLOCK_BYTES), device="cuda").view(torch.int64)
                cache = cache.view(L8B, 2, slots, row // 8)
                for l in range(L8B):
                    src = synth_words(torch, layer_seed(partner, l), layer_bytes).view(2, slots, row // 8)
                    bad += (src[:, s_t] != cache[l][:, d_t]).sum()
            dist.all_reduce(bad, op=dist.ReduceOp.SUM)
            if int(bad.item()) != 0:
                raise RuntimeError("token-granular %s: %d words differ from their source" % (mode, int(bad.item())))

        token = {"tokens_per_launch": ntok, "run_bytes": row, "runs_per_launch": ntok * 2 * L8B}
        k_t = max(2, min(steps, 10))
        for mode in modes:
            if not is_prefill:
                lib.check(L.kvdp_memset(own_cache, 0, n * BLOCK_BYTES))
            token_timed(mode, 2)
            sec_t = token_timed(mode, k_t)
            token_verify(mode)
            per_dir = ntok * row * 2 * L8B * k_t / sec_t / GB * max(1, P // D)
            token[mode] = {"per_direction_GBps": per_dir, "frac_of_nominal_900": per_dir / NVLINK_PEAK_NOMINAL,
                           "tokens_per_s": P * ntok * k_t / sec_t}
        t_plan.close()

    # ---- comparators: (i) local relayout into HBM staging + NCCL send/recv, (ii) plain peer memcpy of the payload
    nccl_value = memcpy_value = None
    nsteps_c = max(2, min(steps, 10))
    if not args.no_comparators:
        staging = torch.empty(pairs * BLOCK_BYTES, dtype=torch.uint8, device="cuda")
        iota = np.arange(pairs, dtype=np.int32)
        nccl_plan = C.c_void_p()
        if is_prefill:
            ptrs = np.array([p.value for p in own_layers], dtype=np.uint64)
            lib.check(L.kvdp_plan_contig(C.byref(nccl_plan), ptrs.ctypes.data, L8B, 0, n, CHUNK,
                                         C.c_void_p(staging.data_ptr()), pairs, 1, local_rank))

        def nccl_steps():
            cur = torch.cuda.current_stream()
            for k in range(nsteps_c):
                if is_prefill:
                    s_ids, _ = setup.ids(k)
                    lib.check(L.kvdp_xfer_run(setup.xfer, nccl_plan, s_ids.ctypes.data, iota.ctypes.data, pairs, 0,
                                              L8B, 0, None, 0))
                    ev = torch.cuda.Event()
                    ev.record(setup.stream)
                    cur.wait_event(ev)
                    dist.send(staging, partner)
                    ev2 = torch.cuda.Event()
                    ev2.record(cur)
                    setup.stream.wait_event(ev2)
                else:
                    dist.recv(staging, partner)

        def memcpy_steps():
            if is_prefill:
                for _ in range(nsteps_c):
                    lib.check(L.cuda_memcpy_async(peer_cache, C.c_void_p(staging.data_ptr()), pairs * BLOCK_BYTES,
                                                  C.c_void_p(torch.cuda.current_stream().cuda_stream)))

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

        nccl_steps()  # warm-up (NCCL connection setup)
        nccl_value = P * pairs * BLOCK_BYTES * nsteps_c / timed_default_stream(nccl_steps) / GB
        # what NCCL delivered in its last step is byte-equal to the source relayout, hence to what the push delivers
        # (SURVEY D.2 "NCCL path byte-equal to push path"): blocks s_ids of the prefill partner, in list order
        bad_n = torch.zeros(1, dtype=torch.int64, device="cuda")
        if not is_prefill:
            sperm_p = np.random.Generator(np.random.PCG64(1234 + partner)).permutation(n).astype(np.int32)
            o_n = ((nsteps_c - 1) % max(1, n // pairs)) * pairs
            s_n = torch.from_numpy(sperm_p[o_n:o_n + pairs].astype(np.int64)).cuda()
            got_n = staging.view(torch.int64).view(pairs, L8B, 2, CHUNK // 8)
            for l in range(L8B):
                src = synth_words(torch, layer_seed(partner, l), 2 * n * CHUNK).view(2, n, CHUNK // 8)
                bad_n += (src[:, s_n].transpose(0, 1) != got_n[:, l]).sum()
        dist.all_reduce(bad_n, op=dist.ReduceOp.SUM)
        if int(bad_n.item()) != 0:
            raise RuntimeError("NCCL comparator: %d words differ from the source relayout" % int(bad_n.item()))
        memcpy_steps()
        memcpy_value = P * pairs * BLOCK_BYTES * nsteps_c / timed_default_stream(memcpy_steps) / GB

    # ---- BASELINE config 3: Llama-3-70B-shaped TP-P -> TP-D head resharding through the connector (bench_reshard.py)
    reshard = []
    if not getattr(args, "no_reshard", False):
        from bench_reshard import configs_for, reshard_leg
        for p_tp, d_tp in configs_for(world):
            res = reshard_leg(torch, dist, lib, rank, local_rank, world, p_tp, d_tp, steps, warmup,
                              clock_sampler=ClockSampler(local_rank))
            if rank == 0:
                reshard.append(res)

    if rank == 0:
        best = max(results, key=lambda m: results[m]["value"])
        r = results[best]
        cfg = workload_config(args)
        topo = {"pairs": ["%d->%d" % (a, b) for a, b in role_pairs], "pairing": args.pairing, "mode": best}
        # the link's practical ceiling is what the copy engine reaches on the same payload in this very run
        peak = memcpy_value / P * max(1, P // D) if memcpy_value else NVLINK_PEAK_FALLBACK
        peak_src = ("cudaMemcpyAsync to the peer measured in this run, per direction per GPU" if memcpy_value else
                    "fallback: peer copy per direction per GPU (B200_PROFILING.md)")
        launches = steps * len(modes) + sum(x["steps"] * 2 for x in reshard)
        line = {
            "metric": "kv_transfer_GBps", "value": r["value"], "unit": "GB/s", "n_gpus": world, "steps": steps,
            "warmup": warmup, "ms_per_step": r["sec"] / steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": cfg, "topology": topo,
            "blocks_per_s": P * pairs * steps / r["sec"],
            "modes": {m: {"value": v["value"], "per_direction_GBps": v["per_direction"],
                          "frac_of_nominal_900": v["per_direction"] / NVLINK_PEAK_NOMINAL,
                          "frac_of_peer_memcpy": v["per_direction"] / peak,
                          "verified_bytes_last_step": v["verified_bytes_last_step"]}
                      for m, v in results.items()},
            "verification": "all bytes: each decode GPU regenerates its partner's 32 layers and compares every block of "
                            "the last timed step (and every row of the token-granular leg)",
            "roofline": {"bound": "nvlink", "achieved": r["per_direction"], "peak": peak,
                         "unit": "GB/s", "frac": r["per_direction"] / peak, "traffic": None,
                         "kernel": "kvdp_copy_tma_tma_kernel over peer-mapped memory (%s)" % best,
                         "peak_source": peak_src + "; nominal 900",
                         "frac_of_nominal_900": r["per_direction"] / NVLINK_PEAK_NOMINAL,
                         "note": "achieved = payload bytes per direction on the busiest port / time"},
            "comparators": {"nccl_send_recv_GBps": nccl_value, "cuda_memcpy_peer_GBps": memcpy_value,
                            "steps": nsteps_c,
                            "nccl_payload_check": None if nccl_value is None else
                            "all bytes of the last NCCL step equal the source relayout (and so the push path's result)"},
            "options": {k: lib.get_option(k) for k in ("copy_path", "tile_bytes", "stages", "ctas_per_sm")},
            "gpu_launches": launches, "clocks": clk, "env": env_info(torch, local_rank),
            "e2e": e2e, "token_transfer": token, "reshard": reshard,
        }
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()
