"""Round-2 tuning sweeps on one B200 (printed as plain tables, kept under profiles/):
  hash   : one chained kvdp_hash_token_blocks call on the router-sweep batch for every `hash_stages` setting
  host   : legacy copy_stream_memcpy between pinned host and HBM, both directions, host_path x host_ctas x tile_bytes
  hosthash : kvdp_hash_token_blocks_host on pinned buffers (pipelined pieces) against the link
Usage: python tools/r02_sweeps.py hash host hosthash
"""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamo_exp_b200 import _lib as lib  # noqa: E402

L = lib.lib()
GB = 1e9
BS = 64


def hash_sweep():
    isls = [512, 1024, 2048, 4096, 8192, 16384, 32768]
    n_req = 7 * 8192
    lens = np.array([isls[i % 7] for i in range(n_req)], dtype=np.int64)
    offs = np.zeros(n_req + 1, dtype=np.int64)
    offs[1:] = np.cumsum(lens)
    total, nblk = int(offs[-1]), int((lens // BS).sum())
    tok = torch.randint(0, 128256, (total,), dtype=torch.int32, device="cuda")
    d_off = torch.from_numpy(offs).cuda()
    d_boff = torch.empty(n_req + 1, dtype=torch.int64, device="cuda")
    d_bh = torch.empty(nblk, dtype=torch.int64, device="cuda")
    d_sh = torch.empty(nblk, dtype=torch.int64, device="cuda")
    ws = L.kvdp_hash_workspace_bytes(n_req)
    d_ws = torch.empty(ws, dtype=torch.uint8, device="cuda")
    st = torch.cuda.Stream()
    ref = None
    print("hash_stages  ms_per_batch  G_hashes/s  frac_of_6445(272 B/block)  same_bits")
    for stages in (1, 2, 3, 4, 5, 6, 8, 12, 16, 0):
        lib.set_option("hash_stages", stages)

        def call(chained=True):
            lib.check(L.kvdp_hash_token_blocks(tok.data_ptr(), d_off.data_ptr(), n_req, total, BS, 1337,
                                               d_boff.data_ptr(), d_bh.data_ptr(), d_sh.data_ptr() if chained else None,
                                               d_ws.data_ptr(), ws, 0, C.c_void_p(st.cuda_stream)))
        for _ in range(3):
            call()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        for _ in range(20):
            call()
        b.record(st)
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 20
        cur = (d_bh.clone(), d_sh.clone())
        ref = ref or cur
        same = torch.equal(cur[0], ref[0]) and torch.equal(cur[1], ref[1])
        print("%11d  %12.4f  %10.2f  %10.3f  %s" % (stages, ms, nblk / ms / 1e6, nblk * 272 / (ms / 1e3) / GB / 6445, same))
    lib.set_option("hash_stages", 0)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for _ in range(20):
        lib.check(L.kvdp_hash_token_blocks(tok.data_ptr(), d_off.data_ptr(), n_req, total, BS, 1337, d_boff.data_ptr(),
                                           d_bh.data_ptr(), None, d_ws.data_ptr(), ws, 0, C.c_void_p(st.cuda_stream)))
    b.record(st)
    torch.cuda.synchronize()
    print("un-chained: %.4f ms" % (a.elapsed_time(b) / 20))


def host_sweep():
    layers, nb, chunk = 8, 512, 128 * 1024
    layer_bytes = 2 * nb * chunk
    hp = []
    for _ in range(layers):
        p = C.c_void_p()
        lib.check(L.cuda_malloc_host(C.byref(p), layer_bytes))
        C.memset(p.value, 1, layer_bytes)
        hp.append(p.value)
    dev = [torch.zeros(layer_bytes, dtype=torch.uint8, device="cuda") for _ in range(layers)]
    cs = C.c_void_p()
    assert L.copy_stream_create(C.byref(cs), layers, nb) == 0
    rng = np.random.Generator(np.random.PCG64(1))
    payload = layers * layer_bytes
    # ceilings
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    for name, dst, src in (("h2d", dev[0].data_ptr(), hp[0]), ("d2h", hp[0], dev[0].data_ptr())):
        best = 0
        for _ in range(3):
            ev[0].record()
            lib.check(L.cuda_memcpy_async(dst, src, layer_bytes, None))
            ev[1].record()
            torch.cuda.synchronize()
            best = max(best, layer_bytes / (ev[0].elapsed_time(ev[1]) / 1e3) / GB)
        print("ceiling %s: %.2f GB/s" % (name, best))
    print("dir  host_path  host_ctas  tile   GB/s")
    for direction in ("h2d", "d2h"):
        for host_path in (0, 2):
            for ctas in (16, 32, 48, 64, 96, 148):
                for tile in (16384, 32768, 65536):
                    if host_path == 2 and tile > 16384:
                        continue
                    lib.set_option("host_path", host_path)
                    lib.set_option("host_ctas", ctas)
                    lib.set_option("tile_bytes", tile)
                    best = 0
                    for rep in range(3):
                        sid, did = rng.permutation(nb).astype(np.int32), rng.permutation(nb).astype(np.int32)
                        torch.cuda.synchronize()
                        t0 = time.perf_counter()
                        lib.check(L.copy_stream_prepare_block_ids(cs, sid.ctypes.data, did.ctypes.data, nb))
                        for l in range(layers):
                            if direction == "h2d":
                                lib.check(L.copy_stream_memcpy(cs, hp[l], dev[l].data_ptr(), 2, chunk // 2, 2, nb, nb))
                            else:
                                lib.check(L.copy_stream_memcpy(cs, dev[l].data_ptr(), hp[l], 2, chunk // 2, 2, nb, nb))
                        lib.check(L.copy_stream_sync(cs))
                        dt = time.perf_counter() - t0
                        if rep:
                            best = max(best, payload / dt / GB)
                    print("%s  %9d  %9d  %5d  %6.2f" % (direction, host_path, ctas, tile, best), flush=True)
    for k in ("host_path", "host_ctas", "tile_bytes"):
        lib.set_option(k, 0)
    L.copy_stream_destroy(cs)


def hosthash_sweep():
    isls = [512, 1024, 2048, 4096, 8192, 16384, 32768]
    for n_e in (7 * 256, 7 * 1024):
        lens = np.array([isls[i % 7] for i in range(n_e)], dtype=np.int64)
        offs = np.zeros(n_e + 1, dtype=np.int64)
        offs[1:] = np.cumsum(lens)
        tot, nb = int(offs[-1]), int((lens // BS).sum())
        h_tok = torch.randint(0, 128256, (tot,), dtype=torch.int32).pin_memory()
        h_boff = torch.empty(n_e + 1, dtype=torch.int64).pin_memory()
        h_bh = torch.empty(nb, dtype=torch.int64).pin_memory()
        h_sh = torch.empty(nb, dtype=torch.int64).pin_memory()

        def run():
            lib.check(L.kvdp_hash_token_blocks_host(h_tok.data_ptr(), offs.ctypes.data, n_e, BS, 1337, h_boff.data_ptr(),
                                                    h_bh.data_ptr(), h_sh.data_ptr(), 0, -1))
        run()
        t0 = time.perf_counter()
        for _ in range(5):
            run()
        dt = (time.perf_counter() - t0) / 5
        print("host hash: %d requests, %.1f MB up, %.1f MB down: %.3f ms, %.1f M hashes/s, %.1f GB/s up" %
              (n_e, tot * 4 / 1e6, nb * 16 / 1e6, dt * 1e3, nb / dt / 1e6, tot * 4 / dt / GB))


if __name__ == "__main__":
    for what in sys.argv[1:]:
        {"hash": hash_sweep, "host": host_sweep, "hosthash": hosthash_sweep}[what]()
