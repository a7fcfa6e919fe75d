"""GPU, two or more devices: the cross-GPU paths over REAL peers (NVLink on the B200 boxes) instead of ranks emulated as
slabs of one GPU. Skipped on a one-GPU box (the driver's GPU-test box has one); run with `gpurun --gpus 2`, log kept
under profiles/. One process, two devices, peer access enabled through kvdp_peer_enable.

  * every byte of a TP-resharding write_blocks / read_blocks between a prefill engine on GPU 0 and a decode engine on GPU 1
    against the canonical head slice;
  * SURVEY D.2 "completion flag observed only after all bytes are visible": the receiving GPU's stream waits on the flag
    the sending kernel's last CTA publishes and copies the data out at once, many times over, fresh data every round;
  * notifications across GPUs in both topologies the advisor named (2 senders -> 1 receiver, 1 sender -> 2 receivers).
"""
import ctypes as C

import numpy as np
import pytest

from dynamo_exp_b200 import CopyPlan, TransferEngine
from dynamo_exp_b200.connector import DynamoKvdpConnector, tp_targets
from tests.util import DevBuf, PinnedBuf, payload

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def two_gpus(gpu):
    n = C.c_int(0)
    gpu.check(gpu.lib().kvdp_device_count(C.byref(n)))
    if n.value < 2:
        pytest.skip("needs two GPUs (run under gpurun --gpus 2)")
    gpu.check(gpu.lib().kvdp_peer_enable(0, 1))
    gpu.check(gpu.lib().kvdp_peer_enable(1, 0))
    return gpu


L, H, D, BS, E = 4, 8, 128, 64, 2          # Llama-3-70B row geometry, 4 layers


def _engine(gpu, engine_id, tp, device, num_blocks, canon=None):
    h = H // tp
    conns, bufs = [], []
    for r in range(tp):
        layers = []
        for l in range(L):
            b = DevBuf(gpu, 2 * num_blocks * BS * h * D * E, device)
            if canon is not None:
                b.upload(np.ascontiguousarray(canon[l][:, :, :, r * h:(r + 1) * h, :]))
            layers.append(b)
        c = DynamoKvdpConnector(engine_id, r, tp, device=device)
        c.register_kv_caches([b.ptr for b in layers], shape=(2, num_blocks, BS, h, D), element_size=E)
        conns.append(c)
        bufs.append(layers)
    return conns, bufs


@pytest.mark.parametrize("ptp,dtp", [(4, 2), (2, 2), (1, 4), (2, 1)])
def test_reshard_between_two_gpus_every_byte(two_gpus, ptp, dtp):
    gpu = two_gpus
    NP, ND = 24, 20
    rng = np.random.Generator(np.random.PCG64(ptp * 16 + dtp))
    canon = rng.integers(-2 ** 15, 2 ** 15, size=(L, 2, NP, BS, H, D), dtype=np.int64).astype(np.int16)
    pre, pre_bufs = _engine(gpu, "prefill", ptp, 0, NP, canon)
    dec, dec_bufs = _engine(gpu, "decode", dtp, 1, ND)
    dec_meta = [c.get_agent_metadata() for c in dec]
    pre_meta = [c.get_agent_metadata() for c in pre]
    sid = [int(x) for x in rng.permutation(NP)[:13]]
    did = [int(x) for x in rng.permutation(ND)[:13]]
    for c in pre:
        c.add_remote_agent("decode", dec_meta, dtp, None, ND)
        c.write_blocks(sid, sid, did, "decode", "x%d" % c.rank)
    for c in pre:
        c.sync()
    hq = H // dtp
    for q in range(dtp):
        for l in range(L):
            got = dec_bufs[q][l].download(np.int16).reshape(2, ND, BS, hq, D)
            want = np.zeros_like(got)
            want[:, did] = canon[l][:, sid][:, :, :, q * hq:(q + 1) * hq, :]
            assert np.array_equal(got, want), ("push", q, l)
        senders = sorted(p for p in range(ptp) if q in tp_targets(ptp, dtp, p))
        assert dec[q].get_new_notifs() == {"prefill:%d" % p: ["x%d" % p] for p in senders}
    # pull the same blocks back into fresh prefill-side blocks (kernel on GPU 0 loading from GPU 1)
    back = [b for b in range(NP) if b not in sid]          # 11 blocks the push did not read: the sources stay intact
    hp = H // ptp
    for c in pre:
        c.read_blocks(back, back, did[:len(back)], "decode")
        for l in range(L):
            got = pre_bufs[c.rank][l].download(np.int16).reshape(2, NP, BS, hp, D)
            assert np.array_equal(got[:, back],
                                  canon[l][:, sid[:len(back)]][:, :, :, c.rank * hp:(c.rank + 1) * hp, :]), ("pull", l)
    # and the decode side pulling from the prefill GPU (kernel on GPU 1 loading from GPU 0), into other blocks
    did2 = [int(x) for x in rng.permutation(ND)[:13]]
    for c in dec:
        c.add_remote_agent("prefill", pre_meta, ptp, None, NP)
        c.read_blocks(did2, did2, sid, "prefill")
        for l in range(L):
            got = dec_bufs[c.rank][l].download(np.int16).reshape(2, ND, BS, hq, D)
            assert np.array_equal(got[:, did2], canon[l][:, sid][:, :, :, c.rank * hq:(c.rank + 1) * hq, :]), l
    for c in pre + dec:
        c.shutdown()


@pytest.mark.parametrize("path", [0, 1, 2, 3])
def test_flag_is_seen_only_after_all_bytes_across_two_gpus(two_gpus, path):
    """Sender: GPU 0 kernel storing into GPU 1 and publishing a flag there. Receiver: a GPU-1 stream that waits on the
    flag and immediately DMAs the destination to the host. 40 rounds, new payload each round, no host synchronisation
    between the two sides inside a round."""
    gpu = two_gpus
    Lr = gpu.lib()
    gpu.set_option("copy_path", path)
    try:
        layers, nb, chunk = 8, 96, 128 * 1024          # 8 layers x K|V x 96 blocks x 128 KiB = 192 MiB per round
        src = [DevBuf(gpu, 2 * nb * chunk, 0) for _ in range(layers)]
        dst = [DevBuf(gpu, 2 * nb * chunk, 1) for _ in range(layers)]
        flag = DevBuf(gpu, 256, 1)
        plan = CopyPlan.block_copy_raw([b.ptr for b in src], [b.ptr for b in dst], 2, nb, nb, chunk, 1, 0)
        sender, receiver = TransferEngine(nb, 0), TransferEngine(nb, 1)
        out = PinnedBuf(gpu, 2 * nb * chunk)
        ids = np.arange(nb, dtype=np.int32)
        probe_layer = layers - 1                          # issued last by the sender
        for rnd in range(1, 41):
            data = payload(2 * nb * chunk, 7000 + rnd)
            for b in src:
                b.upload(data)
            perm = np.random.Generator(np.random.PCG64(rnd)).permutation(nb).astype(np.int32)
            # receiver first: its stream blocks on the flag, then copies the last layer out
            gpu.check(Lr.kvdp_stream_wait_flag(C.c_void_p(receiver.stream), C.c_void_p(flag.ptr), rnd))
            gpu.check(Lr.cuda_memcpy_async(C.c_void_p(out.ptr), C.c_void_p(dst[probe_layer].ptr), 2 * nb * chunk,
                                           C.c_void_p(receiver.stream)))
            sender.run(plan, ids, perm, num_layers=layers, done_flag=flag.ptr, done_value=rnd)
            receiver.sync()
            want = np.empty_like(data).reshape(2, nb, chunk)
            want[:, perm] = data.reshape(2, nb, chunk)[:, ids]
            assert np.array_equal(out.array.reshape(2, nb, chunk), want), rnd
            sender.sync()
    finally:
        gpu.set_option("copy_path", 0)


def test_notifications_across_gpus_two_senders_and_two_receivers(two_gpus):
    gpu = two_gpus
    NP = ND = 8
    pa, _a = _engine(gpu, "prefill-a", 1, 0, NP)
    pb, _b = _engine(gpu, "prefill-b", 1, 0, NP)
    d1, _1 = _engine(gpu, "decode-1", 1, 1, ND)
    d2, _2 = _engine(gpu, "decode-2", 1, 1, ND)
    for c in pa + pb:
        c.add_remote_agent("decode-1", [d1[0].get_agent_metadata()], 1, None, ND)
        c.add_remote_agent("decode-2", [d2[0].get_agent_metadata()], 1, None, ND)
    want = {("decode-1", "prefill-a:0"): [], ("decode-1", "prefill-b:0"): [], ("decode-2", "prefill-a:0"): [],
            ("decode-2", "prefill-b:0"): []}
    for k in range(90):
        snd = pa[0] if k % 2 == 0 else pb[0]
        eng = "decode-1" if (k // 2) % 2 == 0 else "decode-2"
        snd.write_blocks([k % NP], [0], [k % ND], eng, "m%d" % k)
        want[(eng, snd.agent_name)].append("m%d" % k)
    pa[0].sync()
    pb[0].sync()
    for name, c in (("decode-1", d1[0]), ("decode-2", d2[0])):
        assert c.get_new_notifs() == {s: want[(name, s)] for s in ("prefill-a:0", "prefill-b:0")}
    for c in pa + pb + d1 + d2:
        c.shutdown()


def test_onboard_from_a_shared_host_region_over_both_gpus_links(two_gpus, oracle):
    """bench.py --gpus N e2e, --e2e-links both: host blocks in a registered shared region reach the decode GPU's cache over
    two PCIe links at once - GPU 0's CopyStream carries request 0 (copy engines into its staging, kernel stores into GPU 1's
    memory over NVLink), GPU 1's own CopyStream fetches request 1. Every byte of the decode cache against the oracle."""
    import torch

    from dynamo_exp_b200.shared_host import SharedHostRegion, split_for_links
    gpu = two_gpus
    lib = gpu.lib()
    layers, nh, nd, suffix, elem, per_req = 3, 64, 48, 64 * 1024, 2, 16
    chunk = suffix * elem
    layer_bytes = 2 * nh * chunk
    region = SharedHostRegion.create(layers * layer_bytes)
    try:
        arr = region.array()
        arr[:] = payload(arr.size, 5150)
        src_np = [arr[l * layer_bytes:(l + 1) * layer_bytes].copy() for l in range(layers)]
        cache = [DevBuf(gpu, 2 * nd * chunk, 1) for _ in range(layers)]
        for c in cache:
            c.upload(np.zeros(c.nbytes, dtype=np.uint8))
        rng = np.random.Generator(np.random.PCG64(8))
        h_ids = np.concatenate([np.arange(8, 24), np.arange(40, 56)]).astype(np.int32)
        d_ids = rng.permutation(nd)[:2 * per_req].astype(np.int32)
        streams = []
        for link, (r0, r1) in enumerate(split_for_links(2, 2)):
            torch.cuda.set_device(link)
            cs = C.c_void_p()
            assert lib.copy_stream_create(C.byref(cs), layers, per_req) == 0
            streams.append(cs)
            h = h_ids[r0 * per_req:r1 * per_req].copy()
            d = d_ids[r0 * per_req:r1 * per_req].copy()
            assert lib.copy_stream_prepare_block_ids(cs, h.ctypes.data, d.ctypes.data, len(h)) == 0
            for l in range(layers):
                assert lib.copy_stream_memcpy(cs, C.c_void_p(region.ptr + l * layer_bytes), cache[l].ptr, 2, suffix, elem,
                                              nh, nd) == 0, gpu.last_error()
        for cs in streams:
            assert lib.copy_stream_sync(cs) == 0
            assert lib.copy_stream_destroy(cs) == 0
        torch.cuda.set_device(0)
        want = [np.zeros(2 * nd * chunk, dtype=np.uint8) for _ in range(layers)]
        oracle.copy_blocks_layers(src_np, want, h_ids, d_ids, 2, nh, nd, suffix, elem)
        for l in range(layers):
            assert np.array_equal(cache[l].download(), want[l]), l
        del arr
    finally:
        torch.cuda.set_device(0)
        region.close()
