"""GPU: the transfer connector (mirror of the patched vLLM DynamoNixlConnector, patch:1081-1492) — write_blocks (push),
read_blocks (pull), TP resharding in both directions, notifications, MLA replication. All engine ranks live on ONE GPU in
this process (ranks are separate cache slabs), so peers are reached through raw pointers; the cross-process CUDA-IPC
path is covered by test_two_process_ipc_transfer below. Oracle: the formula-independent canonical-slice property
 — every decode rank's cache equals the head slice of a TP1 cache at the mapped block ids."""
import os
import subprocess
import sys

import numpy as np
import pytest

from dynamo_exp_b200.connector import DynamoKvdpConnector, KvdpMetadata, tp_targets
from dynamo_exp_b200._lib import KvdpError
from dynamo_exp_b200.storage import KvError
from tests.util import DevBuf

pytestmark = pytest.mark.gpu

L, H, D, BS, E = 3, 8, 16, 8, 2
NP, ND = 9, 11


def make_engine(gpu, engine_id, tp, num_blocks, canon=None):
    """One connector + cache slabs per TP rank. If `canon` ([L,2,N,BS,H,D]) is given the caches hold its head slices."""
    h = H // tp
    conns, bufs = [], []
    for r in range(tp):
        layers = []
        for l in range(L):
            b = DevBuf(gpu, 2 * num_blocks * BS * h * D * E)
            if canon is not None:
                b.upload(np.ascontiguousarray(canon[l][:, :, :, r * h:(r + 1) * h, :]))
            layers.append(b)
        c = DynamoKvdpConnector(engine_id, r, tp, device=0)
        c.register_kv_caches([b.ptr for b in layers], shape=(2, num_blocks, BS, h, D), element_size=E)
        conns.append(c)
        bufs.append(layers)
    meta = KvdpMetadata(engine_id, [c.get_agent_metadata() for c in conns],
                        [c.kv_caches_base_addr[engine_id] for c in conns], num_blocks)
    return conns, bufs, KvdpMetadata.decode(meta.encode())     # through the wire format, like etcd would


def read_rank(bufs, r, tp, num_blocks):
    h = H // tp
    return [b.download(np.int16).reshape(2, num_blocks, BS, h, D) for b in bufs[r]]


@pytest.mark.parametrize("ptp,dtp", [(1, 1), (1, 2), (2, 4), (1, 4), (4, 2), (2, 1), (4, 1), (2, 2)])
def test_write_blocks_reshards_and_notifies(gpu, ptp, dtp):
    rng = np.random.Generator(np.random.PCG64(ptp * 10 + dtp))
    canon = rng.integers(0, 2 ** 15, size=(L, 2, NP, BS, H, D), dtype=np.int16)
    pre, pre_bufs, pre_meta = make_engine(gpu, "prefill", ptp, NP, canon)
    dec, dec_bufs, dec_meta = make_engine(gpu, "decode", dtp, ND)
    local, remote = [7, 0, 3, 5], [2, 10, 0, 6, 9]          # one extra remote id: truncated like patch:1368
    for c in pre:
        names = c.add_remote_agent("decode", dec_meta.agent_metadata, dtp, dec_meta.kv_caches_base_addr,
                                   dec_meta.num_blocks)
        assert names == [f"decode:{q}" for q in tp_targets(ptp, dtp, c.rank)]
        c.write_blocks(local, [0] * len(local), remote, "decode", f"req-42-from-{c.rank}")
    for c in pre:
        c.sync()
        assert c.get_done_tranfers() == [f"req-42-from-{c.rank}"]
        assert c.get_done_tranfers() == []
    hq = H // dtp
    for q in range(dtp):
        got = read_rank(dec_bufs, q, dtp, ND)
        for l in range(L):
            want = canon[l][:, local][:, :, :, q * hq:(q + 1) * hq, :]
            assert np.array_equal(got[l][:, remote[:4]], want), (q, l)
        senders = sorted(p for p in range(ptp) if q in tp_targets(ptp, dtp, p))
        notifs = dec[q].get_new_notifs()
        assert notifs == {f"prefill:{p}": [f"req-42-from-{p}"] for p in senders}
        assert dec[q].get_new_notifs() == {}
    # empty transfer: notify only (patch:1373-1377)
    pre[0].write_blocks([], [], [1, 2], "decode", "empty")
    pre[0].sync()
    for q in tp_targets(ptp, dtp, 0):
        assert dec[q].get_new_notifs() == {"prefill:0": ["empty"]}
    for c in pre + dec:
        c.shutdown()


@pytest.mark.parametrize("ptp,dtp", [(1, 1), (1, 2), (2, 4), (1, 4), (4, 2), (2, 1), (4, 1), (2, 2)])
def test_write_token_rows_reshards_per_token(gpu, ptp, dtp):
    """The NCCL connector's KV path (patch:1602-1766) as one launch: decode.reshape(2, slots, Hq, D)[kv, dst_slot] =
    prefill.reshape(2, slots, Hp, D)[kv, src_slot, head slice] for every layer; untouched slots keep their content."""
    rng = np.random.Generator(np.random.PCG64(ptp * 31 + dtp))
    canon = rng.integers(0, 2 ** 15, size=(L, 2, NP, BS, H, D), dtype=np.int16)
    pre, pre_bufs, _ = make_engine(gpu, "prefill", ptp, NP, canon)
    dec, dec_bufs, dec_meta = make_engine(gpu, "decode", dtp, ND)
    before = [read_rank(dec_bufs, q, dtp, ND) for q in range(dtp)]
    src_slots = [int(x) for x in rng.permutation(NP * BS)[:37]]
    dst_slots = [int(x) for x in rng.permutation(ND * BS)[:37]]
    for c in pre:
        c.add_remote_agent("decode", dec_meta.agent_metadata, dtp, dec_meta.kv_caches_base_addr, dec_meta.num_blocks)
        c.write_token_rows(src_slots, dst_slots, "decode", f"tok-{c.rank}")
    for c in pre:
        c.sync()
        assert c.get_done_tranfers() == [f"tok-{c.rank}"]
    hq = H // dtp
    flat_canon = canon.reshape(L, 2, NP * BS, H, D)
    for q in range(dtp):
        got = read_rank(dec_bufs, q, dtp, ND)
        for l in range(L):
            want = before[q][l].reshape(2, ND * BS, hq, D).copy()
            want[:, dst_slots] = flat_canon[l][:, src_slots][:, :, q * hq:(q + 1) * hq, :]
            assert np.array_equal(got[l].reshape(2, ND * BS, hq, D), want), (q, l)
        senders = sorted(p for p in range(ptp) if q in tp_targets(ptp, dtp, p))
        assert dec[q].get_new_notifs() == {f"prefill:{p}": [f"tok-{p}"] for p in senders}
    with pytest.raises(KvError):
        pre[0].write_token_rows([0, 1], [0], "decode", "bad")
    for c in pre + dec:
        c.shutdown()


@pytest.mark.parametrize("ptp,dtp", [(1, 1), (1, 2), (2, 4), (4, 2), (2, 1), (4, 1)])
def test_read_blocks_pulls_and_inverts_the_reshard(gpu, ptp, dtp):
    """read_blocks (patch:1296-1359): the prefill side pulls prefix-hit blocks out of the decode caches. Any TP ratio,
    including decode TP < prefill TP (strided source rows), which the reference cannot do."""
    rng = np.random.Generator(np.random.PCG64(ptp * 7 + dtp))
    canon = rng.integers(0, 2 ** 15, size=(L, 2, ND, BS, H, D), dtype=np.int16)
    dec, dec_bufs, dec_meta = make_engine(gpu, "decode", dtp, ND, canon)
    pre, pre_bufs, pre_meta = make_engine(gpu, "prefill", ptp, NP)
    remote, local = [10, 1, 4], [0, 8, 3]
    hp = H // ptp
    for c in pre:
        c.add_remote_agent("decode", dec_meta.agent_metadata, dtp, dec_meta.kv_caches_base_addr, dec_meta.num_blocks)
        c.read_blocks(local, local, remote, "decode")        # blocking
        got = read_rank(pre_bufs, c.rank, ptp, NP)
        for l in range(L):
            want = canon[l][:, remote][:, :, :, c.rank * hp:(c.rank + 1) * hp, :]
            assert np.array_equal(got[l][:, local], want), (c.rank, l)
    with pytest.raises(KvError):
        pre[0].read_blocks([0], [0], [1, 2], "decode")
    with pytest.raises(KvError, match="unknown remote engine"):
        pre[0].write_blocks([0], [0], [0], "nobody", "x")
    for c in pre + dec:
        c.shutdown()


def test_mla_latent_cache_is_replicated(gpu):
    """MLA branch (patch:1138-1160, 1440-1451): one [num_blocks, block_size, head_dim] entry per layer, no head split;
    every target rank receives the full block."""
    nb, bs, hd = 6, 4, 24
    rng = np.random.Generator(np.random.PCG64(5))
    src_h = [rng.integers(0, 255, size=(nb, bs, hd), dtype=np.uint8) for _ in range(L)]
    src = [DevBuf(gpu, a.nbytes) for a in src_h]
    for b, a in zip(src, src_h):
        b.upload(a)
    pre = DynamoKvdpConnector("p", 0, 1, device=0, is_mla=True)
    pre.register_kv_caches([b.ptr for b in src], shape=(nb, bs, hd), element_size=1)
    decs, dbufs = [], []
    for r in range(2):
        bufs = [DevBuf(gpu, nb * bs * hd) for _ in range(L)]
        c = DynamoKvdpConnector("d", r, 2, device=0, is_mla=True)
        c.register_kv_caches([b.ptr for b in bufs], shape=(nb, bs, hd), element_size=1)
        decs.append(c)
        dbufs.append(bufs)
    pre.add_remote_agent("d", [c.get_agent_metadata() for c in decs], 2, None, nb)
    pre.write_blocks([5, 1], [0, 0], [0, 3], "d", "mla")
    pre.sync()
    for r in range(2):
        for l in range(L):
            got = dbufs[r][l].download().reshape(nb, bs, hd)
            assert np.array_equal(got[[0, 3]], src_h[l][[5, 1]])
        assert decs[r].get_new_notifs() == {"p:0": ["mla"]}
    # token-granular form on the latent cache (the NCCL connector's deepseek branch, patch:1662-1665, 1748-1752): whole
    # token rows, replicated to every target rank; untouched slots keep their content
    before = [[b.download().reshape(nb * bs, hd) for b in dbufs[r]] for r in range(2)]
    s_slots, d_slots = [21, 3, 8, 14], [1, 22, 9, 17]
    pre.write_token_rows(s_slots, d_slots, "d", "mla-tok")
    pre.sync()
    for r in range(2):
        for l in range(L):
            want = before[r][l].copy()
            want[d_slots] = src_h[l].reshape(nb * bs, hd)[s_slots]
            assert np.array_equal(dbufs[r][l].download().reshape(nb * bs, hd), want), (r, l)
        assert decs[r].get_new_notifs() == {"p:0": ["mla-tok"]}
    for c in [pre] + decs:
        c.shutdown()


def test_many_notifications_wrap_the_mailbox_ring(gpu):
    dec, dec_bufs, dec_meta = make_engine(gpu, "decode", 1, ND)      # keep the cache slabs alive
    pre, pre_bufs, _ = make_engine(gpu, "prefill", 1, NP)
    pre[0].add_remote_agent("decode", dec_meta.agent_metadata, 1, None, ND)
    sent = []
    for k in range(150):                      # ring depth is 64: drain in between like a polling decode worker
        pre[0].write_blocks([k % NP], [0], [k % ND], "decode", f"r{k}")
        sent.append(f"r{k}")
        if k % 40 == 39:
            pre[0].sync()
            got = dec[0].get_new_notifs().get("prefill:0", [])
            assert got == sent[:len(got)]
            sent = sent[len(got):]
    pre[0].sync()
    assert dec[0].get_new_notifs().get("prefill:0", []) == sent
    for c in pre + dec:
        c.shutdown()
    del dec_bufs, pre_bufs


def test_two_prefill_engines_with_equal_ranks_feed_one_decode_engine(gpu):
    """Two senders whose rank numbers coincide must not share a notification lane or a ring slot (NIXL keeps one queue
    per remote agent). Interleaved writes from both; every message arrives once, under its sender's agent name."""
    rng = np.random.Generator(np.random.PCG64(77))
    canon_a = rng.integers(0, 2 ** 15, size=(L, 2, NP, BS, H, D), dtype=np.int16)
    canon_b = rng.integers(0, 2 ** 15, size=(L, 2, NP, BS, H, D), dtype=np.int16)
    dec, dec_bufs, dec_meta = make_engine(gpu, "decode", 1, ND)
    pa, pa_bufs, _ = make_engine(gpu, "prefill-a", 1, NP, canon_a)
    pb, pb_bufs, _ = make_engine(gpu, "prefill-b", 1, NP, canon_b)
    for c in pa + pb:
        c.add_remote_agent("decode", dec_meta.agent_metadata, 1, None, ND)
    assert pa[0]._remotes["decode"].lane[0] != pb[0]._remotes["decode"].lane[0]
    sent_a, sent_b = [], []
    for k in range(30):
        pa[0].write_blocks([k % NP], [0], [k % 5], "decode", f"a{k}")
        sent_a.append(f"a{k}")
        if k % 3 == 0:
            pb[0].write_blocks([k % NP], [0], [5 + k % 5], "decode", f"b{k}")
            sent_b.append(f"b{k}")
    pa[0].sync()
    pb[0].sync()
    assert dec[0].get_new_notifs() == {"prefill-a:0": sent_a, "prefill-b:0": sent_b}
    got = read_rank(dec_bufs, 0, 1, ND)
    for l in range(L):
        assert np.array_equal(got[l][:, 4], canon_a[l][:, 29 % NP])        # last writer of decode block 4 was a29
        assert np.array_equal(got[l][:, 5 + 27 % 5], canon_b[l][:, 27 % NP])
    # the decode engine can tell where each sender's counter lives (what a stream-wait needs)
    assert dec[0].notif_head_ptr("prefill-a:0") != dec[0].notif_head_ptr("prefill-b:0") != 0
    assert dec[0].notif_head_ptr("nobody:0") == 0
    for c in pa + pb + dec:
        c.shutdown()
    del dec_bufs, pa_bufs, pb_bufs


def test_one_prefill_engine_feeds_two_decode_engines(gpu):
    """The sender's pinned staging ring is shared by all its peers: alternating targets must not let one peer's message
    land in the other's mailbox, also when the ring wraps (more than 64 messages in flight over the session)."""
    rng = np.random.Generator(np.random.PCG64(78))
    canon = rng.integers(0, 2 ** 15, size=(L, 2, NP, BS, H, D), dtype=np.int16)
    pre, pre_bufs, _ = make_engine(gpu, "prefill", 1, NP, canon)
    d1, d1_bufs, m1 = make_engine(gpu, "decode-1", 1, ND)
    d2, d2_bufs, m2 = make_engine(gpu, "decode-2", 1, ND)
    pre[0].add_remote_agent("decode-1", m1.agent_metadata, 1, None, ND)
    pre[0].add_remote_agent("decode-2", m2.agent_metadata, 1, None, ND)
    want1, want2 = [], []
    for k in range(100):
        eng, want = ("decode-1", want1) if k % 2 == 0 else ("decode-2", want2)
        pre[0].write_blocks([k % NP], [0], [k % ND], eng, f"{eng}/r{k}")
        want.append(f"{eng}/r{k}")
        if k == 59:                        # both rings hold 30 unread messages here; drain like a polling worker would
            pre[0].sync()
            assert d1[0].get_new_notifs() == {"prefill:0": want1}
            assert d2[0].get_new_notifs() == {"prefill:0": want2}
            want1, want2 = [], []
    pre[0].sync()
    assert d1[0].get_new_notifs() == {"prefill:0": want1}
    assert d2[0].get_new_notifs() == {"prefill:0": want2}
    assert sorted(pre[0].get_done_tranfers()) == sorted(f"decode-{1 + k % 2}/r{k}" for k in range(100))
    for c in pre + d1 + d2:
        c.shutdown()
    del pre_bufs, d1_bufs, d2_bufs


def test_sender_waits_for_the_receiver_instead_of_overwriting_unread_notifications(gpu):
    dec, dec_bufs, dec_meta = make_engine(gpu, "decode", 1, ND)
    pre, pre_bufs, _ = make_engine(gpu, "prefill", 1, NP)
    pre[0].add_remote_agent("decode", dec_meta.agent_metadata, 1, None, ND)
    pre[0].notify_timeout_s = 0.2
    for k in range(64):
        pre[0].write_blocks([], [], [], "decode", f"n{k}")
    with pytest.raises(KvError, match="unread notifications"):
        pre[0].write_blocks([], [], [], "decode", "one too many")
    assert dec[0].get_new_notifs() == {"prefill:0": [f"n{k}" for k in range(64)]}
    pre[0].write_blocks([], [], [], "decode", "fits again")
    pre[0].sync()
    assert dec[0].get_new_notifs() == {"prefill:0": ["fits again"]}
    for c in pre + dec:
        c.shutdown()
    del dec_bufs, pre_bufs


def test_two_process_ipc_transfer(gpu):
    """Two OS processes on the same GPU exchange metadata through a file and move blocks through CUDA-IPC mappings
    (the path the multi-GPU bench uses across GPUs). Host-side polling only: no kernel waits on another process."""
    script = os.path.join(os.path.dirname(__file__), "mp_connector_check.py")
    r = subprocess.run([sys.executable, script], capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "IPC OK" in r.stdout


@pytest.mark.parametrize("d", [1, 2, 4])
@pytest.mark.parametrize("shape,esz", [((3, 4, 8, 16), 2), ((5, 2, 4, 3), 4), ((2, 64, 8, 128), 2)])
def test_rearrange_tensors_matches_the_triton_index_map(gpu, oracle, d, shape, esz):
    """patch:934-1031; oracle = kvo_rearrange, itself pinned to reshape(N,B,d,H/d,C).transpose(2,0,1,3,4) on CPU."""
    from dynamo_exp_b200.connector import rearrange_tensors
    N, B, H, Cc = shape
    dt = {2: np.int16, 4: np.int32}[esz]
    rng = np.random.Generator(np.random.PCG64(d + N))
    t1 = rng.integers(0, 2 ** 15, size=shape).astype(dt)
    want = np.zeros_like(t1)
    assert oracle.rearrange(t1, want, d, "write") == 0
    d1, d2 = DevBuf(gpu, t1.nbytes), DevBuf(gpu, t1.nbytes)
    d1.upload(t1)
    rearrange_tensors(d1.ptr, d2.ptr, shape, esz, d, "write")
    got = d2.download(dt).reshape(shape)
    assert np.array_equal(got, want)
    assert np.array_equal(got.reshape(d, N, B, H // d, Cc), t1.reshape(N, B, d, H // d, Cc).transpose(2, 0, 1, 3, 4))
    back = DevBuf(gpu, t1.nbytes)
    rearrange_tensors(back.ptr, d2.ptr, shape, esz, d, "read")
    assert np.array_equal(back.download(dt).reshape(shape), t1)
    with pytest.raises(KvError):
        rearrange_tensors(d1.ptr, d2.ptr, shape, esz, 3, "write")
    with pytest.raises(KvError):
        rearrange_tensors(d1.ptr, d2.ptr, shape, esz, d, "sideways")


@pytest.mark.parametrize("H,D,E", [(8, 128, 2), (4, 6, 4), (2, 3, 1)])
def test_token_granular_paged_write_and_gather(gpu, H, D, E):
    """slot_mapping scatter (reshape_and_cache_flash, patch:1721-1746) and the per-layer gather with a
    head slice (patch:1602-1672), checked against numpy indexing."""
    from dynamo_exp_b200.transfer import gather_token_rows, reshape_and_cache
    rng = np.random.Generator(np.random.PCG64(H * D))
    num_blocks, bs, ntok = 6, 4, 13
    row = H * D * E
    key = rng.integers(0, 255, size=(ntok, row), dtype=np.uint8)
    val = rng.integers(0, 255, size=(ntok, row), dtype=np.uint8)
    slots = rng.permutation(num_blocks * bs)[:ntok].astype(np.int64)
    slots[3] = -1                                              # padding token: skipped
    dk, dv = DevBuf(gpu, key.nbytes), DevBuf(gpu, val.nbytes)
    ck, cv = DevBuf(gpu, num_blocks * bs * row), DevBuf(gpu, num_blocks * bs * row)
    dk.upload(key)
    dv.upload(val)
    reshape_and_cache(dk.ptr, dv.ptr, ck.ptr, cv.ptr, slots, num_blocks * bs, row)
    want_k = np.zeros((num_blocks * bs, row), dtype=np.uint8)
    want_v = np.zeros_like(want_k)
    ok = slots >= 0
    want_k[slots[ok]] = key[ok]
    want_v[slots[ok]] = val[ok]
    assert np.array_equal(ck.download().reshape(-1, row), want_k)
    assert np.array_equal(cv.download().reshape(-1, row), want_v)
    # gather heads [h0, h1) of the written tokens back out: cache.reshape(-1, H, D)[slots, h0:h1]
    h0, h1 = (H // 2, H) if H > 1 else (0, 1)
    sel = slots[ok]
    out = DevBuf(gpu, sel.size * (h1 - h0) * D * E)
    gather_token_rows(ck.ptr, out.ptr, sel, num_blocks * bs, row, h0 * D * E, (h1 - h0) * D * E)
    want = want_k.reshape(-1, H, D * E)[sel, h0:h1].reshape(sel.size, -1)
    assert np.array_equal(out.download().reshape(sel.size, -1), want)
    with pytest.raises(KvError):
        gather_token_rows(ck.ptr, out.ptr, [num_blocks * bs], num_blocks * bs, row)


@pytest.mark.parametrize("L,Hs,Hd,D,E,h_src,h_dst,nh", [
    (3, 8, 8, 16, 2, 0, 0, 8),      # same TP degree: whole rows
    (2, 8, 2, 16, 2, 4, 0, 2),      # prefill TP1 -> decode TP4, decode rank 2 takes heads [4, 6)
    (2, 2, 8, 8, 2, 0, 6, 2),       # prefill TP4 rank 3 -> decode TP1: its 2 heads land at heads [6, 8)
    (1, 3, 5, 5, 1, 1, 2, 2),       # odd sizes: byte-granular runs, unaligned offsets
])
def test_token_granular_transfer_all_layers_one_launch(gpu, L, Hs, Hd, D, E, h_src, h_dst, nh):
    """The fused form of the reference's NCCL connector data path (patch:1602-1672, 1676-1766): for every layer, K
    and V, `dst.reshape(2, -1, Hd, D)[kv, dst_slots, h_dst:h_dst+nh] = src.reshape(2, -1, Hs, D)[kv, src_slots,
    h_src:h_src+nh]`, against numpy indexing; untouched destination bytes must stay as they were."""
    from dynamo_exp_b200.transfer import transfer_token_rows
    rng = np.random.Generator(np.random.PCG64(L * 1000 + Hs * 10 + Hd))
    ns_src, ns_dst, ntok = 40, 56, 23
    srow, drow = Hs * D * E, Hd * D * E
    src_np = [rng.integers(0, 255, size=(2, ns_src, srow), dtype=np.uint8) for _ in range(L)]
    dst_np = [rng.integers(0, 255, size=(2, ns_dst, drow), dtype=np.uint8) for _ in range(L)]
    src = [DevBuf(gpu, a.nbytes) for a in src_np]
    dst = [DevBuf(gpu, a.nbytes) for a in dst_np]
    for b, a in zip(src + dst, src_np + dst_np):
        b.upload(a)
    s_slots = rng.permutation(ns_src)[:ntok]
    d_slots = rng.permutation(ns_dst)[:ntok]
    transfer_token_rows([b.ptr for b in src], [b.ptr for b in dst], s_slots, d_slots, ns_src, ns_dst, srow, drow,
                        col_bytes=nh * D * E, src_col_offset=h_src * D * E, dst_col_offset=h_dst * D * E)
    for layer in range(L):
        want = dst_np[layer].copy().reshape(2, ns_dst, Hd, D * E)
        want[:, d_slots, h_dst:h_dst + nh] = src_np[layer].reshape(2, ns_src, Hs, D * E)[:, s_slots, h_src:h_src + nh]
        assert np.array_equal(dst[layer].download().reshape(2, ns_dst, Hd, D * E), want), layer
    with pytest.raises(KvdpError) as dup:                                                           # duplicate dst
        transfer_token_rows([src[0].ptr], [dst[0].ptr], [0, 1], [5, 5], ns_src, ns_dst, srow, drow)
    assert dup.value.code == -73
    with pytest.raises(KvError):
        transfer_token_rows([src[0].ptr], [dst[0].ptr], [0], [ns_dst], ns_src, ns_dst, srow, drow)
    with pytest.raises(KvError):
        transfer_token_rows([src[0].ptr], [dst[0].ptr], [0], [0], ns_src, ns_dst, srow, drow, col_bytes=drow + 1)


def test_disaggregated_flow_end_to_end_two_processes(gpu):
    """tests/mp_disagg_flow.py: request planning, GPU hashing, prefix reuse, IPC push with completion tracking, KV events
    through the engine ABI and the router's prefix index, decode and prefill engines as separate processes."""
    script = os.path.join(os.path.dirname(__file__), "mp_disagg_flow.py")
    r = subprocess.run([sys.executable, script], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "DISAGG FLOW OK" in r.stdout
