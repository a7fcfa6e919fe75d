"""GPU parity: TP head reshard push (scatter TP-up, gather TP-down, equal TP) and the host tier. Bit-exact against the
CPU oracle and against the formula-independent canonical-slice property."""
import ctypes as C

import numpy as np
import pytest

from dynamo_exp_b200 import (CopyPlan, DType, HostTier, KvBlockDetails, KvBlockStorage, KvContiguousBlocks, KvLayout,
                             KvModelDetails, StorageType, TransferEngine)
from tests.util import DevBuf, payload, permutation

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _reset_options(gpu):
    yield
    for k in ("copy_path", "tile_bytes", "stages", "ctas_per_sm", "host_path", "host_ctas"):
        gpu.set_option(k, 0)


def run_reshard(gpu, oracle, src_tp, dst_tp, L, H, D, bs, E, Np, Nd, sid, did, path):
    """All prefill ranks push into all decode ranks on ONE GPU (ranks emulated as separate slabs)."""
    gpu.set_option("copy_path", path)
    rng = np.random.Generator(np.random.PCG64(src_tp * 100 + dst_tp))
    dt = {1: np.uint8, 2: np.int16, 4: np.int32, 8: np.int64}[E]
    canon = rng.integers(0, 120, size=(L, 2, Np, bs, H, D)).astype(dt)
    hp, hq = H // src_tp, H // dst_tp
    src_h = [[np.ascontiguousarray(canon[l][:, :, :, p * hp:(p + 1) * hp, :]) for l in range(L)] for p in range(src_tp)]
    src_d = [[DevBuf(gpu, a.nbytes) for a in rank] for rank in src_h]
    for rank_h, rank_d in zip(src_h, src_d):
        for a, d in zip(rank_h, rank_d):
            d.upload(a)
    dst_bytes = 2 * Nd * bs * hq * D * E
    dst_d = [[DevBuf(gpu, dst_bytes) for _ in range(L)] for _ in range(dst_tp)]
    dst_o = [[np.zeros((2, Nd, bs, hq, D), dtype=dt) for _ in range(L)] for _ in range(dst_tp)]
    eng = TransferEngine(len(sid))
    for p in range(src_tp):
        if dst_tp >= src_tp:
            m = dst_tp // src_tp
            tq = [p * m + g for g in range(m)]
        else:
            m = src_tp // dst_tp
            tq = [p // m]
        plan = CopyPlan.reshard([d.ptr for d in src_d[p]], [[d.ptr for d in dst_d[q]] for q in tq], Np, Nd, bs,
                                hp * D * E, src_tp, dst_tp, p)
        # TP-down: m prefill ranks write disjoint sub-rows of the same decode blocks, so dst ids repeat across plans
        eng.run(plan, sid, did, num_layers=L)
        assert oracle.reshard_push(src_h[p], [dst_o[q] for q in tq], Np, Nd, bs, hp * D * E, src_tp, dst_tp, p, sid,
                                   did) == 0
    eng.sync()
    for q in range(dst_tp):
        for l in range(L):
            got = dst_d[q][l].download(dt).reshape(2, Nd, bs, hq, D)
            assert np.array_equal(got, dst_o[q][l]), (q, l)
            want = canon[l][:, sid][:, :, :, q * hq:(q + 1) * hq, :]
            assert np.array_equal(got[:, did], want), (q, l)


@pytest.mark.parametrize("path", [1, 2, 4])
@pytest.mark.parametrize("src_tp,dst_tp", [(1, 1), (1, 2), (1, 4), (2, 4), (1, 8), (2, 8), (4, 2), (4, 1), (8, 2),
                                           (2, 1), (8, 8)])
def test_reshard_canonical_slice(gpu, oracle, src_tp, dst_tp, path):
    run_reshard(gpu, oracle, src_tp, dst_tp, L=3, H=8, D=16, bs=8, E=2, Np=7, Nd=9, sid=[5, 0, 3, 6],
                did=[2, 8, 0, 4], path=path)


@pytest.mark.parametrize("E,D", [(1, 16), (4, 4), (8, 2), (2, 3), (1, 5)])
def test_reshard_element_sizes_and_odd_head_dims(gpu, oracle, E, D):
    """Runs of 5..32 bytes: alignment drops below 16 and the narrow-vector path must stay exact."""
    run_reshard(gpu, oracle, 2, 4, L=2, H=8, D=D, bs=4, E=E, Np=5, Nd=5, sid=[4, 1], did=[0, 3], path=0)
    run_reshard(gpu, oracle, 4, 2, L=2, H=8, D=D, bs=4, E=E, Np=5, Nd=5, sid=[4, 1], did=[0, 3], path=0)


@pytest.mark.parametrize("bs,D,E", [(1, 8, 2), (2, 16, 2), (4, 64, 2), (16, 32, 2), (3, 5, 1), (2, 128, 4)])
def test_reshard_small_chunks_every_lane_group_width(gpu, oracle, bs, D, E):
    """Chunks of 16 B .. 4 KiB take the lane-group kernel (4, 8, 16 or 32 lanes per chunk): scatter, gather and the
    same-TP copy must stay exact for each width and for narrow vectors."""
    for src_tp, dst_tp in ((1, 2), (2, 1), (2, 2), (1, 4)):
        run_reshard(gpu, oracle, src_tp, dst_tp, L=3, H=4, D=D, bs=bs, E=E, Np=11, Nd=13,
                    sid=[10, 0, 3, 6, 7, 1, 9], did=[2, 8, 0, 4, 12, 5, 6], path=0)


def test_reshard_llama70b_tp4_to_tp2_shape(gpu, oracle):
    """BASELINE config 3 geometry (Llama-3-70B: H=8, D=128, bs=64, bf16; 512-byte runs into 1 KiB rows) on a cut-down
    layer/block count so the oracle finishes in seconds."""
    run_reshard(gpu, oracle, 4, 2, L=4, H=8, D=128, bs=64, E=2, Np=12, Nd=10, sid=list(permutation(12, 3)[:8]),
                did=list(permutation(10, 4)[:8]), path=0)
    run_reshard(gpu, oracle, 2, 4, L=4, H=8, D=128, bs=64, E=2, Np=12, Nd=10, sid=list(permutation(12, 3)[:8]),
                did=list(permutation(10, 4)[:8]), path=0)


def test_reshard_rejects_bad_geometry(gpu):
    a, b = DevBuf(gpu, 1 << 16), DevBuf(gpu, 1 << 16)
    with pytest.raises(gpu.KvdpError):
        CopyPlan.reshard([a.ptr], [[b.ptr]], 4, 4, 8, 64, 3, 4, 0)       # 4 % 3
    with pytest.raises(gpu.KvdpError):
        CopyPlan.reshard([a.ptr], [[b.ptr]], 4, 4, 8, 64, 2, 4, 2)       # rank >= tp
    with pytest.raises(gpu.KvdpError):
        CopyPlan.reshard([a.ptr], [[b.ptr], [b.ptr]], 4, 4, 8, 65, 1, 2, 0)  # row not divisible by m


# ------------------------------------------------------------------------------------------------ host tier
@pytest.mark.parametrize("layout", [KvLayout.KvFirst, KvLayout.BlockFirst])
def test_host_tier_offload_onboard_round_trip(gpu, oracle, layout):
    """SURVEY D.2 offload/onboard: D -> pinned -> D identity under a mixed batch, more blocks than one staging slot,
    non-consecutive host ids (several DMA ranges)."""
    L, nb, nh = 3, 24, 40
    bd = KvBlockDetails(layout, 8, KvModelDetails(L, 4, 16, DType.BF16))
    chunk = bd.bytes_per_token_block_per_layer() // 2
    dev = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    dev2 = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    host = KvContiguousBlocks(nh, L, chunk, StorageType.Pinned)
    src = []
    for l, layer in enumerate(dev.layers):
        data = payload(layer.storage_size(), 700 + l)
        gpu.check(gpu.lib().cuda_memcpy_sync(layer.get_pointer(), data.ctypes.data, data.nbytes))
        src.append(data)
    dev_ids = list(permutation(nb, 5)[:17])
    host_ids = [3, 4, 5, 6, 20, 21, 9, 39, 0, 1, 2, 30, 31, 32, 33, 34, 11]
    tier = HostTier(dev, host, staging_blocks=5)
    tier.offload(dev_ids, host_ids)
    tier.sync()
    want_host = np.zeros(host.storage_size(), dtype=np.uint8)
    oracle.relayout_contig(src, layout.value, nb, chunk, want_host, dev_ids, host_ids, True)
    assert np.array_equal(host.storage.as_numpy(), want_host)
    # onboard into a second device cache at different ids
    tier2 = HostTier(dev2, host, staging_blocks=4)
    new_ids = list(permutation(nb, 6)[:17])
    tier2.onboard(host_ids, new_ids)
    tier2.sync()
    want_dev = [np.zeros_like(s) for s in src]
    oracle.relayout_contig(want_dev, layout.value, nb, chunk, want_host, host_ids, new_ids, False)
    for l, layer in enumerate(dev2.layers):
        got = np.empty(layer.storage_size(), dtype=np.uint8)
        gpu.check(gpu.lib().cuda_memcpy_sync(got.ctypes.data, layer.get_pointer(), got.nbytes))
        assert np.array_equal(got, want_dev[l]), l
    with pytest.raises(gpu.KvdpError):
        tier.offload([nb], [0])
    with pytest.raises(gpu.KvdpError):
        tier.onboard([nh], [0])
    with pytest.raises(gpu.KvdpError):
        tier2.onboard([0, 1], [2, 2])  # duplicate destination blocks
    tier.close()
    tier2.close()


def test_host_tier_under_a_concurrent_copy_stream(gpu, oracle):
    """offload / onboard while another stream keeps the SMs and HBM busy (here: a second engine running
    block copies over unrelated caches, launch after launch, without synchronising) must not corrupt either side."""
    layout = KvLayout.KvFirst
    L, nb, nh = 4, 48, 48
    bd = KvBlockDetails(layout, 16, KvModelDetails(L, 4, 32, DType.BF16))
    chunk = bd.bytes_per_token_block_per_layer() // 2
    dev = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    back = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    host = KvContiguousBlocks(nh, L, chunk, StorageType.Pinned)
    src = []
    for l, layer in enumerate(dev.layers):
        data = payload(layer.storage_size(), 900 + l)
        gpu.check(gpu.lib().cuda_memcpy_sync(layer.get_pointer(), data.ctypes.data, data.nbytes))
        src.append(data)
    # the busy neighbour: its own caches, its own stream
    na = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    nbk = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    n_src = []
    for l, layer in enumerate(na.layers):
        data = payload(layer.storage_size(), 950 + l)
        gpu.check(gpu.lib().cuda_memcpy_sync(layer.get_pointer(), data.ctypes.data, data.nbytes))
        n_src.append(data)
    n_plan = CopyPlan.block_copy(na, nbk)
    n_eng = TransferEngine(nb)
    n_sid, n_did = list(permutation(nb, 21)), list(permutation(nb, 22))
    tier = HostTier(dev, host, staging_blocks=7)
    tier_b = HostTier(back, host, staging_blocks=5)
    dev_ids, host_ids, new_ids = list(permutation(nb, 23)), list(permutation(nh, 24)), list(permutation(nb, 25))
    for _ in range(6):
        n_eng.run(n_plan, n_sid, n_did, num_layers=L)          # queued, not waited for
    tier.offload(dev_ids, host_ids)
    for _ in range(6):
        n_eng.run(n_plan, n_sid, n_did, num_layers=L)
    tier.sync()
    tier_b.onboard(host_ids, new_ids)
    for _ in range(6):
        n_eng.run(n_plan, n_sid, n_did, num_layers=L)
    tier_b.sync()
    n_eng.sync()
    want_host = np.zeros(host.storage_size(), dtype=np.uint8)
    oracle.relayout_contig(src, layout.value, nb, chunk, want_host, dev_ids, host_ids, True)
    assert np.array_equal(host.storage.as_numpy(), want_host)
    want_dev = [np.zeros_like(s) for s in src]
    oracle.relayout_contig(want_dev, layout.value, nb, chunk, want_host, host_ids, new_ids, False)
    for l, layer in enumerate(back.layers):
        got = np.empty(layer.storage_size(), dtype=np.uint8)
        gpu.check(gpu.lib().cuda_memcpy_sync(got.ctypes.data, layer.get_pointer(), got.nbytes))
        assert np.array_equal(got, want_dev[l]), l
    want_n = [np.zeros_like(s) for s in n_src]
    oracle.copy_blocks_layers(n_src, want_n, n_sid, n_did, bd.prefix_dim(), nb, nb, bd.suffix_dim(), bd.elem_size())
    for l, layer in enumerate(nbk.layers):
        got = np.empty(layer.storage_size(), dtype=np.uint8)
        gpu.check(gpu.lib().cuda_memcpy_sync(got.ctypes.data, layer.get_pointer(), got.nbytes))
        assert np.array_equal(got, want_n[l]), l
    tier.close()
    tier_b.close()
    n_eng.close()


@pytest.mark.parametrize("src_tp,dst_tp", [(1, 2), (2, 4), (4, 1), (2, 16), (16, 4)])
def test_reshard_16_heads_head_size_6_as_in_the_reference_scatter_test(gpu, oracle, src_tp, dst_tp):
    """The geometry of the reference's own scatter test (kv/layer.rs:1934-2015: 16 heads, head size 6): 12-byte head rows,
    so runs of 12..96 bytes whose alignment is 4; non-identity ids."""
    run_reshard(gpu, oracle, src_tp, dst_tp, L=2, H=16, D=6, bs=4, E=2, Np=6, Nd=7, sid=[4, 1, 5], did=[0, 6, 3], path=0)
