"""CPU: pins the oracle's layout maps (oracle/layout_oracle.c) against formula-independent numpy statements of the
same maps, including the reference's own test patterns."""
import numpy as np
import pytest


def test_copy_blocks_3d_reference_pattern(oracle):
    # kv/layer.rs:1536-1581: fill-with-block-id, src=[1,2,2,3,5] -> dst=[0,3,2,1,4], both layouts
    nb, bs, h, d = 8, 4, 2, 8
    for prefix, shape in ((2, (2, nb, bs * h * d)), (1, (1, nb, 2 * bs * h * d))):
        src = np.zeros(shape, dtype=np.float32)
        for b in range(nb):
            src[:, b, :] = b
        dst = np.full(shape, -1, dtype=np.float32)
        rc = oracle.copy_blocks_3d(src, dst, [1, 2, 2, 3, 5], [0, 3, 2, 1, 4], prefix, nb, nb, shape[2], 4)
        assert rc == 0
        for s, t in zip([1, 2, 2, 3, 5], [0, 3, 2, 1, 4]):
            assert np.all(dst[:, t, :] == s)
        for t in (5, 6, 7):
            assert np.all(dst[:, t, :] == -1)


def test_copy_blocks_3d_validation(oracle):
    a = np.zeros(16, dtype=np.uint8)
    assert oracle.copy_blocks_3d(a, a.copy(), [], [], 1, 1, 1, 16, 1) == 1  # num_pairs <= 0 (block_copy.cu:262-265)
    assert oracle.copy_blocks_3d(a, a.copy(), [0], [0], 0, 1, 1, 16, 1) == 1


@pytest.mark.parametrize("elem", [1, 2, 4, 8])
@pytest.mark.parametrize("prefix", [1, 2])
def test_copy_blocks_3d_random(oracle, elem, prefix):
    rng = np.random.Generator(np.random.PCG64(elem * 10 + prefix))
    ns, nd, suffix = 13, 9, 24
    src = rng.integers(0, 256, size=(prefix, ns, suffix * elem), dtype=np.uint8)
    dst = np.zeros((prefix, nd, suffix * elem), dtype=np.uint8)
    sid = rng.permutation(ns)[:7]
    did = rng.permutation(nd)[:7]
    assert oracle.copy_blocks_3d(src, dst, sid, did, prefix, ns, nd, suffix, elem) == 0
    want = np.zeros_like(dst)
    want[:, did, :] = src[:, sid, :]
    assert np.array_equal(dst, want)


@pytest.mark.parametrize("block_dim_index", [0, 1])
def test_scatter_is_permuted_axes(oracle, block_dim_index):
    # kv/layer.rs:1934-1939, 2006-2015: value formula + permuted_axes([3,0,1,2,4,5])
    nb, bs, heads, hs, sf = 3, 4, 16, 6, 4
    dnh = heads // sf
    shape5 = (nb, 2, bs, heads, hs) if block_dim_index == 0 else (2, nb, bs, heads, hs)
    src = np.zeros(shape5, dtype=np.float32)
    for i0 in range(shape5[0]):
        for i1 in range(shape5[1]):
            for t in range(bs):
                for hd in range(heads):
                    for c in range(hs):
                        src[i0, i1, t, hd, c] = 1000 * i0 + 100 * i1 + 10 * t + hd + 0.1 * c
    dims6 = [shape5[0], shape5[1], bs, sf, dnh, hs]
    dst = np.zeros(src.size, dtype=np.float32)
    assert oracle.scatter_6d(src, dst, dims6, 4, block_dim_index, nb, nb, nb) == 0
    want = src.reshape(dims6).transpose(3, 0, 1, 2, 4, 5)
    assert np.array_equal(dst.reshape(want.shape), want)


def test_scatter_return_codes(oracle):
    a = np.zeros(64, dtype=np.uint8)
    assert oracle.scatter_6d(a, a.copy(), [2, 1, 1, 1, 1], 1, 1, 1, 1, 1) == -1
    assert oracle.scatter_6d(a, a.copy(), [2, 1, 1, 1, 1, 1], 1, 2, 1, 1, 1) == -2
    assert oracle.scatter_6d(a, a.copy(), [2, 3, 1, 1, 1, 1], 1, 1, 4, 4, 1) == -3
    assert oracle.scatter_6d(a, a.copy(), [3, 1, 1, 1, 1, 1], 1, 1, 1, 1, 1) == -4


@pytest.mark.parametrize("d", [1, 2, 4])
def test_rearrange_is_triton_map(oracle, d):
    # patch:968-1001: t2 viewed [d][N][B][H/d][C] <- t1[n][b][g*H/d+h'][c]  ==  reshape(N,B,d,H/d,C).transpose(2,0,1,3,4)
    N, B, H, Cc = 3, 4, 8, 16
    rng = np.random.Generator(np.random.PCG64(d))
    t1 = rng.integers(0, 2 ** 15, size=(N, B, H, Cc), dtype=np.int16)
    t2 = np.zeros_like(t1)
    assert oracle.rearrange(t1, t2, d, "write") == 0
    want = t1.reshape(N, B, d, H // d, Cc).transpose(2, 0, 1, 3, 4)
    assert np.array_equal(t2.reshape(want.shape), want)
    back = np.zeros_like(t1)
    assert oracle.rearrange(back, t2, d, "read") == 0
    assert np.array_equal(back, t1)


@pytest.mark.parametrize("layout", [0, 1])
def test_relayout_contig(oracle, layout):
    L, N, NC, chunk = 3, 6, 5, 32
    rng = np.random.Generator(np.random.PCG64(layout))
    shape = (2, N, chunk) if layout == 0 else (N, 2, chunk)
    layers = [rng.integers(0, 256, size=shape, dtype=np.uint8) for _ in range(L)]
    contig = np.zeros((NC, L, 2, chunk), dtype=np.uint8)
    sid, did = [4, 0, 5], [1, 3, 0]
    assert oracle.relayout_contig(layers, layout, N, chunk, contig, sid, did, True, threads=2) == 0
    for s, t in zip(sid, did):
        for l in range(L):
            for kv in range(2):
                src = layers[l][kv, s] if layout == 0 else layers[l][s, kv]
                assert np.array_equal(contig[t, l, kv], src)
    # round trip back into fresh layers at new ids is the identity on the moved blocks
    layers2 = [np.zeros(shape, dtype=np.uint8) for _ in range(L)]
    assert oracle.relayout_contig(layers2, layout, N, chunk, contig, did, sid, False) == 0
    for s in sid:
        for l in range(L):
            a = layers[l][:, s] if layout == 0 else layers[l][s]
            b = layers2[l][:, s] if layout == 0 else layers2[l][s]
            assert np.array_equal(a, b)


@pytest.mark.parametrize("src_tp,dst_tp", [(1, 1), (1, 2), (1, 4), (2, 4), (2, 8), (4, 2), (4, 1), (8, 2), (2, 2), (8, 8)])
def test_reshard_canonical_slice_property(oracle, src_tp, dst_tp):
    """Formula-independent oracle check: after pushing every prefill rank's blocks, decode rank q's cache
    equals the head slice [q*H/dst_tp, (q+1)*H/dst_tp) of a canonical TP1 cache at the mapped block ids."""
    L, H, D, bs, E = 2, 8, 4, 3, 2
    Np, Nd = 7, 9
    rng = np.random.Generator(np.random.PCG64(src_tp * 16 + dst_tp))
    canon = rng.integers(0, 2 ** 15, size=(L, 2, Np, bs, H, D), dtype=np.int16)  # TP1 prefill-side content
    hp, hq = H // src_tp, H // dst_tp
    src = [[np.ascontiguousarray(canon[l][:, :, :, p * hp:(p + 1) * hp, :]) for l in range(L)] for p in range(src_tp)]
    dst = [[np.zeros((2, Nd, bs, hq, D), dtype=np.int16) for _ in range(L)] for _ in range(dst_tp)]
    sid = np.array([5, 0, 3, 6], dtype=np.int32)
    did = np.array([2, 8, 0, 4], dtype=np.int32)
    for p in range(src_tp):
        if dst_tp >= src_tp:
            m = dst_tp // src_tp
            targets = [dst[p * m + g] for g in range(m)]
        else:
            m = src_tp // dst_tp
            targets = [dst[p // m]]
        rc = oracle.reshard_push(src[p], targets, Np, Nd, bs, hp * D * E, src_tp, dst_tp, p, sid, did)
        assert rc == 0
    for q in range(dst_tp):
        for l in range(L):
            want = canon[l][:, sid][:, :, :, q * hq:(q + 1) * hq, :]
            assert np.array_equal(dst[q][l][:, did], want), (q, l)
