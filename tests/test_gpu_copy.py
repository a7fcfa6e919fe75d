"""GPU parity: block copy / relayout / legacy ABI through libkvdp.so against the CPU oracle and, where it was built,
against the reference's own block_copy.cu (oracle/_ref/libblock_copy_ref.so) on the same inputs. Bit-exact."""
import ctypes as C

import numpy as np
import pytest

from dynamo_exp_b200 import (CopyPlan, CopyStream, CopyStreamBlockMap, DType, KvBlockDetails, KvBlockStorage,
                             KvContiguousBlocks, KvError, KvLayout, KvModelDetails, StorageType, TransferEngine)
from tests.util import DevBuf, PinnedBuf, payload, permutation

pytestmark = pytest.mark.gpu

PATHS = [1, 2, 3, 4]   # 4 = lane-group-per-chunk kernel, taken when a chunk is <= 4 KiB (else the launcher falls back to auto)


@pytest.fixture(autouse=True)
def _reset_options(gpu):
    yield
    for k in ("copy_path", "tile_bytes", "stages", "ctas_per_sm", "host_path", "host_ctas", "vec32"):
        gpu.set_option(k, 0)


def details(layout, L=3, H=4, D=16, bs=8, dtype=DType.BF16, tp=1):
    return KvBlockDetails(layout=layout, block_size=bs,
                          model_details=KvModelDetails(L, H, D, dtype), tp_size=tp)


def fill_layers(gpu, storage: KvBlockStorage, seed):
    """Write a distinct payload into every layer; return the host copies (numpy uint8)."""
    out = []
    for l, layer in enumerate(storage.layers):
        n = layer.storage_size()
        data = payload(n, seed * 1000 + l)
        if storage.storage_type().kind == "pinned":
            layer.storage.storage.as_numpy()[:] = data
        else:
            gpu.check(gpu.lib().cuda_memcpy_sync(C.c_void_p(layer.get_pointer()), C.c_void_p(data.ctypes.data), n))
        out.append(data)
    return out


def read_layers(gpu, storage: KvBlockStorage):
    out = []
    for layer in storage.layers:
        n = layer.storage_size()
        buf = np.empty(n, dtype=np.uint8)
        gpu.check(gpu.lib().cuda_memcpy_sync(C.c_void_p(buf.ctypes.data), C.c_void_p(layer.get_pointer()), n))
        out.append(buf)
    return out


# ------------------------------------------------------------------------------------------------ legacy ABI
@pytest.mark.parametrize("layout", [KvLayout.KvFirst, KvLayout.BlockFirst])
def test_copy_blocks_to_reference_pattern(gpu, layout):
    """kv/layer.rs:1412-1584: fill-with-block-id, D->H with src=[1,2,2,3,5] dst=[0,3,2,1,4] (duplicate src allowed)."""
    nb = 8
    bd = details(layout, L=1, dtype=DType.F32)
    dev = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    host = KvBlockStorage.allocate(nb, bd, StorageType.Pinned)
    shape = dev.layer(0).layer_shape()
    src = np.zeros(shape, dtype=np.float32)
    for b in range(nb):
        if layout == KvLayout.KvFirst:
            src[:, b] = b
        else:
            src[b] = b
    gpu.check(gpu.lib().cuda_memcpy_sync(C.c_void_p(dev.layer(0).get_pointer()), C.c_void_p(src.ctypes.data),
                                         src.nbytes))
    dev.layer(0).copy_blocks_to([1, 2, 2, 3, 5], host.layer(0), [0, 3, 2, 1, 4])
    got = host.layer(0).view().as_numpy(np.float32)
    for s, t in zip([1, 2, 2, 3, 5], [0, 3, 2, 1, 4]):
        blk = got[:, t] if layout == KvLayout.KvFirst else got[t]
        assert np.all(blk == s)
    for t in (5, 6, 7):
        blk = got[:, t] if layout == KvLayout.KvFirst else got[t]
        assert np.all(blk == 0)


def test_copy_blocks_to_rejects_like_the_reference(gpu):
    bd = details(KvLayout.KvFirst, L=1)
    dev = KvBlockStorage.allocate(4, bd, StorageType.Device(0))
    dev2 = KvBlockStorage.allocate(4, bd, StorageType.Device(0))
    host = KvBlockStorage.allocate(4, bd, StorageType.Pinned)
    host2 = KvBlockStorage.allocate(4, bd, StorageType.Pinned)
    with pytest.raises(KvError, match="Device to Device"):
        dev.layer(0).copy_blocks_to([0], dev2.layer(0), [0])
    with pytest.raises(KvError, match="Pinned to Pinned"):
        host.layer(0).copy_blocks_to([0], host2.layer(0), [0])
    with pytest.raises(KvError, match="same length"):
        host.layer(0).copy_blocks_to([0, 1], dev.layer(0), [0])
    # Q9: an empty id list is Ok(()) in the reference (validation failure is swallowed by copy_blocks_3d)
    host.layer(0).copy_blocks_to([], dev.layer(0), [])


@pytest.mark.parametrize("layout", [KvLayout.KvFirst, KvLayout.BlockFirst])
@pytest.mark.parametrize("host_path", [0, 1])
def test_copy_stream_h2d_then_reversed_d2h(gpu, oracle, layout, host_path):
    """kv/layer.rs:1590-1735: CopyStream H2D with ids, then D2H with the ids reversed; every element checked."""
    gpu.set_option("host_path", host_path)
    nb_h, nb_d, L = 12, 16, 4
    bd = details(layout, L=L)
    host = KvBlockStorage.allocate(nb_h, bd, StorageType.Pinned)
    dev = KvBlockStorage.allocate(nb_d, bd, StorageType.Device(0))
    back = KvBlockStorage.allocate(nb_h, bd, StorageType.Pinned)
    src_data = fill_layers(gpu, host, 1)
    h_ids = [0, 3, 5, 7, 11, 2]
    d_ids = [15, 1, 4, 9, 0, 8]

    cs = CopyStream.new(L, 16)
    cs.prepare_block_map(CopyStreamBlockMap.new(host, dev))
    cs.prepare_block_ids(h_ids, d_ids)
    with pytest.raises(KvError, match="already loaded"):
        cs.prepare_block_ids(h_ids, d_ids)
    cs.trigger_all_layers()
    cs.trigger_layer(0)  # doorbell: second trigger is a no-op
    cs.sync_stream()
    cs.reset()

    # oracle on the host copies
    bdx = bd
    want_dev = [np.zeros(dev.layer(l).storage_size(), dtype=np.uint8) for l in range(L)]
    oracle.copy_blocks_layers(src_data, want_dev, h_ids, d_ids, bdx.prefix_dim(), nb_h, nb_d, bdx.suffix_dim(),
                              bdx.elem_size())
    got_dev = read_layers(gpu, dev)
    for l in range(L):
        assert np.array_equal(got_dev[l], want_dev[l]), l

    cs.prepare_block_map(CopyStreamBlockMap.new(dev, back))
    cs.prepare_block_ids(d_ids, h_ids)
    cs.trigger_all_layers()
    cs.sync_stream()
    cs.on_return()
    want_back = [np.zeros(back.layer(l).storage_size(), dtype=np.uint8) for l in range(L)]
    oracle.copy_blocks_layers(want_dev, want_back, d_ids, h_ids, bdx.prefix_dim(), nb_d, nb_h, bdx.suffix_dim(),
                              bdx.elem_size())
    for l in range(L):
        got = back.layer(l).storage.storage.as_numpy()
        assert np.array_equal(got, want_back[l]), l
    cs.close()


def test_copy_stream_state_machine_errors(gpu):
    bd = details(KvLayout.KvFirst, L=2)
    a = KvBlockStorage.allocate(4, bd, StorageType.Pinned)
    b = KvBlockStorage.allocate(4, bd, StorageType.Device(0))
    cs = CopyStream.new(2, 4)
    with pytest.raises(KvError, match="Layers must be loaded before preparing block ids"):
        cs.prepare_block_ids([0], [0])
    with pytest.raises(KvError, match="Layers must be loaded before triggering"):
        cs.trigger_layer(0)
    cs.prepare_block_map(CopyStreamBlockMap.new(a, b))
    with pytest.raises(KvError, match="Layers already loaded"):
        cs.prepare_block_map(CopyStreamBlockMap.new(a, b))
    with pytest.raises(KvError, match="Block ids must be loaded"):
        cs.trigger_layer(0)
    with pytest.raises(KvError, match="exceeds max number of blocks"):
        cs.prepare_block_ids([0, 1, 2, 3, 4], [4, 3, 2, 1, 0])
    with pytest.raises(KvError, match="dst_block_ids must be unique"):
        cs.prepare_block_ids([0, 1], [2, 2])
    with pytest.raises(KvError, match="src_block_ids must be unique"):
        cs.prepare_block_ids([1, 1], [2, 3])
    cs.prepare_block_ids([0, 1], [2, 3])
    with pytest.raises(KvError, match="out of bounds"):
        cs.trigger_layer(2)
    cs.trigger_all_layers()
    cs.sync_stream()
    big = CopyStream.new(1, 4)
    with pytest.raises(KvError, match="exceeds max number of layers"):
        big.prepare_block_map(CopyStreamBlockMap.new(a, b))


def test_legacy_return_codes(gpu):
    """cudaErrorInvalidValue (1) for NULLs / non-positive dims; copy_blocks_3d swallows them (Q9)."""
    L = gpu.lib()
    a, b = DevBuf(gpu, 4096), DevBuf(gpu, 4096)
    ids = np.zeros(4, dtype=np.int32)
    ip = C.c_void_p(ids.ctypes.data)
    for fn in (L.copy_blocks_memcpy_3d, L.copy_blocks_launcher_3d):
        assert fn(None, b.ptr, ip, ip, 1, 1, 16, 1, 4, 4, None) == 1
        assert fn(a.ptr, None, ip, ip, 1, 1, 16, 1, 4, 4, None) == 1
        assert fn(a.ptr, b.ptr, None, ip, 1, 1, 16, 1, 4, 4, None) == 1
        assert fn(a.ptr, b.ptr, ip, ip, 0, 1, 16, 1, 4, 4, None) == 1
        assert fn(a.ptr, b.ptr, ip, ip, 1, 0, 16, 1, 4, 4, None) == 1
        assert fn(a.ptr, b.ptr, ip, ip, 1, 1, 0, 1, 4, 4, None) == 1
        assert fn(a.ptr, b.ptr, ip, ip, 1, 1, 16, 0, 4, 4, None) == 1
    assert L.copy_blocks_3d(None, b.ptr, ip, ip, 1, 1, 4, 4, 16, 1) == 0
    assert L.copy_blocks_3d(a.ptr, b.ptr, ip, ip, 0, 1, 4, 4, 16, 1) == 0
    h = C.c_void_p()
    assert L.copy_stream_create(C.byref(h), 1, 4) == 0 and h.value
    dims = np.array([2, 4, 2, 2, 1, 8], dtype=np.uint32)
    dp = C.c_void_p(dims.ctypes.data)
    assert L.copy_stream_scatter(h, a.ptr, b.ptr, dp, 5, 1, 1, 4, 4) == -1
    assert L.copy_stream_scatter(h, a.ptr, b.ptr, dp, 6, 1, 2, 4, 4) == -2
    assert L.copy_stream_scatter(h, a.ptr, b.ptr, dp, 6, 1, 1, 5, 4) == -3
    bad = np.array([3, 4, 2, 2, 1, 8], dtype=np.uint32)
    assert L.copy_stream_scatter(h, a.ptr, b.ptr, C.c_void_p(bad.ctypes.data), 6, 1, 1, 4, 4) == -4
    assert L.copy_stream_sync(h) == 0
    assert L.copy_stream_destroy(h) == 0
    p = C.c_void_p()
    assert L.cuda_malloc_host(C.byref(p), 1024) == 0 and p.value
    assert L.cuda_memcpy_sync(a.ptr, p, 1024) == 0
    assert L.cuda_memcpy_async(p, a.ptr, 1024, None) == 0
    assert L.copy_blocks_3d(a.ptr, b.ptr, ip, ip, 1, 1, 4, 4, 16, 1) == 0  # also synchronises stream 0
    assert L.cuda_free_host(p) == 0


@pytest.mark.parametrize("block_dim_index", [0, 1])
@pytest.mark.parametrize("elem,hs", [(4, 6), (2, 8), (1, 16), (8, 2), (2, 3)])
def test_legacy_scatter_matches_reference_formula(gpu, oracle, block_dim_index, elem, hs):
    """kv/layer.rs:1865-2065: tp1 -> tp4, 16 heads; equals permuted_axes([3,0,1,2,4,5]). Identity block ids (Q4)."""
    nb_s, nb_d, bs, heads, sf = 5, 7, 4, 16, 4
    dnh = heads // sf
    dims5 = [nb_s, 2, bs, heads, hs] if block_dim_index == 0 else [2, nb_s, bs, heads, hs]
    n_src = int(np.prod(dims5)) * elem
    n_dst = sf * 2 * nb_d * bs * dnh * hs * elem
    src_h = payload(n_src, 17 + elem)
    src, dst = DevBuf(gpu, n_src), DevBuf(gpu, n_dst)
    src.upload(src_h)
    L = gpu.lib()
    h = C.c_void_p()
    assert L.copy_stream_create(C.byref(h), 1, 8) == 0
    ids = np.array([3, 1, 0], dtype=np.int32)  # staged ids are ignored by scatter; only their count (3) matters
    assert L.copy_stream_prepare_block_ids(h, ids.ctypes.data, ids.ctypes.data, 3) == 0
    dims6 = np.array([dims5[0], dims5[1], bs, sf, dnh, hs], dtype=np.uint32)
    assert L.copy_stream_scatter(h, src.ptr, dst.ptr, dims6.ctypes.data, 6, elem, block_dim_index, nb_s, nb_d) == 0
    assert L.copy_stream_sync(h) == 0
    want = np.zeros(n_dst, dtype=np.uint8)
    assert oracle.scatter_6d(src_h, want, dims6, elem, block_dim_index, nb_s, nb_d, 3) == 0
    assert np.array_equal(dst.download(), want)
    assert L.copy_stream_destroy(h) == 0


def test_legacy_symbols_match_the_real_reference_library(gpu, oracle):
    """oracle/_ref/libblock_copy_ref.so is the reference's own block_copy.cu compiled unmodified (oracle/Makefile).
    Same inputs through both libraries' legacy entry points must give identical bytes."""
    ref = oracle.ref_lib()
    if ref is None:
        pytest.skip("oracle/_ref was not built (reference tree absent at build time)")
    L = gpu.lib()
    for fn in ("copy_blocks_memcpy_3d", "copy_stream_create", "copy_stream_prepare_block_ids", "copy_stream_memcpy",
               "copy_stream_scatter", "copy_stream_sync"):
        getattr(ref, fn).restype = C.c_int
    prefix, ns, nd, suffix, elem = 2, 9, 11, 96, 2
    n_src, n_dst = prefix * ns * suffix * elem, prefix * nd * suffix * elem
    src_h = payload(n_src, 5)
    src = DevBuf(gpu, n_src)
    src.upload(src_h)
    sid = np.array([8, 0, 3, 3, 5], dtype=np.int32)
    did = np.array([1, 10, 4, 6, 0], dtype=np.int32)
    outs = []
    for lib_ in (ref, L):
        dst = DevBuf(gpu, n_dst)
        rc = lib_.copy_blocks_memcpy_3d(C.c_void_p(src.ptr), C.c_void_p(dst.ptr), C.c_void_p(sid.ctypes.data),
                                        C.c_void_p(did.ctypes.data), C.c_int(5), C.c_int(prefix), C.c_int(suffix),
                                        C.c_int(elem), C.c_int(ns), C.c_int(nd), C.c_void_p(0))
        assert rc == 0
        outs.append(dst.download())
    assert np.array_equal(outs[0], outs[1])
    want = np.zeros(n_dst, dtype=np.uint8)
    oracle.copy_blocks_3d(src_h, want, sid, did, prefix, ns, nd, suffix, elem)
    assert np.array_equal(outs[0], want)  # the oracle restatement agrees with the real reference too

    # scatter through a CopyStream of each library
    nb, bs, sf, dnh, hs, e = 4, 4, 4, 2, 8, 2
    dims6 = np.array([2, nb, bs, sf, dnh, hs], dtype=np.uint32)
    n = int(np.prod(dims6)) * e
    s_h = payload(n, 9)
    s = DevBuf(gpu, n)
    s.upload(s_h)
    ids = np.arange(nb, dtype=np.int32)
    outs = []
    for lib_ in (ref, L):
        d = DevBuf(gpu, n)
        h = C.c_void_p()
        assert lib_.copy_stream_create(C.byref(h), C.c_int(1), C.c_int(nb)) == 0
        assert lib_.copy_stream_prepare_block_ids(h, C.c_void_p(ids.ctypes.data), C.c_void_p(ids.ctypes.data),
                                                  C.c_int(nb)) == 0
        assert lib_.copy_stream_scatter(h, C.c_void_p(s.ptr), C.c_void_p(d.ptr), C.c_void_p(dims6.ctypes.data),
                                        C.c_uint32(6), C.c_uint32(e), C.c_uint32(1), C.c_uint32(nb),
                                        C.c_uint32(nb)) == 0
        assert lib_.copy_stream_sync(h) == 0
        outs.append(d.download())
    assert np.array_equal(outs[0], outs[1])
    want = np.zeros(n, dtype=np.uint8)
    oracle.scatter_6d(s_h, want, dims6, e, 1, nb, nb, nb)
    assert np.array_equal(outs[0], want)


# ------------------------------------------------------------------------------------------------ v2 plans
@pytest.mark.parametrize("path", PATHS)
@pytest.mark.parametrize("layout", [KvLayout.KvFirst, KvLayout.BlockFirst])
@pytest.mark.parametrize("dtype", [DType.U8, DType.BF16, DType.F32, DType.U64])
def test_plan_block_copy_d2d(gpu, oracle, path, layout, dtype):
    gpu.set_option("copy_path", path)
    gpu.set_option("tile_bytes", 4096)  # several tiles per chunk even at test sizes
    L, ns, nd = 3, 10, 14
    bd = details(layout, L=L, H=4, D=32, bs=16, dtype=dtype)
    a = KvBlockStorage.allocate(ns, bd, StorageType.Device(0))
    b = KvBlockStorage.allocate(nd, bd, StorageType.Device(0))
    src = fill_layers(gpu, a, 3)
    sid = permutation(ns, 1)[:7]
    did = permutation(nd, 2)[:7]
    plan = CopyPlan.block_copy(a, b)
    eng = TransferEngine(16)
    eng.run(plan, sid, did, num_layers=L)
    eng.sync()
    want = [np.zeros(b.layer(l).storage_size(), dtype=np.uint8) for l in range(L)]
    oracle.copy_blocks_layers(src, want, sid, did, bd.prefix_dim(), ns, nd, bd.suffix_dim(), bd.elem_size())
    got = read_layers(gpu, b)
    for l in range(L):
        assert np.array_equal(got[l], want[l]), (path, l)
    assert plan.bytes_per_pair() == bd.bytes_per_token_block_per_layer() * L


@pytest.mark.parametrize("elem,suffix", [(1, 37), (2, 37), (4, 5), (8, 3), (2, 1), (1, 1), (2, 8)])
def test_plan_block_copy_unaligned_suffix_falls_back_to_narrow_vectors(gpu, oracle, elem, suffix):
    """Odd element counts make chunk starts 1/2/4/8-byte aligned only; result must stay exact (legacy Q8 fixed)."""
    prefix, ns, nd, L = 2, 6, 5, 2
    n_src, n_dst = prefix * ns * suffix * elem, prefix * nd * suffix * elem
    srcs = [DevBuf(gpu, n_src) for _ in range(L)]
    dsts = [DevBuf(gpu, n_dst) for _ in range(L)]
    data = [payload(n_src, 40 + l) for l in range(L)]
    for s, d in zip(srcs, data):
        s.upload(d)
    plan = CopyPlan.block_copy_raw([s.ptr for s in srcs], [d.ptr for d in dsts], prefix, ns, nd, suffix, elem)
    eng = TransferEngine(8)
    sid, did = [5, 0, 2], [1, 4, 0]
    eng.run(plan, sid, did, num_layers=L)
    eng.sync()
    want = [np.zeros(n_dst, dtype=np.uint8) for _ in range(L)]
    oracle.copy_blocks_layers(data, want, sid, did, prefix, ns, nd, suffix, elem)
    for l in range(L):
        assert np.array_equal(dsts[l].download(), want[l])


@pytest.mark.parametrize("path", PATHS)
@pytest.mark.parametrize("layout", [KvLayout.KvFirst, KvLayout.BlockFirst])
def test_contig_relayout_and_round_trip(gpu, oracle, path, layout):
    gpu.set_option("copy_path", path)
    gpu.set_option("tile_bytes", 2048)
    L, nb, nc = 4, 9, 6
    bd = details(layout, L=L, H=2, D=32, bs=16)
    chunk = bd.bytes_per_token_block_per_layer() // 2
    ls = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    ls2 = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    contig = KvContiguousBlocks(nc, L, chunk, StorageType.Device(0))
    src = fill_layers(gpu, ls, 8)
    sid = [7, 0, 3, 8]
    cid = [2, 5, 0, 1]
    eng = TransferEngine(8)
    eng.run(CopyPlan.contig(ls, contig, True), sid, cid, num_layers=L)
    eng.sync()
    want = np.zeros(contig.storage_size(), dtype=np.uint8)
    oracle.relayout_contig(src, layout.value, nb, chunk, want, sid, cid, True)
    got = np.empty_like(want)
    gpu.check(gpu.lib().cuda_memcpy_sync(got.ctypes.data, contig.get_pointer(), got.nbytes))
    assert np.array_equal(got, want)
    # and back: contig[cid] -> layers2[sid]; the moved blocks are identical to the originals
    eng.run(CopyPlan.contig(ls2, contig, False), cid, sid, num_layers=L)
    eng.sync()
    want2 = [np.zeros_like(s) for s in src]
    oracle.relayout_contig(want2, layout.value, nb, chunk, want, cid, sid, False)
    got2 = read_layers(gpu, ls2)
    for l in range(L):
        assert np.array_equal(got2[l], want2[l])


def test_layer_range_and_doorbell_like_partial_launch(gpu, oracle):
    """kvdp_xfer_run moves only [layer_begin, layer_end): the v2 form of per-layer triggering."""
    L, nb = 5, 6
    bd = details(KvLayout.KvFirst, L=L)
    a = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    b = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    src = fill_layers(gpu, a, 2)
    plan = CopyPlan.block_copy(a, b)
    eng = TransferEngine(8)
    eng.run(plan, [0, 1], [3, 4], layer_begin=1, layer_end=3)
    eng.sync()
    got = read_layers(gpu, b)
    want = [np.zeros_like(s) for s in src]
    oracle.copy_blocks_layers(src[1:3], want[1:3], [0, 1], [3, 4], 2, nb, nb, bd.suffix_dim(), 2)
    for l in range(L):
        assert np.array_equal(got[l], want[l])
    with pytest.raises(Exception, match="layer range"):
        eng.run(plan, [0], [1], layer_begin=3, layer_end=9)


def test_v2_rejects_bad_ids_and_duplicate_destinations(gpu):
    bd = details(KvLayout.KvFirst, L=2)
    a = KvBlockStorage.allocate(4, bd, StorageType.Device(0))
    b = KvBlockStorage.allocate(4, bd, StorageType.Device(0))
    plan = CopyPlan.block_copy(a, b)
    eng = TransferEngine(2)
    with pytest.raises(gpu.KvdpError) as e:
        eng.run(plan, [0, 1], [2, 2], num_layers=2)
    assert e.value.code == gpu.KVDP_EDUP
    with pytest.raises(gpu.KvdpError) as e:
        eng.run(plan, [0, 4], [2, 3], num_layers=2)
    assert e.value.code == gpu.KVDP_EINVAL
    with pytest.raises(gpu.KvdpError):
        eng.run(plan, [0, -1], [2, 3], num_layers=2)
    eng.run(plan, [1, 1], [2, 3], num_layers=2)  # repeated SOURCE ids are fine (kv/layer.rs:1536-1537)
    eng.run(plan, [0, 1, 2], [1, 2, 3], num_layers=2)  # more pairs than announced: staging grows
    eng.run(plan, [], [], num_layers=2)  # empty is a no-op in v2
    eng.sync()


@pytest.mark.parametrize("path", PATHS)
def test_completion_flag_published_by_the_kernel(gpu, path):
    gpu.set_option("copy_path", path)
    L, nb = 2, 8
    bd = details(KvLayout.KvFirst, L=L, H=4, D=64, bs=32)
    a = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    b = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    fill_layers(gpu, a, 4)
    flag = PinnedBuf(gpu, 64)
    plan = CopyPlan.block_copy(a, b)
    eng = TransferEngine(8)
    for it in range(1, 4):  # the CTA counter must re-arm itself between launches
        eng.run(plan, list(range(nb)), list(range(nb))[::-1], num_layers=L, done_flag=flag.ptr, done_value=it)
        eng.sync()
        assert int(flag.array[:4].view(np.uint32)[0]) == it
    eng.run(plan, [], [], num_layers=L, done_flag=flag.ptr, done_value=77)  # "notify only" (patch:1373-1377)
    eng.sync()
    assert int(flag.array[:4].view(np.uint32)[0]) == 77
    eng.signal(flag.ptr + 4, 5)
    eng.sync()
    assert int(flag.array[4:8].view(np.uint32)[0]) == 5
    # receiver side: a stream that waits for the flag value
    dflag = DevBuf(gpu, 4)
    eng2 = TransferEngine(1)
    gpu.check(gpu.lib().kvdp_stream_wait_flag(C.c_void_p(eng2.stream), C.c_void_p(dflag.ptr), 3))
    eng2.signal(flag.ptr + 8, 9)  # ordered after the wait
    eng.signal(dflag.ptr, 3)      # releases it
    eng.sync()
    eng2.sync()
    assert int(flag.array[8:12].view(np.uint32)[0]) == 9


def test_full_size_round_trip_llama8b_isl4k(gpu):
    """BASELINE config 2 size: Llama-3-8B shape, 64 blocks (512 MiB). Property: layer-separate -> contiguous ->
    layer-separate (fresh ids) reproduces the source blocks bit-for-bit; every chunk carries a distinct payload."""
    L, H, D, bs, nb = 32, 8, 128, 64, 64
    bd = KvBlockDetails(KvLayout.KvFirst, bs, KvModelDetails(L, H, D, DType.BF16))
    chunk = bd.bytes_per_token_block_per_layer() // 2
    assert chunk == 128 * 1024
    a = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    b = KvBlockStorage.allocate(nb, bd, StorageType.Device(0))
    contig = KvContiguousBlocks(nb, L, chunk, StorageType.Device(0))
    src = fill_layers(gpu, a, 11)
    sid = permutation(nb, 1234)
    cid = permutation(nb, 1235)
    did = permutation(nb, 1236)
    eng = TransferEngine(nb)
    eng.run(CopyPlan.contig(a, contig, True), sid, cid, num_layers=L)
    eng.run(CopyPlan.contig(b, contig, False), cid, did, num_layers=L)
    eng.sync()
    got = read_layers(gpu, b)
    for l in range(L):
        s = src[l].reshape(2, nb, chunk)
        g = got[l].reshape(2, nb, chunk)
        assert np.array_equal(g[:, did], s[:, sid]), l


@pytest.mark.parametrize("src_kind,dst_kind", [("pinned", "pinned"), ("pinned", "device"), ("device", "pinned")])
def test_v2_plan_with_pinned_endpoints(gpu, oracle, src_kind, dst_kind):
    """directions H->D, D->H and P->P (v2 only: the reference rejects pinned-to-pinned, kv/layer.rs:385-387).
    Pinned memory is reached by the SMs directly (zero-copy), all layers in one launch."""
    L, ns, nd = 3, 7, 6
    bd = details(KvLayout.KvFirst, L=L, H=2, D=32, bs=8)
    kind = {"pinned": StorageType.Pinned, "device": StorageType.Device(0)}
    a = KvBlockStorage.allocate(ns, bd, kind[src_kind])
    b = KvBlockStorage.allocate(nd, bd, kind[dst_kind])
    src = fill_layers(gpu, a, 21)
    sid, did = [6, 0, 3, 3], [1, 5, 0, 2]
    eng = TransferEngine(8)
    eng.run(CopyPlan.block_copy(a, b), sid, did, num_layers=L)
    eng.sync()
    want = [np.zeros(b.layer(l).storage_size(), dtype=np.uint8) for l in range(L)]
    oracle.copy_blocks_layers(src, want, sid, did, 2, ns, nd, bd.suffix_dim(), 2)
    for l in range(L):
        got = b.layer(l).storage.storage.as_numpy() if dst_kind == "pinned" else read_layers(gpu, b)[l]
        assert np.array_equal(got, want[l]), l


def test_v2_plan_rejects_pageable_host_memory(gpu):
    buf = np.zeros(4096, dtype=np.uint8)
    d = DevBuf(gpu, 4096)
    with pytest.raises(gpu.KvdpError, match="pageable"):
        CopyPlan.block_copy_raw([buf.ctypes.data], [d.ptr], 1, 4, 4, 1024, 1)


def test_legacy_memcpy_falls_back_to_dma_for_pageable_memory(gpu, oracle):
    """copy_blocks_memcpy_3d accepts any memory cudaMemcpyDefault accepts (block_copy.cu:246-305); pageable host
    buffers cannot be read by a kernel, so that case keeps the reference's per-chunk DMA."""
    prefix, ns, nd, suffix, elem = 2, 5, 5, 64, 2
    n = prefix * ns * suffix * elem
    src_h = payload(n, 77)
    dst = DevBuf(gpu, prefix * nd * suffix * elem)
    sid = np.array([4, 0, 2], dtype=np.int32)
    did = np.array([1, 3, 0], dtype=np.int32)
    rc = gpu.lib().copy_blocks_memcpy_3d(src_h.ctypes.data, dst.ptr, sid.ctypes.data, did.ctypes.data, 3, prefix, suffix,
                                         elem, ns, nd, None)
    assert rc == 0
    want = np.zeros(prefix * nd * suffix * elem, dtype=np.uint8)
    oracle.copy_blocks_3d(src_h, want, sid, did, prefix, ns, nd, suffix, elem)
    assert np.array_equal(dst.download(), want)


def test_range_limits_are_reported_not_wrapped(gpu):
    """32-bit fields inside the kernels (chunk bytes, tile count) are guarded by the host: geometry beyond them is
    refused with KVDP_ERANGE before anything is launched (the buffers here are far smaller than the geometry claims)."""
    buf, ids = DevBuf(gpu, 1 << 20), DevBuf(gpu, 4096 * 4)
    ptr = np.array([buf.ptr], dtype=np.uint64)

    def desc(run_bytes):
        d = gpu.PlanDesc()
        d.num_layers, d.row_groups, d.prefix_dim, d.rows = 1, 1, 1, 1
        d.run_bytes = run_bytes
        d.src_block_stride = d.dst_block_stride = d.dst_row_stride = run_bytes
        d.src_layer_ptrs, d.dst_layer_ptrs, d.device = ptr.ctypes.data, ptr.ctypes.data, -1
        return d
    h = C.c_void_p()
    assert gpu.lib().kvdp_plan_create(C.byref(h), C.byref(desc(1 << 32))) == gpu.KVDP_ERANGE      # chunk >= 4 GiB
    assert "4 GiB" in gpu.last_error()
    gpu.check(gpu.lib().kvdp_plan_create(C.byref(h), C.byref(desc(1 << 31))))                    # 2 GiB chunks: fine
    gpu.set_option("copy_path", 1)
    gpu.set_option("tile_bytes", 1024)                       # 2^21 tiles per chunk
    rc = gpu.lib().kvdp_plan_launch(h, ids.ptr, ids.ptr, 1024, 0, 1, None)                       # x 1024 pairs = 2^31
    assert rc == gpu.KVDP_ERANGE and "tiles in one launch" in gpu.last_error()
    assert gpu.lib().kvdp_plan_launch(h, ids.ptr, ids.ptr, 0, 0, 1, None) == 0                   # nothing to do
    gpu.lib().kvdp_plan_destroy(h)


def _model_plan(desc, src_bufs, dst_bufs, sid, did, layer_begin, layer_end):
    """Plain-Python statement of the kvdp_plan_desc address algebra (include/kvdp.h): for every layer, prefix, pair,
    row and group one run of run_bytes."""
    L, G = desc["num_layers"], desc["row_groups"]
    for l in range(layer_begin, layer_end):
        for p in range(desc["prefix_dim"]):
            for s, d in zip(sid, did):
                for r in range(desc["rows"]):
                    for g in range(G):
                        if desc["src_row_stride"]:
                            so = (p * desc["src_prefix_stride"] + s * desc["src_block_stride"] + r * desc["src_row_stride"]
                                  + desc["src_col_offset"] + g * desc["run_bytes"])
                        else:
                            so = (p * desc["src_prefix_stride"] + s * desc["src_block_stride"]
                                  + (r * G + g) * desc["run_bytes"])
                        do = (p * desc["dst_prefix_stride"] + d * desc["dst_block_stride"] + r * desc["dst_row_stride"]
                              + desc["dst_col_offset"])
                        dst_bufs[g * L + l][do:do + desc["run_bytes"]] = src_bufs[l][so:so + desc["run_bytes"]]


@pytest.mark.parametrize("seed", range(30))
def test_random_plan_geometries_against_the_address_algebra(gpu, seed):
    """Fuzz over the whole descriptor space the kernels accept — run sizes 1 B..6 KiB at every alignment, 1..3 row groups,
    dense and strided rows on either side, column offsets, several layers and prefixes, partial layer ranges — on every
    copy path; destinations are pre-filled so that a stray write shows."""
    rng = np.random.default_rng(1000 + seed)
    for case in range(6):
        align = int(rng.choice([1, 2, 4, 8, 16, 16, 32, 32]))
        run = int(rng.integers(1, 6144 // align + 1)) * align
        if rng.random() < 0.4:
            run = int(rng.choice([16, 64, 256, 512, 2048, 4096]))
        G = int(rng.choice([1, 1, 2, 3]))
        rows = int(rng.choice([1, 1, 2, 5, 16]))
        L, prefix = int(rng.integers(1, 4)), int(rng.integers(1, 3))
        n_src, n_dst, npairs = int(rng.integers(3, 9)), int(rng.integers(3, 9)), 0
        npairs = int(rng.integers(1, min(n_src, n_dst) + 1))
        src_strided = G == 1 and rng.random() < 0.35
        pad = lambda v: (v + align - 1) // align * align
        if src_strided:
            src_row = pad(run + int(rng.integers(0, 3)) * align * 3)
            src_col = int(rng.integers(0, (src_row - run) // align + 1)) * align
            src_block = rows * src_row
        else:
            src_row, src_col, src_block = 0, 0, rows * G * run
        dst_row = pad(run + int(rng.integers(0, 3)) * align * 5) if rng.random() < 0.6 else run
        dst_col = int(rng.integers(0, (dst_row - run) // align + 1)) * align
        dst_block = rows * dst_row
        desc = dict(num_layers=L, row_groups=G, prefix_dim=prefix, rows=rows, run_bytes=run,
                    src_prefix_stride=n_src * src_block, src_block_stride=src_block,
                    dst_prefix_stride=n_dst * dst_block, dst_block_stride=dst_block, dst_row_stride=dst_row,
                    dst_col_offset=dst_col, src_row_stride=src_row, src_col_offset=src_col)
        src_np = [rng.integers(0, 255, size=prefix * n_src * src_block, dtype=np.uint8) for _ in range(L)]
        dst_np = [rng.integers(0, 255, size=prefix * n_dst * dst_block, dtype=np.uint8) for _ in range(G * L)]
        src_d = [DevBuf(gpu, a.nbytes) for a in src_np]
        dst_d = [DevBuf(gpu, a.nbytes) for a in dst_np]
        sid = [int(x) for x in rng.integers(0, n_src, size=npairs)]
        did = [int(x) for x in rng.permutation(n_dst)[:npairs]]
        lb = int(rng.integers(0, L))
        le = int(rng.integers(lb + 1, L + 1))
        d = gpu.PlanDesc()
        for k, v in desc.items():
            setattr(d, k, v)
        sp = np.array([b.ptr for b in src_d], dtype=np.uint64)
        dp = np.array([b.ptr for b in dst_d], dtype=np.uint64)
        d.src_blocks, d.dst_blocks = n_src, n_dst
        d.src_layer_ptrs, d.dst_layer_ptrs, d.device = sp.ctypes.data, dp.ctypes.data, -1
        plan = CopyPlan.from_desc(d, keepalive=(sp, dp))
        want = [a.copy() for a in dst_np]
        _model_plan(desc, src_np, want, sid, did, lb, le)
        eng = TransferEngine(npairs)
        for path in (0, 1, 2, 3, 4):
            gpu.set_option("copy_path", path)
            gpu.set_option("tile_bytes", int(rng.choice([0, 1024, 4096])))
            gpu.set_option("vec32", int(rng.integers(0, 2)))     # 256-bit accesses where everything is 32-byte aligned
            for b, a in zip(src_d, src_np):
                b.upload(a)
            for b, a in zip(dst_d, dst_np):
                b.upload(a)
            eng.run(plan, sid, did, layer_begin=lb, layer_end=le)
            eng.sync()
            for k, (b, w) in enumerate(zip(dst_d, want)):
                assert np.array_equal(b.download(), w), (seed, case, path, k, desc)
        plan.close()
        eng.close()


@pytest.mark.parametrize("host_path", [0, 3, 1])
def test_copy_stream_coalesced_dma_host_path(gpu, oracle, host_path):
    """Pinned-host endpoint with 128 KiB chunks whose host-side ids form long runs (two requests of 16 consecutive host
    blocks, listed in shuffled order): host_path 0 takes the copy-engine path (one DMA per run through HBM staging + one
    kernel), 3 the zero-copy kernel, 1 the reference's per-chunk DMA loop. Same bytes from all three, both directions,
    and a scattered id set afterwards on the same handle (falls back to the kernel) must still be ordered behind them."""
    from tests.util import PinnedBuf
    L = gpu.lib()
    gpu.set_option("host_path", host_path)
    layers, nh, nd, suffix, elem = 3, 64, 48, 64 * 1024, 2
    chunk = suffix * elem
    rng = np.random.Generator(np.random.PCG64(host_path + 10))
    host = [PinnedBuf(gpu, 2 * nh * chunk) for _ in range(layers)]
    back = [PinnedBuf(gpu, 2 * nh * chunk) for _ in range(layers)]
    dev = [DevBuf(gpu, 2 * nd * chunk) for _ in range(layers)]
    src_np = []
    for l, h in enumerate(host):
        h.array[:] = payload(h.nbytes, 300 + l)
        src_np.append(h.array.copy())
    pairs = 32
    h_ids = np.concatenate([np.arange(8, 24), np.arange(40, 56)]).astype(np.int32)
    d_ids = rng.permutation(nd)[:pairs].astype(np.int32)
    shuffle = rng.permutation(pairs)
    h_ids, d_ids = h_ids[shuffle].copy(), d_ids[shuffle].copy()
    cs = C.c_void_p()
    assert L.copy_stream_create(C.byref(cs), layers, pairs) == 0
    # onboard
    assert L.copy_stream_prepare_block_ids(cs, h_ids.ctypes.data, d_ids.ctypes.data, pairs) == 0
    for l in range(layers):
        assert L.copy_stream_memcpy(cs, host[l].ptr, dev[l].ptr, 2, suffix, elem, nh, nd) == 0, gpu.last_error()
    assert L.copy_stream_sync(cs) == 0
    want_dev = [np.zeros(2 * nd * chunk, dtype=np.uint8) for _ in range(layers)]
    oracle.copy_blocks_layers(src_np, want_dev, h_ids, d_ids, 2, nh, nd, suffix, elem)
    for l in range(layers):
        assert np.array_equal(dev[l].download(), want_dev[l]), ("onboard", l)
    # offload into other host blocks (runs again), twice in a row without a sync in between: the second id set reuses the
    # staging slots and the sorted-id buffer of the first
    h2 = np.concatenate([np.arange(0, 16), np.arange(24, 40)]).astype(np.int32)[rng.permutation(pairs)].copy()
    want_back = [np.zeros(2 * nh * chunk, dtype=np.uint8) for _ in range(layers)]
    for ids_h in (h2, ((h2 + 7) % nh).astype(np.int32)):
        assert L.copy_stream_prepare_block_ids(cs, d_ids.ctypes.data, ids_h.ctypes.data, pairs) == 0
        for l in range(layers):
            assert L.copy_stream_memcpy(cs, dev[l].ptr, back[l].ptr, 2, suffix, elem, nd, nh) == 0, gpu.last_error()
        oracle.copy_blocks_layers(want_dev, want_back, d_ids, ids_h, 2, nd, nh, suffix, elem)
    # scattered ids right behind (no runs: kernel path) overwrite a few of the same host blocks
    s_dev = rng.permutation(nd)[:6].astype(np.int32)
    s_host = np.array([1, 9, 17, 33, 41, 63], dtype=np.int32)
    assert L.copy_stream_prepare_block_ids(cs, s_dev.ctypes.data, s_host.ctypes.data, 6) == 0
    for l in range(layers):
        assert L.copy_stream_memcpy(cs, dev[l].ptr, back[l].ptr, 2, suffix, elem, nd, nh) == 0
    oracle.copy_blocks_layers(want_dev, want_back, s_dev, s_host, 2, nd, nh, suffix, elem)
    assert L.copy_stream_sync(cs) == 0
    for l in range(layers):
        assert np.array_equal(back[l].array, want_back[l]), ("offload", l)
    assert L.copy_stream_destroy(cs) == 0
    gpu.set_option("host_path", 0)


def test_copy_stream_dma_path_is_ordered_behind_zero_copy_kernels_of_the_same_handle(gpu, oracle):
    """Offload a few scattered blocks with the kernel path (SMs write host memory), then, with no sync in between, onboard a
    long run that covers those host blocks with the copy-engine path: the DMA must see what the kernel wrote."""
    L = gpu.lib()
    layers, nh, nd, suffix, elem = 2, 32, 40, 64 * 1024, 2
    chunk = suffix * elem
    host = [PinnedBuf(gpu, 2 * nh * chunk) for _ in range(layers)]
    dev_a = [DevBuf(gpu, 2 * nd * chunk) for _ in range(layers)]
    dev_b = [DevBuf(gpu, 2 * nd * chunk) for _ in range(layers)]
    a_np = []
    for l, d in enumerate(dev_a):
        a = payload(d.nbytes, 900 + l)
        d.upload(a)
        a_np.append(a)
    cs = C.c_void_p()
    assert L.copy_stream_create(C.byref(cs), layers, nh) == 0
    for rnd in range(5):
        s_dev = permutation(nd, 50 + rnd)[:5].copy()
        s_host = np.array([3, 8, 13, 21, 30], dtype=np.int32)
        assert L.copy_stream_prepare_block_ids(cs, s_dev.ctypes.data, s_host.ctypes.data, 5) == 0
        for l in range(layers):
            assert L.copy_stream_memcpy(cs, dev_a[l].ptr, host[l].ptr, 2, suffix, elem, nd, nh) == 0
        run_h = np.arange(0, 32, dtype=np.int32)
        run_d = permutation(nd, 70 + rnd)[:32].copy()
        assert L.copy_stream_prepare_block_ids(cs, run_h.ctypes.data, run_d.ctypes.data, 32) == 0
        for l in range(layers):
            assert L.copy_stream_memcpy(cs, host[l].ptr, dev_b[l].ptr, 2, suffix, elem, nh, nd) == 0
        assert L.copy_stream_sync(cs) == 0
        for l in range(layers):
            got = dev_b[l].download().reshape(2, nd, chunk)
            src = a_np[l].reshape(2, nd, chunk)
            for hb, db in zip(s_host, s_dev):
                where = int(run_d[hb])                     # host block hb went to device block run_d[hb]
                assert np.array_equal(got[:, where], src[:, int(db)]), (rnd, l, int(hb))
    assert L.copy_stream_destroy(cs) == 0


@pytest.mark.parametrize("piece_kib", [0, 64, 384, 8192])
def test_copy_stream_dma_runs_in_pieces(gpu, oracle, piece_kib):
    """host_dma_piece_kib: the copy-engine path issues each host-side run in pieces of one size. Runs of 16 and 11 chunks
    of 128 KiB against pieces of: whole runs, less than a chunk (one chunk per DMA), 3 chunks (ragged tails), larger
    than any run. Same bytes as the oracle in both directions."""
    L = gpu.lib()
    gpu.set_option("host_path", 0)
    gpu.set_option("host_dma_piece_kib", piece_kib)
    try:
        layers, nh, nd, suffix, elem = 2, 64, 40, 64 * 1024, 2
        chunk = suffix * elem
        rng = np.random.Generator(np.random.PCG64(piece_kib + 3))
        host = [PinnedBuf(gpu, 2 * nh * chunk) for _ in range(layers)]
        back = [PinnedBuf(gpu, 2 * nh * chunk) for _ in range(layers)]
        dev = [DevBuf(gpu, 2 * nd * chunk) for _ in range(layers)]
        src_np = []
        for l, h in enumerate(host):
            h.array[:] = payload(h.nbytes, 700 + l)
            src_np.append(h.array.copy())
        pairs = 27
        sh = rng.permutation(pairs)
        h_ids = np.concatenate([np.arange(5, 21), np.arange(30, 41)]).astype(np.int32)[sh].copy()
        d_ids = rng.permutation(nd)[:pairs].astype(np.int32)
        cs = C.c_void_p()
        assert L.copy_stream_create(C.byref(cs), layers, pairs) == 0
        assert L.copy_stream_prepare_block_ids(cs, h_ids.ctypes.data, d_ids.ctypes.data, pairs) == 0
        for l in range(layers):
            assert L.copy_stream_memcpy(cs, host[l].ptr, dev[l].ptr, 2, suffix, elem, nh, nd) == 0, gpu.last_error()
        assert L.copy_stream_sync(cs) == 0
        want_dev = [np.zeros(2 * nd * chunk, dtype=np.uint8) for _ in range(layers)]
        oracle.copy_blocks_layers(src_np, want_dev, h_ids, d_ids, 2, nh, nd, suffix, elem)
        for l in range(layers):
            assert np.array_equal(dev[l].download(), want_dev[l]), ("onboard", piece_kib, l)
        h2 = np.concatenate([np.arange(0, 11), np.arange(48, 64)]).astype(np.int32)[rng.permutation(pairs)].copy()
        assert L.copy_stream_prepare_block_ids(cs, d_ids.ctypes.data, h2.ctypes.data, pairs) == 0
        for l in range(layers):
            assert L.copy_stream_memcpy(cs, dev[l].ptr, back[l].ptr, 2, suffix, elem, nd, nh) == 0, gpu.last_error()
        assert L.copy_stream_sync(cs) == 0
        want_back = [np.zeros(2 * nh * chunk, dtype=np.uint8) for _ in range(layers)]
        oracle.copy_blocks_layers(want_dev, want_back, d_ids, h2, 2, nd, nh, suffix, elem)
        for l in range(layers):
            assert np.array_equal(back[l].array, want_back[l]), ("offload", piece_kib, l)
        assert L.copy_stream_destroy(cs) == 0
    finally:
        gpu.set_option("host_dma_piece_kib", 8192)


def test_copy_stream_on_a_node_shared_registered_host_region(gpu, oracle):
    """kvdp_host_register: a memfd mapping registered with CUDA behaves like cuda_malloc_host memory for the CopyStream
    (copy-engine path on host-side runs, zero-copy kernel on scattered ids), in the process that created it and in a second
    process that opened and registered the same memory (dynamo_exp_b200/shared_host.py). The second process onboards from the
    first half of the region and offloads into the second half; this process checks both against the oracle."""
    import os
    import subprocess
    import sys

    from dynamo_exp_b200.shared_host import SharedHostRegion
    L = gpu.lib()
    layers, nh, nd, suffix, elem = 2, 48, 40, 64 * 1024, 2
    chunk = suffix * elem
    layer_bytes = 2 * nh * chunk
    region = SharedHostRegion.create(2 * layers * layer_bytes)
    try:
        arr = region.array()
        arr[:layers * layer_bytes] = payload(layers * layer_bytes, 4100)
        src_np = [arr[l * layer_bytes:(l + 1) * layer_bytes].copy() for l in range(layers)]
        dev = [DevBuf(gpu, 2 * nd * chunk) for _ in range(layers)]
        rng = np.random.Generator(np.random.PCG64(91))
        for case, h_ids in (("runs", np.concatenate([np.arange(4, 20), np.arange(30, 46)])),
                            ("scattered", rng.permutation(nh)[:32])):
            h_ids = h_ids.astype(np.int32)[rng.permutation(32)].copy()
            d_ids = rng.permutation(nd)[:32].astype(np.int32)
            cs = C.c_void_p()
            assert L.copy_stream_create(C.byref(cs), layers, 32) == 0
            assert L.copy_stream_prepare_block_ids(cs, h_ids.ctypes.data, d_ids.ctypes.data, 32) == 0
            for l in range(layers):
                assert L.copy_stream_memcpy(cs, C.c_void_p(region.ptr + l * layer_bytes), dev[l].ptr, 2, suffix, elem,
                                            nh, nd) == 0, gpu.last_error()
            assert L.copy_stream_sync(cs) == 0
            want = [np.zeros(2 * nd * chunk, dtype=np.uint8) for _ in range(layers)]
            oracle.copy_blocks_layers(src_np, want, h_ids, d_ids, 2, nh, nd, suffix, elem)
            for l in range(layers):
                got = dev[l].download().reshape(2, nd, chunk)[:, d_ids]
                assert np.array_equal(got, want[l].reshape(2, nd, chunk)[:, d_ids]), (case, l)
            assert L.copy_stream_destroy(cs) == 0
        # second process: open + register, onboard blocks 0..31 of the first half, offload them into the second half
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        code = """
import ctypes as C, sys
sys.path.insert(0, %r)
import numpy as np
from dynamo_exp_b200 import _lib as lib
from dynamo_exp_b200.shared_host import SharedHostRegion
L = lib.lib()
layers, nh, nd, suffix, elem = %d, %d, %d, %d, %d
chunk = suffix * elem
layer_bytes = 2 * nh * chunk
r = SharedHostRegion.open(%r)
dev = []
for l in range(layers):
    p = C.c_void_p()
    lib.check(L.kvdp_device_alloc(C.byref(p), 2 * nd * chunk, 0))
    dev.append(p)
h = np.arange(32, dtype=np.int32)
d = np.arange(32, dtype=np.int32)[::-1].copy()
h2 = (np.arange(32, dtype=np.int32) + 16).copy()
cs = C.c_void_p()
assert L.copy_stream_create(C.byref(cs), layers, 32) == 0
lib.check(L.copy_stream_prepare_block_ids(cs, h.ctypes.data, d.ctypes.data, 32))
for l in range(layers):
    lib.check(L.copy_stream_memcpy(cs, C.c_void_p(r.ptr + l * layer_bytes), dev[l], 2, suffix, elem, nh, nd))
lib.check(L.copy_stream_prepare_block_ids(cs, d.ctypes.data, h2.ctypes.data, 32))
for l in range(layers):
    lib.check(L.copy_stream_memcpy(cs, dev[l], C.c_void_p(r.ptr + (layers + l) * layer_bytes), 2, suffix, elem, nd, nh))
lib.check(L.copy_stream_sync(cs))
L.copy_stream_destroy(cs)
r.close()
print("ok")
""" % (root, layers, nh, nd, suffix, elem, region.handle())
        out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
        assert out.returncode == 0 and out.stdout.strip().endswith("ok"), out.stderr[-2000:]
        for l in range(layers):
            a = arr[l * layer_bytes:(l + 1) * layer_bytes].reshape(2, nh, chunk)
            b = arr[(layers + l) * layer_bytes:(layers + l + 1) * layer_bytes].reshape(2, nh, chunk)
            assert np.array_equal(b[:, 16:48], a[:, 0:32]), l      # host block i went to host block i + 16
            assert not b[:, :16].any()
        del arr, a, b
    finally:
        region.close()
