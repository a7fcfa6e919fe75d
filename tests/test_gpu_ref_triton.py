"""GPU: the TP head-reshard rows pinned to the REFERENCE'S OWN CODE.

  * the committed golden vectors (tests/golden/rearrange_vectors.npz: output of the reference's Triton kernels and of its
    connector's descriptor arithmetic, see tests/golden/make_rearrange_vectors.py) against `rearrange_tensors`, the fused
    push (`write_blocks`) and the pull (`read_blocks`) of this library;
  * where oracle/_ref/kv_rearrange.py travelled with the snapshot (it is cut out of the reference's patch by
    oracle/ref_triton.py at build time), the reference's Triton kernels are compiled and RUN ON THIS GPU at Llama-3-70B
    shapes [N, 64, 8/tp, 128] and compared with the same three paths;
  * the NCCL connector's gather / paged write (patch:1602-1672, 1721-1746) executed as the patch's own torch expressions
    against `transfer_token_rows`.
"""
import os

import numpy as np
import pytest

from dynamo_exp_b200.connector import DynamoKvdpConnector, rearrange_tensors
from tests.golden.rearrange_cases import CONNECTOR_CASES, KERNEL_CASES, case_input, connector_inputs
from tests.util import DevBuf

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rearrange_vectors.npz")


@pytest.fixture(scope="module")
def golden():
    return dict(np.load(GOLDEN))


# ------------------------------------------------------------------------------------------ golden vectors (always run)
@pytest.mark.parametrize("i", range(len(KERNEL_CASES)))
def test_rearrange_tensors_equals_the_reference_kernel_output(gpu, golden, i):
    name, shape, d = KERNEL_CASES[i]
    t1 = case_input(10 + i, shape)
    d1, d2 = DevBuf(gpu, t1.nbytes), DevBuf(gpu, t1.nbytes)
    d1.upload(t1)
    rearrange_tensors(d1.ptr, d2.ptr, shape, 2, d, "write")
    assert np.array_equal(d2.download(np.int16).reshape(shape), golden[name + "_write"])
    t2 = case_input(50 + i, shape)
    d2.upload(t2)
    gpu.check(gpu.lib().kvdp_memset(d1.ptr, 0, t1.nbytes))
    rearrange_tensors(d1.ptr, d2.ptr, shape, 2, d, "read")
    assert np.array_equal(d1.download(np.int16).reshape(shape), golden[name + "_read"])


def _engines(gpu, layers, nl, nr, B, H, C, d, local, remote):
    """Prefill engine TP1 (rank 0) and decode engine TP d on this GPU, caches preloaded with the case inputs."""
    lb = [DevBuf(gpu, a.nbytes) for a in local]
    for b, a in zip(lb, local):
        b.upload(a)
    pre = DynamoKvdpConnector("prefill", 0, 1, device=0)
    pre.register_kv_caches([b.ptr for b in lb], shape=(2, nl, B, H, C), element_size=2)
    decs, rb = [], []
    for g in range(d):
        bufs = [DevBuf(gpu, a.nbytes) for a in remote[g]]
        for b, a in zip(bufs, remote[g]):
            b.upload(a)
        c = DynamoKvdpConnector("decode", g, d, device=0)
        c.register_kv_caches([b.ptr for b in bufs], shape=(2, nr, B, H // d, C), element_size=2)
        decs.append(c)
        rb.append(bufs)
    pre.add_remote_agent("decode", [c.get_agent_metadata() for c in decs], d, None, nr)
    return pre, decs, lb, rb


@pytest.mark.parametrize("i", range(len(CONNECTOR_CASES)))
def test_write_blocks_equals_the_reference_write_blocks_output(gpu, golden, i):
    """One fused launch, no staging blocks, against rearrange("write") + per-descriptor transfer of the reference:
    every byte of every decode rank's cache, untouched blocks included."""
    name, layers, nl, nr, B, H, C, d, lid, sid, rid = CONNECTOR_CASES[i]
    local, remote = connector_inputs(i, layers, nl, nr, B, H, C, d)
    pre, decs, lb, rb = _engines(gpu, layers, nl, nr, B, H, C, d, local, remote)
    pre.write_blocks(lid, sid, rid, "decode", "golden")
    pre.sync()
    for g in range(d):
        for l in range(layers):
            got = rb[g][l].download(np.int16).reshape(2, nr, B, H // d, C)
            assert np.array_equal(got, golden["%s_write_g%d_l%d" % (name, g, l)]), (g, l)
    for l in range(layers):      # unlike the reference, the local cache is left alone (no staging blocks are written)
        assert np.array_equal(lb[l].download(np.int16).reshape(local[l].shape), local[l])
    for c in [pre] + decs:
        c.shutdown()


@pytest.mark.parametrize("i", range(len(CONNECTOR_CASES)))
def test_read_blocks_equals_the_reference_read_blocks_output(gpu, golden, i):
    name, layers, nl, nr, B, H, C, d, lid, sid, rid = CONNECTOR_CASES[i]
    local, remote = connector_inputs(i, layers, nl, nr, B, H, C, d)
    pre, decs, lb, rb = _engines(gpu, layers, nl, nr, B, H, C, d, local, remote)
    pre.read_blocks(lid, sid, rid, "decode")
    for l in range(layers):
        got = lb[l].download(np.int16).reshape(2, nl, B, H, C)
        assert np.array_equal(got[:, lid], golden["%s_read_l%d" % (name, l)]), l
        keep = [b for b in range(nl) if b not in lid]
        assert np.array_equal(got[:, keep], local[l][:, keep]), l      # no staging blocks touched here
    for c in [pre] + decs:
        c.shutdown()


# ------------------------------------------------------------------------------------- the reference run on this GPU
@pytest.fixture(scope="module")
def ref_kernel():
    """The reference's rearrange_tensors compiled by Triton for this GPU (module-level: one JIT per direction)."""
    from oracle import ref_triton
    if not ref_triton.available():
        pytest.skip("oracle/_ref/kv_rearrange.py did not travel with this snapshot (reference tree absent at build "
                    "time); the golden-vector tests above still pin the same maps")
    assert os.environ.get("TRITON_INTERPRET", "0") in ("", "0"), "this test wants the compiled kernel"
    import torch
    mod = ref_triton.load()
    probe = torch.zeros(1, 8, 8, 16, dtype=torch.int16, device="cuda")
    mod.rearrange_tensors(probe, probe.clone(), 2, "write")      # JIT now: a Triton toolchain failure shows up here
    torch.cuda.synchronize()
    return mod.rearrange_tensors


@pytest.mark.parametrize("ptp,d", [(1, 1), (1, 2), (1, 4), (2, 2), (2, 4), (4, 2), (1, 8)])
def test_rearrange_tensors_matches_reference_triton_at_70b_shapes(gpu, oracle, ref_kernel, ptp, d):
    """Llama-3-70B: 8 KV heads x 128, 64-token blocks, bf16 (viewed as int16). [N, 64, 8/ptp, 128], N*B*H*C % 1024 == 0."""
    import torch
    N, B, H, C = 24, 64, 8 // ptp, 128
    g = torch.Generator(device="cuda")
    g.manual_seed(ptp * 10 + d)
    t1 = torch.randint(-2 ** 15, 2 ** 15, (N, B, H, C), dtype=torch.int16, device="cuda", generator=g)
    ref_out = torch.zeros_like(t1)
    ref_kernel(t1, ref_out, d, "write")
    ours = torch.zeros_like(t1)
    rearrange_tensors(t1.data_ptr(), ours.data_ptr(), (N, B, H, C), 2, d, "write")
    torch.cuda.synchronize()
    assert torch.equal(ours, ref_out)
    want = np.zeros((N, B, H, C), dtype=np.int16)
    assert oracle.rearrange(t1.cpu().numpy(), want, d, "write") == 0
    assert np.array_equal(ref_out.cpu().numpy(), want)                # the C restatement agrees with the reference too
    back_ref, back = torch.zeros_like(t1), torch.zeros_like(t1)
    ref_kernel(back_ref, ref_out, d, "read")
    rearrange_tensors(back.data_ptr(), ours.data_ptr(), (N, B, H, C), 2, d, "read")
    torch.cuda.synchronize()
    assert torch.equal(back, back_ref) and torch.equal(back, t1)


@pytest.mark.parametrize("ptp,d", [(4, 1), (2, 2), (1, 4), (2, 4), (1, 2)])
def test_fused_push_and_pull_match_reference_connector_at_70b_shapes(gpu, ref_kernel, ptp, d):
    """The whole write_blocks / read_blocks data movement of one prefill rank (patch:1296-1421) with the reference's
    kernel running on this GPU, against one fused launch of this library. 3 layers of the 70B geometry."""
    import torch

    from oracle import ref_triton
    layers, nl, nr, B, H, C = 3, 40, 36, 64, 8 // ptp, 128
    rank = ptp - 1
    gen = torch.Generator(device="cuda")
    gen.manual_seed(100 * ptp + d)

    def rnd(*shape):
        return torch.randint(-2 ** 15, 2 ** 15, shape, dtype=torch.int16, device="cuda", generator=gen)
    local = [rnd(2, nl, B, H, C) for _ in range(layers)]
    remote = [[rnd(2, nr, B, H // d, C) for _ in range(layers)] for _ in range(d)]
    lid = [3, 4, 5, 17, 9, 10, 30]
    sid = [32, 33, 34, 35, 36, 38, 39]            # staging: the reference needs spare blocks, split into three ranges
    rid = [20, 0, 7, 8, 35, 2, 11]

    ours_local = [t.clone() for t in local]
    ours_remote = [[t.clone() for t in grp] for grp in remote]
    pre = DynamoKvdpConnector("prefill", rank, ptp, device=0)
    pre.register_kv_caches(ours_local)
    decs = []
    for q in range(ptp * d):                       # the whole decode engine; only ranks rank*d+i hold real caches
        c = DynamoKvdpConnector("decode", q, ptp * d, device=0)
        i = q - rank * d
        c.register_kv_caches(ours_remote[i] if 0 <= i < d else [torch.zeros_like(t) for t in remote[0]])
        decs.append(c)
    pre.add_remote_agent("decode", [c.get_agent_metadata() for c in decs], ptp * d, None, nr)

    # write
    ref_local = [t.clone() for t in local]
    ref_remote = [[t.clone() for t in grp] for grp in remote]
    ref_triton.reference_write_blocks(ref_kernel, ref_local, ref_remote, lid, sid, rid, d)
    pre.write_blocks(lid, sid, rid, "decode", "w")
    pre.sync()
    torch.cuda.synchronize()
    for i in range(d):
        for l in range(layers):
            assert torch.equal(ours_remote[i][l], ref_remote[i][l]), ("write", i, l)
    # read (fresh local caches; the remote ones now hold what was written plus their random background)
    ref_local = [t.clone() for t in local]
    for t, s in zip(ours_local, local):
        t.copy_(s)
    ref_triton.reference_read_blocks(ref_kernel, ref_local, ref_remote, lid, sid, rid, d)
    pre.read_blocks(lid, sid, rid, "decode")
    torch.cuda.synchronize()
    for l in range(layers):
        assert torch.equal(ours_local[l][:, lid], ref_local[l][:, lid]), ("read", l)
    for c in [pre] + decs:
        c.shutdown()


# ------------------------------------------------------------------------------------- N1: the NCCL connector's data path
@pytest.mark.parametrize("tp_mult", [1, 2, 4])
def test_transfer_token_rows_equals_the_patch_expressions(gpu, tp_mult):
    """patch:1641-1666: `key_cache = kv_cache[0].reshape(-1, num_heads, head_size)`,
    `keys.append(key_cache[current_slot_mapping, head_start:head_end])` per layer and target rank, sent, then
    `reshape_and_cache_flash(keys[i], values[i], key_cache, value_cache, slot_mapping)` on the decode side
    (patch:1721-1746), i.e. cache[slot] = row. Executed here with those torch expressions; this library does all layers,
    K and V of one target rank in ONE launch."""
    import torch

    from dynamo_exp_b200.transfer import transfer_token_rows
    layers, nb_p, nb_d, bs, heads, hd = 4, 12, 15, 16, 8, 128
    gen = torch.Generator(device="cuda")
    gen.manual_seed(tp_mult)
    kv_p = [torch.randint(-2 ** 15, 2 ** 15, (2, nb_p, bs, heads, hd), dtype=torch.int16, device="cuda", generator=gen)
            for _ in range(layers)]
    hpr = heads // tp_mult
    ntok = 97
    slot_p = torch.randperm(nb_p * bs, device="cuda", generator=gen)[:ntok]
    slot_d = torch.randperm(nb_d * bs, device="cuda", generator=gen)[:ntok]
    for target_rank in range(tp_mult):
        kv_d = [torch.randint(-2 ** 15, 2 ** 15, (2, nb_d, bs, hpr, hd), dtype=torch.int16, device="cuda",
                              generator=gen) for _ in range(layers)]
        want = [t.clone() for t in kv_d]
        head_start, head_end = target_rank * hpr, (target_rank + 1) * hpr
        for layer_id in range(layers):
            key_cache = kv_p[layer_id][0].reshape(-1, heads, hd)
            value_cache = kv_p[layer_id][1].reshape(-1, heads, hd)
            keys = key_cache[slot_p, head_start:head_end]
            values = value_cache[slot_p, head_start:head_end]
            want[layer_id][0].reshape(-1, hpr, hd)[slot_d] = keys          # reshape_and_cache_flash, auto dtype
            want[layer_id][1].reshape(-1, hpr, hd)[slot_d] = values
        torch.cuda.synchronize()
        transfer_token_rows([t.data_ptr() for t in kv_p], [t.data_ptr() for t in kv_d], slot_p.cpu().numpy(),
                            slot_d.cpu().numpy(), nb_p * bs, nb_d * bs, heads * hd * 2, hpr * hd * 2,
                            col_bytes=hpr * hd * 2, src_col_offset=head_start * hd * 2)
        torch.cuda.synchronize()
        for layer_id in range(layers):
            assert torch.equal(kv_d[layer_id], want[layer_id]), (target_rank, layer_id)
