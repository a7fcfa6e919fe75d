"""GPU parity: batched XXH3 token-block hashing (csrc/token_hash.cu) against the reference's known answers, the
committed golden vectors and the CPU oracle. Bit-exact (u64)."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from dynamo_exp_b200 import (TokenBlockSequence, Tokens, TwinTokens, compute_block_hash_for_seq, compute_hash,
                             hash_token_blocks)
from tests.util import DevBuf, tokens_for

pytestmark = pytest.mark.gpu
GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "hash_vectors.json")))


def test_reference_known_answers_on_gpu(gpu):
    # lib/tokens/src/lib.rs:539-545
    seq = TokenBlockSequence(Tokens(list(range(1, 11))), 4, 1337)
    assert len(seq.blocks()) == 2
    assert seq.blocks()[0].block_hash == 14643705804678351452
    assert seq.blocks()[0].sequence_hash == 14643705804678351452
    assert seq.blocks()[0].parent_sequence_hash is None
    assert seq.blocks()[1].block_hash == 16777012769546811212
    assert seq.blocks()[1].sequence_hash == 4945711292740353085
    assert seq.blocks()[1].parent_sequence_hash == 14643705804678351452
    assert len(seq.current_block().tokens()) == 2 and seq.current_block().remaining_tokens() == 2
    # doctest at lib/tokens/src/lib.rs:116-128
    s2 = Tokens([1, 2, 3, 4, 5]).into_sequence(4, 1337)
    assert len(s2.blocks()) == 1 and len(s2.current_block().tokens()) == 1
    assert s2.current_block().remaining_tokens() == 3
    assert s2.blocks()[0].sequence_hash == 14643705804678351452


def test_incremental_push_token_matches_batch(gpu):
    # lib/tokens/src/lib.rs:574-622
    seq = TokenBlockSequence(Tokens(), 4, 1337)
    out = [seq.push_token(t) for t in range(1, 11)]
    assert [o is not None for o in out] == [False, False, False, True, False, False, False, True, False, False]
    assert seq.blocks()[0].sequence_hash == 14643705804678351452
    assert seq.blocks()[1].block_hash == 16777012769546811212
    assert seq.blocks()[1].sequence_hash == 4945711292740353085
    assert seq.last().sequence_hash == 4945711292740353085
    assert seq.salt_hash() == 1337


def test_every_length_class_on_gpu(gpu):
    """compute_hash over every XXH3 length class with four seeds, vs the golden vectors (python xxhash)."""
    for v in GOLD["raw"]:
        data = bytes.fromhex(v["hex"])
        assert compute_hash(data, int(v["seed"])) == int(v["digest"]), (len(data), v["seed"])


def test_batch_kernel_many_lengths_vs_oracle(gpu, oracle):
    rng = np.random.Generator(np.random.PCG64(11))
    L = gpu.lib()
    for length in list(range(0, 70)) + [127, 128, 129, 200, 239, 240, 241, 256, 300, 511, 512, 513, 1024, 1500, 2048]:
        n = 33
        data = rng.integers(0, 256, size=n * max(length, 1), dtype=np.uint8)
        d_in, d_out = DevBuf(gpu, data.nbytes), DevBuf(gpu, n * 8)
        d_in.upload(data)
        for seed in (0, 1337, 0xFEEDFACECAFEBEEF):
            gpu.check(L.kvdp_xxh3_64_batch(d_in.ptr, n, length, seed, d_out.ptr, None))
            got = d_out.download(np.uint64)
            want = [oracle.xxh3_64(data[i * length:(i + 1) * length].tobytes(), seed) for i in range(n)]
            assert [int(x) for x in got] == want, (length, seed)


def test_sequence_golden_vectors_on_gpu(gpu):
    for s in GOLD["sequences"]:
        toks = np.array(s["tokens"], dtype=np.int64)
        _, bh, sh = hash_token_blocks(toks, np.array([0, toks.size]), s["block_size"], int(s["salt"]))
        assert [[str(int(a)), str(int(b))] for a, b in zip(bh, sh)] == s["blocks"], s["block_size"]


@pytest.mark.parametrize("bs", [4, 8, 11, 16, 32, 64, 128, 48, 1, 256, 300])
@pytest.mark.parametrize("salt", [0, 1337, 0x1234567890ABCDEF])
def test_ragged_batch_vs_oracle(gpu, oracle, bs, salt):
    """Ragged CSR batch incl. empty / sub-block / exact / +1 requests; TMA rows for bs 64 / 128 / 256, lane-parts kernel for bs 4 / 8 / 16 / 32, generic kernel else."""
    lens = [0, 1, bs - 1, bs, bs + 1, 2 * bs + 1, 0, 5 * bs, 3 * bs + bs // 2, 512, 1000, 33 * bs, 0]
    lens = [max(0, x) for x in lens]
    offs = np.zeros(len(lens) + 1, dtype=np.int64)
    offs[1:] = np.cumsum(lens)
    toks = np.concatenate([tokens_for(r, n) for r, n in enumerate(lens)] + [np.zeros(0, dtype=np.uint32)])
    boff, bh, sh = hash_token_blocks(toks, offs, bs, salt)
    w_off, w_bh, w_sh = oracle.hash_batch(toks, offs, bs, salt)
    assert np.array_equal(boff, w_off)
    assert np.array_equal(bh, w_bh)
    assert np.array_equal(sh, w_sh)
    _, bh_only, none = hash_token_blocks(toks, offs, bs, salt, chained=False)
    assert np.array_equal(bh_only, w_bh) and none.size == 0


@pytest.mark.parametrize("bs", [4, 8, 16, 32, 256])
def test_short_and_1k_block_fast_paths_on_a_wide_batch(gpu, oracle, bs):
    """Block sizes of 4..32 tokens take the lane-parts kernel (several lanes per block, many warps and CTAs with their
    own block ranges), 256 tokens the 1 KiB TMA rows; 3000 ragged requests with aligned and unaligned starts."""
    rng = np.random.Generator(np.random.PCG64(bs))
    lens = rng.integers(0, 40 * bs, size=3000)
    lens[rng.integers(0, lens.size, size=300)] = 0
    lens[::7] = (lens[::7] // 4) * 4                       # some requests start on 16-byte boundaries, most do not
    offs = np.zeros(lens.size + 1, dtype=np.int64)
    offs[1:] = np.cumsum(lens)
    toks = tokens_for(7, int(offs[-1]))
    boff, bh, sh = hash_token_blocks(toks, offs, bs, 1337)
    w_off, w_bh, w_sh = oracle.hash_batch(toks, offs, bs, 1337, threads=8)
    assert np.array_equal(boff, w_off) and np.array_equal(bh, w_bh) and np.array_equal(sh, w_sh)


def test_unaligned_request_starts_use_the_lane_fallback(gpu, oracle):
    """Requests whose first token is not on a 16-byte boundary cannot use cp.async.bulk; results must not change."""
    lens = [64 * 3 + 1, 64 * 2 + 2, 64 + 3, 64 * 4, 64 * 40 + 5, 64]
    offs = np.zeros(len(lens) + 1, dtype=np.int64)
    offs[1:] = np.cumsum(lens)
    toks = np.concatenate([tokens_for(100 + r, n) for r, n in enumerate(lens)])
    boff, bh, sh = hash_token_blocks(toks, offs, 64, 1337)
    w_off, w_bh, w_sh = oracle.hash_batch(toks, offs, 64, 1337)
    assert np.array_equal(boff, w_off) and np.array_equal(bh, w_bh) and np.array_equal(sh, w_sh)


def test_indexer_and_twin_semantics(gpu, oracle):
    # kv_router/indexer.rs:1202-1225: counts 1,1,2; values = seed-1337 block hashes
    for bs, n, want in ((11, 11, 1), (32, 37, 1), (64, 128, 2), (64, 63, 0)):
        toks = np.arange(n, dtype=np.uint32)
        got = compute_block_hash_for_seq(toks, bs)
        assert len(got) == want
        assert got == [int(x) for x in oracle.indexer_block_hashes(toks, bs)]
    assert TwinTokens.compute_block_hash([1, 2, 3, 4], 4) == [14643705804678351452]
    # twin quirk Q5: push_token chains the FIRST block onto parent 0; split_tokens does not
    p = TwinTokens.partial(4)
    blk = [p.push_token(t) for t in (1, 2, 3, 4)][-1]
    assert blk.block_hash == 14643705804678351452 and blk.sequence_hash == 15321365105636534287
    blocks, _ = TwinTokens.split_tokens([1, 2, 3, 4, 5, 6, 7, 8, 9, 10], 4)
    assert blocks[0].sequence_hash == 14643705804678351452 and blocks[1].sequence_hash == 4945711292740353085
    with pytest.raises(IndexError):
        TwinTokens.split_tokens([1, 2, 3], 4)
    _, bh, sh = hash_token_blocks(np.array([1, 2, 3, 4], dtype=np.uint32), np.array([0, 4]), 4, 1337,
                                  twin_first_block=True)
    assert int(sh[0]) == 15321365105636534287


def test_negative_i32_tokens(gpu, oracle):
    a = np.array([-1, -2, 3, 4, -5, 6, 7, -2147483648], dtype=np.int32)
    _, bh, sh = hash_token_blocks(a, np.array([0, 8]), 4, 7)
    w_bh, w_sh = oracle.split_tokens(a.view(np.uint32), 4, 7)
    assert np.array_equal(bh, w_bh) and np.array_equal(sh, w_sh)


def test_device_resident_api_and_large_batch_properties(gpu, oracle):
    """Router-sweep shape (ISL 512..32k, bs 64) at a size the oracle finishes in seconds, through the device-pointer
    entry point; plus properties that hold at any size: prefix stability of the chain and request independence."""
    isls = [512, 1024, 2048, 4096, 8192, 16384, 32768]
    reqs = [isls[i % len(isls)] for i in range(70)]
    offs = np.zeros(len(reqs) + 1, dtype=np.int64)
    offs[1:] = np.cumsum(reqs)
    toks = np.concatenate([tokens_for(r, n) for r, n in enumerate(reqs)])
    L = gpu.lib()
    n_req, total = len(reqs), int(offs[-1])
    nblk = total // 64
    d_tok, d_off = DevBuf(gpu, toks.nbytes), DevBuf(gpu, offs.nbytes)
    d_boff, d_bh, d_sh = DevBuf(gpu, (n_req + 1) * 8), DevBuf(gpu, nblk * 8), DevBuf(gpu, nblk * 8)
    ws_bytes = L.kvdp_hash_workspace_bytes(n_req)
    d_ws = DevBuf(gpu, ws_bytes)
    d_tok.upload(toks)
    d_off.upload(offs)
    for salt in (1337, 99):
        gpu.check(L.kvdp_hash_token_blocks(d_tok.ptr, d_off.ptr, n_req, total, 64, salt, d_boff.ptr, d_bh.ptr,
                                           d_sh.ptr, d_ws.ptr, ws_bytes, 0, None))
        boff, bh, sh = d_boff.download(np.int64), d_bh.download(np.uint64), d_sh.download(np.uint64)
        w_off, w_bh, w_sh = oracle.hash_batch(toks, offs, 64, salt, threads=4)
        assert np.array_equal(boff, w_off) and np.array_equal(bh, w_bh) and np.array_equal(sh, w_sh)
    # prefix stability: hashing only the first half of request 6 (32k tokens) gives the first half of its hashes
    r = 6
    half = toks[offs[r]:offs[r] + reqs[r] // 2]
    _, hb, hs = hash_token_blocks(half, np.array([0, half.size]), 64, 99)
    assert np.array_equal(hb, bh[boff[r]:boff[r] + hb.size]) and np.array_equal(hs, sh[boff[r]:boff[r] + hs.size])
    # bad workspace is rejected, not overrun
    assert L.kvdp_hash_token_blocks(d_tok.ptr, d_off.ptr, n_req, total, 64, 1, d_boff.ptr, d_bh.ptr, d_sh.ptr,
                                    d_ws.ptr, 8, 0, None) == gpu.KVDP_EINVAL


def test_stored_event_from_parts_like_the_c_bindings(gpu, oracle):
    """lib/bindings/c/src/lib.rs:175-227: blocks are published until the first one that is not full; tokens_hash is the
    seed-1337 un-chained hash of each block; block_hash is the engine's id."""
    from dynamo_exp_b200.tokens import kv_event_create_stored_from_parts
    bs = 16
    toks = tokens_for(3, 5 * bs)
    ev = kv_event_create_stored_from_parts(7, toks, [bs, bs, bs, 5, bs], [100, 101, 102, 103, 104], bs, parent_hash=99)
    blocks = ev["data"]["stored"]["blocks"]
    assert ev["event_id"] == 7 and ev["data"]["stored"]["parent_hash"] == 99
    assert [b["block_hash"] for b in blocks] == [100, 101, 102]          # stops at the short block
    want = oracle.indexer_block_hashes(toks[:3 * bs], bs)
    assert [b["tokens_hash"] for b in blocks] == [int(x) for x in want]
    assert kv_event_create_stored_from_parts(8, toks, [3], [1], bs)["data"]["stored"]["blocks"] == []


@pytest.mark.parametrize("seed", range(24))
def test_random_batches_block_sizes_and_salts_vs_oracle(gpu, oracle, seed):
    """Fuzz: block sizes 1..300 tokens (every kernel: lane-parts, TMA rows, generic), random request lengths with many
    empties and sub-block requests, random 64-bit salts."""
    rng = np.random.default_rng(5000 + seed)
    bs = int(rng.choice([1, 2, 3, 4, 5, 7, 8, 12, 16, 24, 32, 33, 48, 64, 100, 128, 200, 256, 300]))
    n_req = int(rng.integers(1, 400))
    lens = rng.integers(0, 12 * bs, size=n_req)
    lens[rng.random(n_req) < 0.2] = 0
    lens[rng.random(n_req) < 0.1] = bs - 1 if bs > 1 else 0
    if seed % 5 == 0:
        lens[int(rng.integers(0, n_req))] = 3000 * bs // max(1, bs // 4)       # one long request
    offs = np.zeros(n_req + 1, dtype=np.int64)
    offs[1:] = np.cumsum(lens)
    toks = rng.integers(0, 2 ** 32, size=int(offs[-1]), dtype=np.uint64).astype(np.uint32)
    salt = int(rng.integers(0, 2 ** 63)) * 2 + int(rng.integers(0, 2))
    boff, bh, sh = hash_token_blocks(toks, offs, bs, salt)
    w_off, w_bh, w_sh = oracle.hash_batch(toks, offs, bs, salt, threads=4)
    assert np.array_equal(boff, w_off) and np.array_equal(bh, w_bh) and np.array_equal(sh, w_sh), (bs, n_req, salt)


@pytest.mark.parametrize("bs,twin", [(64, False), (16, False), (24, True), (128, False)])
def test_staged_chained_call_and_pipelined_host_form_vs_oracle(gpu, oracle, bs, twin):
    """A batch big enough (6 000 requests, ~19 M tokens) that (a) the device-pointer call cuts it into request ranges and
    runs their chain kernels on side streams under the next range's block hashing, for every `hash_stages` setting, and
    (b) the host-buffer call moves it in ~18 pieces with upload, hashing and download overlapped. Ragged lengths with
    empty requests and starts off the 16-byte grid; every output word against the oracle."""
    rng = np.random.Generator(np.random.PCG64(bs))
    lens = rng.integers(0, 6400, size=6000).astype(np.int64)
    lens[rng.integers(0, lens.size, size=200)] = 0
    lens[::5] = (lens[::5] // 64) * 64
    offs = np.zeros(lens.size + 1, dtype=np.int64)
    offs[1:] = np.cumsum(lens)
    toks = tokens_for(11, int(offs[-1]))
    want = oracle.hash_batch(toks, offs, bs, 1337, oracle.hw_threads())
    if twin:          # the oracle's batch form has no twin flag: apply the quirk per request (first block onto parent 0)
        want = None
    L = gpu.lib()
    n_req, total = lens.size, int(offs[-1])
    nblk = int((lens // bs).sum())
    d_tok, d_off = DevBuf(gpu, toks.nbytes), DevBuf(gpu, offs.nbytes)
    d_boff, d_bh, d_sh = DevBuf(gpu, (n_req + 1) * 8), DevBuf(gpu, nblk * 8), DevBuf(gpu, nblk * 8)
    ws_bytes = L.kvdp_hash_workspace_bytes(n_req)
    d_ws = DevBuf(gpu, ws_bytes)
    d_tok.upload(toks)
    d_off.upload(offs)
    flags = gpu.HASH_TWIN_FIRST if twin else 0
    results = []
    try:
        for stages in (1, 0, 2, 7, 16):
            gpu.set_option("hash_stages", stages)
            gpu.check(L.kvdp_memset(d_bh.ptr, 0, nblk * 8))
            gpu.check(L.kvdp_memset(d_sh.ptr, 0, nblk * 8))
            gpu.check(L.kvdp_hash_token_blocks(d_tok.ptr, d_off.ptr, n_req, total, bs, 1337, d_boff.ptr, d_bh.ptr,
                                               d_sh.ptr, d_ws.ptr, ws_bytes, flags, None))
            gpu.check(L.kvdp_stream_synchronize(None))
            results.append((d_boff.download(np.int64), d_bh.download(np.uint64), d_sh.download(np.uint64)))
    finally:
        gpu.set_option("hash_stages", 0)
    for got in results[1:]:
        assert all(np.array_equal(a, b) for a, b in zip(got, results[0]))      # staged == single stream, bit for bit
    host = hash_token_blocks(toks, offs, bs, 1337, twin_first_block=twin)       # pipelined host-buffer form
    assert all(np.array_equal(a, b) for a, b in zip(host, results[0]))
    if want is not None:
        assert all(np.array_equal(a, b) for a, b in zip(results[0], want))
    else:
        ref_off, ref_bh, ref_sh = oracle.hash_batch(toks, offs, bs, 1337, oracle.hw_threads())
        assert np.array_equal(results[0][0], ref_off) and np.array_equal(results[0][1], ref_bh)
        r = int(np.argmax(lens // bs >= 3))                                      # one request spelled out with the quirk
        b0 = int(ref_off[r])
        _, _, one = hash_token_blocks(toks[offs[r]:offs[r + 1]], np.array([0, lens[r]]), bs, 1337, twin_first_block=True)
        assert np.array_equal(results[0][2][b0:b0 + one.size], one)
        assert one[0] != ref_sh[b0]
