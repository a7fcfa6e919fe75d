"""CPU: pins the oracle's XXH3 / lib-tokens restatement (oracle/xxh3_oracle.c) against
 (1) the reference's own known-answer values, (2) the committed golden vectors, (3) python xxhash on random input."""
import json
import os
import random

import numpy as np
import pytest

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "hash_vectors.json")))


def test_reference_known_answers(oracle):
    # /root/reference/lib/tokens/src/lib.rs:539-545 (also :127, :298, :330, :604, :621)
    k = GOLD["reference_known_answers"]
    bh, sh = oracle.split_tokens(k["tokens"], k["block_size"], k["salt"])
    assert len(bh) == 2
    assert int(bh[0]) == int(sh[0]) == 14643705804678351452 == int(k["block0_hash_and_seq"])
    assert int(bh[1]) == 16777012769546811212 == int(k["block1_hash"])
    assert int(sh[1]) == 4945711292740353085 == int(k["block1_seq"])


def test_raw_golden_vectors(oracle):
    for v in GOLD["raw"]:
        assert oracle.xxh3_64(bytes.fromhex(v["hex"]), int(v["seed"])) == int(v["digest"]), (len(v["hex"]) // 2, v["seed"])


def test_sequence_golden_vectors(oracle):
    for s in GOLD["sequences"]:
        bh, sh = oracle.split_tokens(np.array(s["tokens"], dtype=np.int64), s["block_size"], int(s["salt"]))
        assert [[str(int(a)), str(int(b))] for a, b in zip(bh, sh)] == s["blocks"]


def test_every_length_vs_python_xxhash(oracle):
    xxhash = pytest.importorskip("xxhash")
    rnd = random.Random(7)
    lens = list(range(0, 301)) + [511, 512, 513, 1023, 1024, 1025, 2047, 2048, 2049, 4096, 5000]
    for L in lens:
        data = bytes(rnd.getrandbits(8) for _ in range(L))
        for seed in (0, 1337, rnd.getrandbits(64), (1 << 64) - 1):
            assert oracle.xxh3_64(data, seed) == xxhash.xxh3_64_intdigest(data, seed=seed), (L, seed)


def test_salt_none_is_zero_and_remainder_not_hashed(oracle):
    toks = np.arange(100, dtype=np.uint32)
    bh, sh = oracle.split_tokens(toks, 32, 0)
    assert len(bh) == 3  # 100 // 32; the 4 trailing tokens stay in the partial block
    bh2, _ = oracle.split_tokens(toks[:96], 32, 0)
    assert np.array_equal(bh, bh2)
    assert len(oracle.split_tokens(toks[:31], 32, 0)[0]) == 0
    assert len(oracle.split_tokens(toks[:0], 32, 0)[0]) == 0


def test_indexer_semantics(oracle):
    # kv_router/indexer.rs:1202-1225 pins only the COUNTS for block sizes 11/32/64; values = block_hash at seed 1337
    for bs, n, want in ((11, 11, 1), (32, 32 + 5, 1), (64, 128, 2), (64, 63, 0)):
        toks = np.arange(n, dtype=np.uint32)
        got = oracle.indexer_block_hashes(toks, bs)
        assert len(got) == want
        bh, _ = oracle.split_tokens(toks, bs, 1337)
        assert np.array_equal(got, bh)


def test_twin_first_block_quirk(oracle):
    # lib/llm/src/tokens.rs:131-134: push_token hashes [0, block_hash] for the first block; split_tokens does not
    bh, sh = oracle.split_tokens([1, 2, 3, 4], 4, 1337, twin_first_block=True)
    assert int(bh[0]) == 14643705804678351452
    assert int(sh[0]) == int(GOLD["twin_first_block_push"]["seq"]) == 15321365105636534287


def test_batch_matches_single_and_threads(oracle):
    rng = np.random.Generator(np.random.PCG64(3))
    lens = [0, 1, 63, 64, 65, 129, 512, 1000, 64 * 7]
    offs = np.zeros(len(lens) + 1, dtype=np.int64)
    offs[1:] = np.cumsum(lens)
    toks = rng.integers(0, 128256, size=int(offs[-1]), dtype=np.uint32)
    out_off, bh, sh = oracle.hash_batch(toks, offs, 64, 99, threads=1)
    out_off4, bh4, sh4 = oracle.hash_batch(toks, offs, 64, 99, threads=4)
    assert np.array_equal(bh, bh4) and np.array_equal(sh, sh4) and np.array_equal(out_off, out_off4)
    for r, L in enumerate(lens):
        b, s = oracle.split_tokens(toks[offs[r]:offs[r + 1]], 64, 99)
        assert np.array_equal(bh[out_off[r]:out_off[r + 1]], b)
        assert np.array_equal(sh[out_off[r]:out_off[r + 1]], s)


def test_negative_i32_tokens_are_reinterpreted(oracle):
    # lib/tokens/src/lib.rs:91-101: `t as u32`
    a = np.array([-1, -2, 3, 4], dtype=np.int32)
    bh, _ = oracle.split_tokens(a.view(np.uint32), 4, 5)
    bh2, _ = oracle.split_tokens(np.array([0xFFFFFFFF, 0xFFFFFFFE, 3, 4], dtype=np.uint32), 4, 5)
    assert np.array_equal(bh, bh2)
