"""The engine-facing KV-event ABI (include/kvdp.h §I): the five libdynamo_llm_capi.so symbols served by libkvdp.so.

CPU part: everything that needs no hashing (state machine, removed events, short blocks, outbox / JSON wire format).
GPU part: stored events — tokens_hash against the oracle's compute_block_hash_for_seq, the reference's own
`test_event_handler` scenario (lib/bindings/python/tests/test_kv_bindings.py:66-99) through the ctypes fixture shape it
uses (:141-212), and a random event stream applied to the prefix index through the sink, against the oracle tree.
One publisher per process (like the reference's OnceCell), so every test shuts it down again."""
import ctypes
import json
import threading

import numpy as np
import pytest

from dynamo_exp_b200 import _lib
from dynamo_exp_b200.publisher import KvEventPublisher, poll_raw


@pytest.fixture
def capi(kvlib):
    L = kvlib.lib()
    L.kvdp_event_sink_callback(None, None)
    L.kvdp_event_sink_index(None, 0)
    while poll_raw() is not None:
        pass
    yield L
    L.dynamo_llm_shutdown()                                      # ERR (and a line on stderr) if the test already did
    L.kvdp_event_sink_callback(None, None)
    L.kvdp_event_sink_index(None, 0)


def test_state_machine_without_init(capi, capfd):
    assert capi.dynamo_llm_shutdown() == 1                       # "Runtime not initialized" (lib.rs:117-123)
    assert capi.dynamo_kv_event_publish_removed(1, None, 0) == 1
    assert capi.dynamo_llm_load_publisher_create() == 0          # lib.rs:130-133
    assert capi.dynamo_llm_init(None, b"c", 1, 32) == 1
    assert capi.dynamo_llm_init(b"ns", b"c", 1, 0) == 1
    assert "Runtime not initialized" in capfd.readouterr().err


def test_removed_events_and_wire_format(capi):
    assert capi.dynamo_llm_init(b"kv_test", b"event", 233, 32) == 0
    assert capi.dynamo_llm_init(b"other", b"other", 7, 64) == 0  # OnceCell semantics: OK, nothing changes
    ids = (ctypes.c_uint64 * 3)(5, 2 ** 64 - 1, 7)
    assert capi.dynamo_kv_event_publish_removed(9, ids, 3) == 0
    assert capi.dynamo_kv_event_publish_removed(10, None, 0) == 0
    assert capi.kvdp_event_flush() == 0
    a, b = poll_raw(), poll_raw()
    assert poll_raw() is None
    # byte-for-byte what serde_json prints for RouterEvent (indexer.rs:139-145, protocols.rs:79-121)
    assert a == '{"worker_id":233,"event":{"event_id":9,"data":{"removed":{"block_hashes":[5,18446744073709551615,7]}}}}'
    assert b == '{"worker_id":233,"event":{"event_id":10,"data":{"removed":{"block_hashes":[]}}}}'
    assert capi.kvdp_event_stat(0) == capi.kvdp_event_stat(1)
    assert capi.dynamo_llm_shutdown() == 0
    assert capi.dynamo_llm_shutdown() == 1


def test_short_block_ends_the_list_and_event_is_still_published(capi, capfd):
    """lib.rs:191-210: a block whose token count differs from kv_block_size stops the batch; nothing to hash here."""
    assert capi.dynamo_llm_init(b"ns", b"comp", -5, 4) == 0
    toks = (ctypes.c_uint32 * 3)(1, 2, 3)
    counts = (ctypes.c_size_t * 1)(3)
    ids = (ctypes.c_uint64 * 1)(77)
    parent = (ctypes.c_uint64 * 1)(12)
    before = capi.kvdp_event_stat(4)
    assert capi.dynamo_kv_event_publish_stored(3, toks, counts, ids, 1, parent, 0) == 0
    capi.kvdp_event_flush()
    assert poll_raw() == '{"worker_id":-5,"event":{"event_id":3,"data":{"stored":{"parent_hash":12,"blocks":[]}}}}'
    assert capi.kvdp_event_stat(4) == before + 1
    assert "Block size must be 4 tokens" in capfd.readouterr().err
    assert capi.dynamo_llm_shutdown() == 0


def test_callback_sink_and_subject(capi):
    pub = KvEventPublisher("dynamo/VllmWorker", 42, 16)
    got = []
    pub.on_event(lambda subject, js: got.append((subject, js, threading.get_ident())))
    pub.publish_removed(1, [10, 20])
    pub.flush()
    assert got and got[0][0] == "namespace.dynamo.component.VllmWorker.kv_events" == pub.subject
    assert json.loads(got[0][1]) == {"worker_id": 42, "event": {"event_id": 1, "data": {"removed": {"block_hashes": [10, 20]}}}}
    assert got[0][2] != threading.get_ident()                   # delivered on the publisher thread
    assert pub.poll() is None                                   # the outbox is only used when no sink is set
    pub.on_event(None)
    pub.shutdown()


# ------------------------------------------------------------------------------------------------ GPU
class CtypesEventPublisher:
    """Same calls, argtypes and bookkeeping as the reference fixture (test_kv_bindings.py:141-212)."""

    def __init__(self, lib, namespace, component, worker_id, kv_block_size):
        self.event_id_counter = 0
        self.block_ids = []
        self.lib = lib
        assert lib.dynamo_llm_init(namespace.encode(), component.encode(), worker_id, kv_block_size) == 0

    def store_event(self, tokens, lora_id):
        parent_hash = (ctypes.c_uint64 * 1)(self.event_id_counter) if self.event_id_counter > 0 else None
        result = self.lib.dynamo_kv_event_publish_stored(
            self.event_id_counter, (ctypes.c_uint32 * len(tokens))(*tokens), (ctypes.c_size_t * 1)(len(tokens)),
            (ctypes.c_uint64 * 1)(self.event_id_counter), 1, parent_hash, lora_id)
        self.block_ids.append(self.event_id_counter)
        self.event_id_counter += 1
        assert result == 0

    def remove_event(self):
        result = self.lib.dynamo_kv_event_publish_removed(
            self.event_id_counter, (ctypes.c_uint64 * 1)(self.block_ids[-1]), 1)
        self.event_id_counter += 1
        assert result == 0

    def shutdown(self):
        assert self.lib.dynamo_llm_shutdown() == 0


@pytest.mark.gpu
def test_reference_event_handler_scenario(gpu, capi):
    """test_kv_bindings.py:66-99: store one block -> the worker scores 1 for that request; remove it -> no scores."""
    from dynamo_exp_b200.indexer import KvIndexer
    kv_block_size, worker_id = 32, 233
    indexer = KvIndexer(kv_block_size, max_nodes=1024)
    event_publisher = CtypesEventPublisher(capi, "kv_test", "event", worker_id, kv_block_size)
    capi.kvdp_event_sink_index(indexer.tree._h, indexer.tree._slot_of(worker_id, True))
    test_token = [3] * kv_block_size
    assert not indexer.find_matches_for_request(test_token).scores
    event_publisher.store_event(test_token, 0)
    capi.kvdp_event_flush()                                      # the reference sleeps a second here
    scores = indexer.find_matches_for_request(test_token)
    assert scores.scores == {worker_id: 1}
    event_publisher.remove_event()
    capi.kvdp_event_flush()
    assert not indexer.find_matches_for_request(test_token).scores
    capi.kvdp_event_sink_index(None, 0)
    event_publisher.shutdown()
    indexer.tree.close()


@pytest.mark.gpu
@pytest.mark.parametrize("bs", [4, 32, 64])
def test_stored_event_hashes_match_the_oracle(gpu, capi, oracle, bs):
    rng = np.random.default_rng(bs)
    pub = KvEventPublisher(("ns", "c"), 7, bs)
    expect = []
    for ev in range(40):
        nblk = int(rng.integers(0, 6))
        short_at = int(rng.integers(0, nblk + 1)) if rng.random() < 0.3 else None
        counts = [bs] * nblk
        if short_at is not None and short_at < nblk:
            counts[short_at] = int(rng.integers(1, bs)) if bs > 1 else 0
        toks = rng.integers(0, 128256, size=sum(counts) + 3, dtype=np.uint32)     # trailing tokens are ignored
        ids = [int(x) for x in rng.integers(0, 2 ** 63, size=nblk, dtype=np.uint64)]
        parent = int(rng.integers(0, 2 ** 63)) if rng.random() < 0.5 else None
        pub.publish_stored(ev, toks, counts, ids, 0, parent)
        keep = nblk if short_at is None or short_at >= nblk else short_at
        th = [int(x) for x in oracle.indexer_block_hashes(toks[:keep * bs], bs)] if keep else []
        expect.append({"worker_id": 7, "event": {"event_id": ev, "data": {"stored": {
            "parent_hash": parent, "blocks": [{"block_hash": ids[i], "tokens_hash": th[i]} for i in range(keep)]}}}})
    pub.flush()
    for want in expect:
        ev = pub.poll()
        assert ev is not None and json.loads(ev.to_json()) == want
    assert pub.poll() is None
    st = pub.stats()
    assert st["submitted"] == st["delivered"] and st["failed"] == 0
    pub.shutdown()


@pytest.mark.gpu
def test_known_answer_through_the_abi(gpu, capi):
    """tokens [1,2,3,4], seed 1337 -> 14643705804678351452 (lib/tokens/src/lib.rs:539-540)."""
    pub = KvEventPublisher("a/b", 1, 4)
    pub.publish_stored(0, [1, 2, 3, 4, 5, 6, 7, 8], [4, 4], [100, 200], 0)
    pub.flush()
    ev = pub.poll()
    assert [b.tokens_hash for b in ev.event.data.blocks] == [14643705804678351452, 16777012769546811212]
    pub.shutdown()


@pytest.mark.gpu
def test_event_stream_into_the_prefix_index_matches_the_oracle_tree(gpu, capi, oracle):
    """Many publishers' worth of traffic funnelled through one worker id: stored chains with shared prefixes, removals,
    unknown parents; the index fed by the sink must score like the oracle tree fed with oracle-hashed events."""
    from dynamo_exp_b200.indexer import RadixTree
    from oracle.radix_oracle import RadixTree as OracleTree
    bs, worker = 16, 5
    rng = np.random.default_rng(11)
    tree, ref = RadixTree(max_nodes=1 << 14), OracleTree()
    pub = KvEventPublisher("ns/c", worker, bs)
    pub.attach_index(tree)
    prompts = [rng.integers(0, 1000, size=bs * 12, dtype=np.uint32) for _ in range(6)]
    live, next_id, ev = [], 1, 0
    for step in range(300):
        r = rng.random()
        if r < 0.6 or not live:
            p = prompts[int(rng.integers(0, len(prompts)))]
            depth = int(rng.integers(1, 12))
            start = int(rng.integers(0, depth))
            chain = [b for b in live if b[1] is p and b[2] == start - 1]
            parent = chain[0][0] if (start > 0 and chain) else (None if start == 0 else 999999999 + step)
            n = depth - start
            ids = list(range(next_id, next_id + n))
            next_id += n
            toks = p[start * bs:depth * bs]
            pub.publish_stored(ev, toks, [bs] * n, ids, 0, parent)
            th = [int(x) for x in oracle.indexer_block_hashes(toks, bs)]
            if ref.apply_stored(worker, parent, list(zip(ids, th))):
                live += [(ids[i], p, start + i) for i in range(n)]
        else:
            k = int(rng.integers(1, 4))
            victims = [live.pop(int(rng.integers(0, len(live))))[0] for _ in range(min(k, len(live)))]
            pub.publish_removed(ev, victims)
            ref.apply_removed(worker, victims)
        ev += 1
    pub.flush()
    seqs = [[int(x) for x in oracle.indexer_block_hashes(p, bs)] for p in prompts]
    got = tree.find_matches_batch(seqs)
    for s, g in zip(seqs, got):
        assert g.scores == {w: c for w, c in ref.find_matches(s, False).items() if c}
    assert pub.stats()["parent_missing"] > 0
    pub.attach_index(None)
    pub.shutdown()
    tree.close()


@pytest.mark.gpu
def test_destroying_an_attached_index_detaches_it(gpu, capi):
    """An engine that tears its index down while the publisher still points at it must not crash: destroy detaches
    the sink after the batch in flight; later events go to the outbox."""
    from dynamo_exp_b200.indexer import RadixTree
    tree = RadixTree(max_nodes=256)
    pub = KvEventPublisher("ns/c", 3, 4)
    pub.attach_index(tree)
    for ev in range(50):
        pub.publish_stored(ev, [ev, 1, 2, 3], [4], [ev + 1], 0)
    tree.close()                                                 # waits for the queued events, then detaches
    pub.publish_removed(99, [1])
    pub.flush()
    got = pub.poll()
    assert got is not None and got.event.event_id == 99
    pub.shutdown()
