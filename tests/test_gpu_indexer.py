"""GPU: the prefix index (RadixTree / KvIndexer mirror over kvdp_index_*). Parity against the CPU oracle
(oracle/radix_oracle.py, itself pinned to the reference's unit tests) on the reference's own scenarios and on random
event streams with sharing, removals, worker removal and early exit; wire-format tests follow protocols.rs:160-250."""
import json

import numpy as np
import pytest

from dynamo_exp_b200.indexer import (KvCacheEvent, KvCacheEvents, KvCacheRemoveData, KvCacheStoreData,
                                     KvCacheStoredBlockData, KvIndexer, KvIndexerSharded, RadixTree, RouterEvent)
from oracle.radix_oracle import RadixTree as OracleTree
from tests.util import tokens_for

pytestmark = pytest.mark.gpu


def store_event(worker, event_id, hashes, parent=None):
    # make_blocks (indexer.rs:900-908): tokens_hash = i, block_hash = i * 100
    return RouterEvent(worker, KvCacheEvent(event_id, KvCacheStoreData(
        parent, [KvCacheStoredBlockData(i * 100, i) for i in hashes])))


def remove_event(worker, event_id, hashes):
    return RouterEvent(worker, KvCacheEvent(event_id, KvCacheRemoveData([i * 100 for i in hashes])))


def root_child(trie, tokens_hash):
    return trie.stat(4, tokens_hash)


def test_radix_tree_reference_scenario(gpu):
    """indexer.rs:949-1150, assertion for assertion (lookup sizes, root child's workers/children, shared block)."""
    trie = RadixTree(max_nodes=1024)
    w1, w2 = 0, 1
    trie.apply_event(store_event(w1, 1, [1, 2, 3]))
    assert trie.find_matches([1, 2, 3], False).scores[w1] == 3
    assert trie.stat(5) == 1 and trie.lookup_len(w1) == 3
    assert trie.stat(3, 0) == 0 and trie.stat(1, 0) == 1
    c = root_child(trie, 1)
    assert trie.stat(3, c) == 1 and trie.stat(1, c) == 1
    trie.apply_event(store_event(w2, 1, [1, 4, 5]))
    s = trie.find_matches([1, 2, 3], False).scores
    assert s[w1] == 3 and s[w2] == 1
    assert trie.stat(5) == 2 and trie.lookup_len(w1) == 3 and trie.lookup_len(w2) == 3
    assert trie.stat(3, c) == 2 and trie.stat(1, c) == 2
    trie.apply_event(remove_event(w2, 2, [5]))
    assert trie.lookup_len(w1) == 3 and trie.lookup_len(w2) == 2
    assert trie.stat(3, c) == 2 and trie.stat(1, c) == 2
    trie.apply_event(remove_event(w2, 3, [4]))
    assert trie.lookup_len(w2) == 1 and trie.stat(3, c) == 2 and trie.stat(1, c) == 2
    trie.apply_event(store_event(w2, 4, [2, 6, 7], parent=100))
    s = trie.find_matches([1, 2, 3], False).scores
    assert s[w1] == 3 and s[w2] == 2
    assert trie.lookup_len(w1) == 3 and trie.lookup_len(w2) == 4
    assert trie.stat(3, c) == 2 and trie.stat(1, c) == 2
    trie.close()


def test_remove_worker_and_early_stopping_reference_scenarios(gpu):
    trie = RadixTree(max_nodes=256)       # indexer.rs:1152-1175
    assert trie.find_matches([0], False).scores == {}
    trie.apply_event(store_event(0, 0, [0]))
    trie.apply_event(store_event(1, 0, [0]))
    assert trie.find_matches([0], False).scores == {0: 1, 1: 1}
    trie.remove_worker(0)
    assert trie.find_matches([0], False).scores == {1: 1}
    trie.close()
    trie = RadixTree(max_nodes=256)       # indexer.rs:1177-1200
    trie.apply_event(store_event(0, 0, [0, 1, 2]))
    trie.apply_event(store_event(1, 0, [0]))
    assert trie.find_matches([0, 1, 2], True).scores == {0: 2, 1: 1}
    assert trie.find_matches([0, 1], True).scores == {0: 2, 1: 1}
    assert trie.find_matches([0, 1, 2], False).scores == {0: 3, 1: 1}
    trie.close()


def test_unknown_parent_and_cleared_children(gpu):
    trie = RadixTree(max_nodes=256)
    trie.apply_event(store_event(0, 0, [1, 2], parent=999))         # skipped: unknown parent
    assert trie.find_matches([1, 2], False).scores == {}
    trie.apply_event(remove_event(0, 1, [42]))                       # ignored
    trie.apply_event(store_event(0, 2, [1, 2]))
    trie.apply_event(remove_event(0, 3, [1]))                        # last worker leaves block 1 -> children dropped
    assert trie.find_matches([1, 2], False).scores == {}
    trie.apply_event(store_event(0, 4, [1, 9]))                      # re-populating the emptied block works
    assert trie.find_matches([1, 9], False).scores == {0: 2}
    assert trie.find_matches([1, 2], False).scores == {0: 1}         # the old edge 1->2 must stay dead
    trie.close()


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_random_event_streams_match_the_oracle(gpu, seed):
    """Random stores (fresh, extending, sharing engine block ids across parents), removes, worker removals; after every
    few events a batch of random queries is scored on the GPU and by the oracle, early_exit on and off."""
    rng = np.random.Generator(np.random.PCG64(seed))
    trie, ora = RadixTree(max_nodes=1 << 14), OracleTree()
    workers = [10, -3, 7, 2 ** 40, 5]
    vocab = 12                          # small alphabet -> lots of shared prefixes
    stored = {w: [] for w in workers}   # (external hash, tokens hash) chains this worker has stored
    next_ext = [1000]

    def queries():
        qs = [list(rng.integers(0, vocab, size=int(rng.integers(0, 9)))) for _ in range(40)]
        for early in (False, True):
            got = trie.find_matches_batch(qs, early)
            for q, g in zip(qs, got):
                assert g.scores == ora.find_matches([int(x) for x in q], early), (q, early)

    for step in range(160):
        w = workers[int(rng.integers(len(workers)))]
        op = rng.random()
        if op < 0.6:
            n = int(rng.integers(1, 5))
            parent = None
            if stored[w] and rng.random() < 0.6:
                parent = stored[w][int(rng.integers(len(stored[w])))][0]
            blocks = []
            for _ in range(n):
                if stored[w] and rng.random() < 0.15:        # reuse an engine block id under a new parent (sharing)
                    ext = stored[w][int(rng.integers(len(stored[w])))][0]
                else:
                    ext = next_ext[0]
                    next_ext[0] += 1
                blocks.append((ext, int(rng.integers(vocab))))
            if rng.random() < 0.05:
                parent = 987654321                            # unknown parent: both sides must skip the event
            ev = RouterEvent(w, KvCacheEvent(step, KvCacheStoreData(
                parent, [KvCacheStoredBlockData(e, t) for e, t in blocks])))
            trie.apply_event(ev)
            if ora.apply_stored(w, parent, blocks):
                stored[w] += blocks
        elif op < 0.9:
            if stored[w]:
                k = int(rng.integers(1, 4))
                victims = [stored[w][int(rng.integers(len(stored[w])))][0] for _ in range(k)] + [31337]
                trie.apply_event(RouterEvent(w, KvCacheEvent(step, KvCacheRemoveData(victims))))
                ora.apply_removed(w, victims)
                stored[w] = [b for b in stored[w] if b[0] not in victims]
        else:
            trie.remove_worker(w)
            ora.remove_worker(w)
            stored[w] = []
        if step % 8 == 7:
            queries()
    queries()
    trie.close()


def test_find_matches_for_request_hashes_on_the_gpu(gpu, oracle):
    """indexer.rs:639-651 + the sharded/actor tests' shape (block sizes 11/32/64): tokens in, scores out."""
    for bs in (11, 32, 64):
        idx = KvIndexer(bs, max_nodes=4096)
        toks = tokens_for(1, 5 * bs + 3)
        hashes = [int(x) for x in oracle.indexer_block_hashes(toks, bs)]
        idx.apply_event(RouterEvent(0, KvCacheEvent(1, KvCacheStoreData(None, [
            KvCacheStoredBlockData(1000 + i, h) for i, h in enumerate(hashes[:4])]))))
        idx.apply_event(RouterEvent(1, KvCacheEvent(1, KvCacheStoreData(None, [
            KvCacheStoredBlockData(2000 + i, h) for i, h in enumerate(hashes[:2])]))))
        assert idx.find_matches_for_request(toks).scores == {0: 4, 1: 2}
        assert idx.find_matches(hashes).scores == {0: 4, 1: 2}
        other = tokens_for(2, 3 * bs)
        out = idx.find_matches_for_requests([toks, other, toks[:bs - 1], toks[:2 * bs]])
        assert [o.scores for o in out] == [{0: 4, 1: 2}, {}, {}, {0: 2, 1: 2}]
        idx.tree.close()


def test_router_event_wire_format(gpu):
    """protocols.rs:160-250 and the serde derives: snake_case variant keys, hashes as bare u64 numbers."""
    ev = RouterEvent(7, KvCacheEvent(1, KvCacheStoreData(1, [KvCacheStoredBlockData(2, 3)])))
    assert ev.to_json() == ('{"worker_id":7,"event":{"event_id":1,"data":{"stored":{"parent_hash":1,'
                            '"blocks":[{"block_hash":2,"tokens_hash":3}]}}}}')
    assert RouterEvent.from_json(ev.to_json()) == ev
    rm = RouterEvent(-2, KvCacheEvent(9, KvCacheRemoveData([4, 5])))
    assert json.loads(rm.to_json())["event"]["data"] == {"removed": {"block_hashes": [4, 5]}}
    assert RouterEvent.from_json(rm.to_json()) == rm
    evs = KvCacheEvents([ev.event], False)
    back = KvCacheEvents.from_json(evs.to_json())
    assert len(back.events) == 1 and back.events[0].data.parent_hash == 1 and not back.shutdown
    assert back.events[0].data.blocks[0].block_hash == 2 and back.events[0].data.blocks[0].tokens_hash == 3
    big = RouterEvent(1, KvCacheEvent(2 ** 63, KvCacheStoreData(None, [KvCacheStoredBlockData(2 ** 64 - 1, 2 ** 63 + 5)])))
    assert RouterEvent.from_json(big.to_json()) == big


@pytest.mark.parametrize("n_shards,max_workers", [(3, 64), (1, 128), (2, 512)])
def test_sharded_indexer_matches_sharded_oracle_with_100_workers(gpu, n_shards, max_workers):
    """KvIndexerSharded (indexer.rs:677-884): workers bound to the emptiest shard at their first event, lookups merged
    over shards. 100 workers over 3 shards of the default 64-worker width, over ONE index created for 128 workers (two
    64-worker words per set) and over two of the widest (512), removals and re-binding."""
    rng = np.random.default_rng(5)
    n_workers = 100
    ix = KvIndexerSharded(n_shards, 16, max_nodes=1 << 14, max_workers=max_workers)
    ref = [OracleTree() for _ in range(n_shards)]
    assign, counts = {}, [0] * n_shards

    def ref_shard(w):
        if w not in assign:
            s = min(range(n_shards), key=lambda i: counts[i])
            assign[w] = s
            counts[s] += 1
        return assign[w]
    docs = [[int(x) for x in rng.integers(1, 1 << 40, size=24)] for _ in range(12)]
    eid = 0
    for step in range(600):
        w = int(rng.integers(0, n_workers))
        r = rng.random()
        if r < 0.75:
            d = int(rng.integers(0, len(docs)))
            depth = int(rng.integers(1, 24))
            th = docs[d][:depth]
            bh = [(h ^ (w << 48)) & (2 ** 64 - 1) for h in th]          # engine-side ids are per worker
            ix.apply_event(RouterEvent(w, KvCacheEvent(eid, KvCacheStoreData(
                None, [KvCacheStoredBlockData(b, t) for b, t in zip(bh, th)]))))
            ref[ref_shard(w)].apply_stored(w, None, list(zip(bh, th)))
        elif r < 0.93:
            d = int(rng.integers(0, len(docs)))
            k = int(rng.integers(0, 24))
            victims = [(docs[d][k] ^ (w << 48)) & (2 ** 64 - 1)]
            ix.apply_event(RouterEvent(w, KvCacheEvent(eid, KvCacheRemoveData(victims))))
            ref[ref_shard(w)].apply_removed(w, victims)
        else:
            ix.remove_worker(w)
            if w in assign:
                s = assign.pop(w)
                counts[s] -= 1
                ref[s].remove_worker(w)
        eid += 1
    assert ix.worker_assignments == assign and ix.worker_counts == counts
    queries = [d[:int(rng.integers(1, 25))] for d in docs] + [[7, 8, 9]]
    got = ix.find_matches_batch(queries)
    for q, g in zip(queries, got):
        want = {}
        for t in ref:
            want.update({w: c for w, c in t.find_matches(q, False).items() if c})
        assert g.scores == want
    assert len({w for g in got for w in g.scores}) > 64             # more workers than one 64-bit set word holds
    early = ix.shards[0].find_matches_batch(queries, True)           # early exit needs the set's total count
    for q, g in zip(queries, early):
        assert g.scores == {w: c for w, c in ref[0].find_matches(q, True).items() if c}
    ix.shutdown()


def test_worker_slots_are_recycled_when_nothing_stale_is_left(gpu):
    """One index serves any number of worker ids over its lifetime as long as at most 64 are alive at once; a removed
    worker that left a stale membership behind (overwritten lookup entry, indexer.rs:380-386) keeps its slot."""
    trie, ref = RadixTree(max_nodes=1 << 12), OracleTree()
    for w in range(1000, 1300):                      # 300 distinct ids, one after the other
        ev = store_event(w, 1, [1, 2, w])
        trie.apply_event(ev)
        ref.apply_stored(w, None, [(i * 100, i) for i in [1, 2, w]])
        assert trie.find_matches([1, 2, w], False).scores == {w: 3}
        trie.remove_worker(w)
        ref.remove_worker(w)
    assert trie.find_matches([1, 2, 1299], False).scores == {}
    # a stale membership: worker 7 re-uses external hash 700 for a block that another worker already created under
    # the root, so its lookup entry for 700 moves to that node and the first node is never visited by remove_worker
    def both(worker, parent, blocks):
        trie.apply_event(RouterEvent(worker, KvCacheEvent(1, KvCacheStoreData(
            parent, [KvCacheStoredBlockData(b, t) for b, t in blocks]))))
        ref.apply_stored(worker, parent, blocks)
    both(8, None, [(900, 9)])
    both(7, None, [(700, 5)])
    both(7, None, [(700, 9)])
    trie.remove_worker(7)
    ref.remove_worker(7)
    for q in ([5], [9], [5, 9]):
        assert trie.find_matches(q, False).scores == {w: c for w, c in ref.find_matches(q, False).items() if c}
    assert trie.find_matches([5], False).scores == {7: 1}          # the stale block still reports worker 7
    assert 7 in trie._slot and trie.stat(6, trie._slot[7]) == 1   # ... so its slot is not handed out again
    assert 8 in trie._slot
    trie.remove_worker(8)
    assert 8 not in trie._slot
    trie.close()


@pytest.mark.parametrize("num_shards", [1, 3])
def test_frequency_reference_scenario_and_random_parity(gpu, num_shards):
    """indexer.rs:1305-1363 (test_frequency, for KvIndexer and KvIndexerSharded) with an explicit clock, then a random
    stream of lookups against the oracle's bookkeeping."""
    exp = 0.050
    ix = (KvIndexer.new_with_frequency(exp, 4, max_nodes=1024) if num_shards == 1
          else KvIndexerSharded(num_shards, 4, max_nodes=1024, expiration_duration=exp))
    ix.apply_event(store_event(0, 0, [1, 2, 3, 4]))
    seq, now = [1, 2, 3, 4], 0.005
    assert ix.find_matches(seq, now=now).frequencies == []
    assert ix.find_matches(seq, now=now + 0.0001).frequencies == [1, 1, 1, 1]
    now += 0.100
    assert ix.find_matches(seq, now=now).frequencies == []
    assert ix.find_matches(seq, now=now + 0.0001).frequencies == [1, 1, 1, 1]
    assert ix.find_matches(seq[:3], now=now + 0.0002).frequencies == [2, 2, 2]
    got = ix.find_matches(seq, now=now + 0.0003)
    assert got.frequencies == [3, 3, 3, 2] and got.scores == {0: 4}
    if num_shards > 1:
        ix.shutdown()
        return
    # random parity on one tree: two workers, shared prefixes, lookups at random times
    rng = np.random.default_rng(17)
    tree, ref = RadixTree(max_nodes=4096, expiration_duration=exp), OracleTree()
    docs = [[int(x) for x in rng.integers(1, 50, size=10)] for _ in range(5)]
    for w in (0, 1):
        for d in docs[w:w + 4]:
            tree.apply_event(RouterEvent(w, KvCacheEvent(0, KvCacheStoreData(
                None, [KvCacheStoredBlockData((h << 8) | w, h) for h in d]))))
            ref.apply_stored(w, None, [((h << 8) | w, h) for h in d])
    clock = 1.0
    for _ in range(200):
        clock += float(rng.choice([0.001, 0.004, 0.03, 0.2]))
        q = docs[int(rng.integers(0, 5))][:int(rng.integers(1, 11))]
        early = bool(rng.integers(0, 2))
        want_scores, want_freq = ref.find_matches_with_frequency(q, early, clock, exp)
        got = tree.find_matches(q, early, now=clock)
        assert got.scores == {w: c for w, c in want_scores.items() if c} and got.frequencies == want_freq
    tree.close()


def test_nodes_are_reclaimed_under_churn(gpu):
    """The reference frees a dropped subtree through Rc (indexer.rs:366-369). Here node ids come back through a free
    list: far more distinct blocks than `max_nodes` pass through the index over time without exhausting it, lookups stay
    right, and nothing is left behind when every worker has dropped everything."""
    max_nodes = 256
    trie, ref = RadixTree(max_nodes=max_nodes), OracleTree()
    rng = np.random.default_rng(11)
    serial = 1
    for rnd in range(120):                               # 120 x 40 = 4 800 distinct blocks through 256 node slots
        w = int(rng.integers(0, 3))
        chain = list(range(serial, serial + 40))
        serial += 40
        blocks = [(h * 100 + w, h) for h in chain]
        trie.apply_event(RouterEvent(w, KvCacheEvent(rnd, KvCacheStoreData(
            None, [KvCacheStoredBlockData(b, t) for b, t in blocks]))))
        ref.apply_stored(w, None, blocks)
        assert trie.find_matches(chain, False).scores == {w: 40}
        if rnd % 3 == 2:      # the head goes first: the rest hangs off a dead block, alive only through the lookup table
            batches = [[blocks[0][0]], [b for b, _ in blocks[1:]]]
        else:
            batches = [[b for b, _ in reversed(blocks)]]
        for victims in batches:
            trie.apply_event(RouterEvent(w, KvCacheEvent(rnd, KvCacheRemoveData(victims))))
            ref.apply_removed(w, victims)
            assert trie.find_matches(chain, False).scores == {c_w: c for c_w, c in
                                                              ref.find_matches(chain, False).items() if c}
        # everything below a chain's first block was reclaimed; the first block itself stays in the root's child map
        # with an empty worker set, as in the reference (nothing ever removes an entry from a parent's map)
        assert trie.stat(0) <= 2 + rnd
    for w in range(3):
        trie.remove_worker(w)
        ref.remove_worker(w)
    assert trie.stat(0) + trie.stat(7) <= max_nodes       # live + reclaimed ids never exceed what was created
    assert trie.find_matches(list(range(1, 41)), False).scores == {}
    # one more store after everything was dropped reuses reclaimed ids and is found
    trie.apply_event(store_event(9, 1, [1, 2, 3]))
    assert trie.find_matches([1, 2, 3], False).scores == {9: 3}


def test_wide_index_random_parity_with_300_workers(gpu):
    """An index created for 320 workers (five 64-worker words per set): random stores with shared prefixes, removals and
    worker removals against the oracle, early exit included (it needs the total size of a node's set, which the edge
    record carries when the set itself no longer fits in it)."""
    rng = np.random.default_rng(21)
    trie, ref = RadixTree(max_nodes=1 << 13, max_workers=320), OracleTree()
    docs = [[int(x) for x in rng.integers(1, 1 << 40, size=20)] for _ in range(10)]
    for step in range(900):
        w = int(rng.integers(0, 300))
        r = rng.random()
        d = docs[int(rng.integers(0, len(docs)))]
        if r < 0.8:
            th = d[:int(rng.integers(1, 21))]
            blocks = [((h ^ (w << 44)) & (2 ** 64 - 1), h) for h in th]
            trie.apply_event(RouterEvent(w, KvCacheEvent(step, KvCacheStoreData(
                None, [KvCacheStoredBlockData(b, t) for b, t in blocks]))))
            ref.apply_stored(w, None, blocks)
        elif r < 0.95:
            victims = [(d[int(rng.integers(0, 20))] ^ (w << 44)) & (2 ** 64 - 1)]
            trie.apply_event(RouterEvent(w, KvCacheEvent(step, KvCacheRemoveData(victims))))
            ref.apply_removed(w, victims)
        else:
            trie.remove_worker(w)
            ref.remove_worker(w)
    queries = [d[:int(rng.integers(1, 21))] for d in docs] * 3 + [[5, 6]]
    for early in (False, True):
        got = trie.find_matches_batch(queries, early)
        for q, g in zip(queries, got):
            assert g.scores == {w: c for w, c in ref.find_matches(q, early).items() if c}, (early, q[:3])
    assert max(len(g.scores) for g in trie.find_matches_batch(queries, False)) > 64
