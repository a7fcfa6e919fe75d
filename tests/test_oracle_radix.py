"""CPU: pins oracle/radix_oracle.py to the reference's own RadixTree unit tests (kv_router/indexer.rs:949-1200):
same events, same assertions on scores, lookup-table sizes, child counts and shared blocks."""
from oracle.radix_oracle import RadixTree


def store(trie, worker, hashes, parent=None):
    # make_blocks (indexer.rs:900-908): tokens_hash = i, block_hash = i * 100
    return trie.apply_stored(worker, parent, [(i * 100, i) for i in hashes])


def remove(trie, worker, hashes):
    trie.apply_removed(worker, [i * 100 for i in hashes])


def test_radix_tree_like_the_reference():
    trie = RadixTree()
    w1, w2 = 0, 1
    store(trie, w1, [1, 2, 3])
    assert trie.find_matches([1, 2, 3], False)[w1] == 3
    assert len(trie.lookup) == 1 and len(trie.lookup[w1]) == 3
    assert len(trie.root.workers) == 0 and len(trie.root.children) == 1
    assert len(trie.root.children[1].workers) == 1 and len(trie.root.children[1].children) == 1

    store(trie, w2, [1, 4, 5])
    s = trie.find_matches([1, 2, 3], False)
    assert s[w1] == 3 and s[w2] == 1
    assert len(trie.lookup) == 2 and len(trie.lookup[w1]) == 3 and len(trie.lookup[w2]) == 3
    assert len(trie.root.children) == 1
    assert len(trie.root.children[1].workers) == 2 and len(trie.root.children[1].children) == 2

    remove(trie, w2, [5])
    assert len(trie.lookup[w1]) == 3 and len(trie.lookup[w2]) == 2
    assert len(trie.root.children[1].workers) == 2 and len(trie.root.children[1].children) == 2

    remove(trie, w2, [4])
    assert len(trie.lookup[w2]) == 1
    assert len(trie.root.children[1].workers) == 2 and len(trie.root.children[1].children) == 2

    store(trie, w2, [2, 6, 7], parent=100)
    s = trie.find_matches([1, 2, 3], False)
    assert s[w1] == 3 and s[w2] == 2
    assert len(trie.lookup[w1]) == 3 and len(trie.lookup[w2]) == 4
    assert len(trie.root.children[1].workers) == 2 and len(trie.root.children[1].children) == 2
    assert len(trie.lookup[w1][200].workers) == 2 and len(trie.lookup[w2][200].workers) == 2
    assert trie.lookup[w1][200] is trie.lookup[w2][200]


def test_remove_worker_like_the_reference():
    trie = RadixTree()
    assert trie.find_matches([0], False) == {}
    store(trie, 0, [0])
    store(trie, 1, [0])
    assert trie.find_matches([0], False) == {0: 1, 1: 1}
    trie.remove_worker(0)
    assert trie.find_matches([0], False) == {1: 1}


def test_early_stopping_like_the_reference():
    trie = RadixTree()
    store(trie, 0, [0, 1, 2])
    store(trie, 1, [0])
    assert trie.find_matches([0, 1, 2], True) == {0: 2, 1: 1}
    assert trie.find_matches([0, 1], True) == {0: 2, 1: 1}


def test_unknown_parent_skips_the_store_and_unknown_remove_is_ignored():
    trie = RadixTree()
    assert store(trie, 0, [1, 2], parent=999) is False          # indexer.rs:305-315
    assert trie.find_matches([1, 2], False) == {}
    remove(trie, 0, [42])                                       # indexer.rs:354-364
    store(trie, 0, [1, 2])
    remove(trie, 0, [1])                                        # last worker leaves block 1: its children are cleared
    assert trie.find_matches([1, 2], False) == {}
    assert 200 in trie.lookup[0]                                # block 2 lives on in the lookup table only


def test_frequency_like_the_reference():
    """indexer.rs:1305-1363 (test_frequency) with the clock passed in: 50 ms window; sleeps of 5 ms and 100 ms."""
    t = RadixTree()
    t.apply_stored(0, None, [(i * 100, i) for i in (1, 2, 3, 4)])
    seq, exp = [1, 2, 3, 4], 0.050
    now = 0.005
    assert t.find_matches_with_frequency(seq, False, now, exp)[1] == []
    assert t.find_matches_with_frequency(seq, False, now + 0.0001, exp)[1] == [1, 1, 1, 1]
    now += 0.100
    assert t.find_matches_with_frequency(seq, False, now, exp)[1] == []
    assert t.find_matches_with_frequency(seq, False, now + 0.0001, exp)[1] == [1, 1, 1, 1]
    assert t.find_matches_with_frequency(seq[:3], False, now + 0.0002, exp)[1] == [2, 2, 2]
    assert t.find_matches_with_frequency(seq, False, now + 0.0003, exp)[1] == [3, 3, 3, 2]
