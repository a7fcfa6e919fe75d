"""Reuse pool / reserved blocks / prefill planning (mirror of lib/llm/src/kv/{reuse,reserved,manager}.rs).
CPU tests replay the reference's pool tests (reuse.rs:684-1062) with hand-made sequence hashes (the pool never looks
inside a hash); the GPU test runs the whole chain tokens -> GPU hashes -> manager -> block ids -> relayout kernel."""
import numpy as np
import pytest

from dynamo_exp_b200.reuse import (AvailableBlocks, KvBlock, KvStorageManager, ReservedBlocks, UpdateBlock)
from dynamo_exp_b200.storage import KvError
from dynamo_exp_b200.tokens import TokenBlock, Tokens


def make_blocks(values, block_size, base_hash):
    """create_blocks (reuse.rs:751-766) with synthetic chained hashes: block i of a sequence gets base_hash + i."""
    out = []
    for i in range(len(values) // block_size):
        toks = Tokens(values[i * block_size:(i + 1) * block_size])
        out.append(KvBlock(TokenBlock(toks, 1337, 7000 + base_hash + i, base_hash + i,
                                      None if i == 0 else base_hash + i - 1)))
    return out


def test_priority_key_order():
    # reuse.rs:686-748: ordered by priority, then return tick; the sequence hash does not take part
    pool = AvailableBlocks()
    pool.priority_set[(0, 1)] = 1
    pool.priority_set[(1, 0)] = 2
    pool.priority_set[(0, 2)] = 3
    assert [pool.priority_set.popitem(0) for _ in range(3)] == [((0, 1), 1), ((0, 2), 3), ((1, 0), 2)]


def test_basic_sequence_matching():
    pool = AvailableBlocks()                                   # reuse.rs:768-808
    blocks = make_blocks([1, 2, 3, 4], 2, 100)
    hashes = [b.token_block.sequence_hash for b in blocks]
    for b in blocks:
        pool.insert(b)
    pool.fence()
    assert pool.total_blocks() == 2 and pool.available_blocks() == 2
    matched = pool.match_blocks(hashes)
    assert len(matched) == 2 and pool.total_blocks() == 2 and pool.available_blocks() == 0
    assert [m.token_block.sequence_hash for m in matched] == hashes
    for m in reversed(matched):
        m.drop()                                               # return to pool
    assert pool.total_blocks() == 2 and pool.available_blocks() == 2
    assert pool.match_blocks([hashes[0], 999, hashes[1]])[0].token_block.sequence_hash == hashes[0]


def _two_sequences(p1, p2):
    b1, b2 = make_blocks([1, 2, 3, 4], 2, 100), make_blocks([5, 6, 7, 8], 2, 200)
    for b in b1:
        b.priority = p1
    for b in b2:
        b.priority = p2
    return b1, b2


def test_equal_priority_taking():
    pool = AvailableBlocks()                                   # reuse.rs:810-845: oldest return tick first
    b1, b2 = _two_sequences(1, 1)
    for b in reversed(b2):
        pool.insert(b)
    for b in reversed(b1):
        pool.insert(b)
    taken = pool.take_blocks(4)
    assert [t.token_block.tokens[0] for t in taken] == [7, 5, 3, 1]


def test_priority_taking():
    pool = AvailableBlocks()                                   # reuse.rs:847-882: lower priority value goes first
    b1, b2 = _two_sequences(1, 2)
    for b in reversed(b2):
        pool.insert(b)
    for b in reversed(b1):
        pool.insert(b)
    taken = pool.take_blocks(4)
    assert [t.token_block.tokens[0] for t in taken] == [3, 1, 7, 5]


def test_priority_taking_after_update():
    pool = AvailableBlocks()                                   # reuse.rs:884-935
    b1, b2 = _two_sequences(1, 1)
    hashes2 = [b.token_block.sequence_hash for b in b2]
    for b in reversed(b2):
        pool.insert(b)
    for b in reversed(b1):
        pool.insert(b)
    pool.update_multiple([UpdateBlock(h, 2) for h in hashes2])
    taken = pool.take_blocks(4)
    assert [t.token_block.tokens[0] for t in taken] == [3, 1, 7, 5]


def test_reset_all_and_reset_some():
    pool = AvailableBlocks()                                   # reuse.rs:937-975
    b1, b2 = _two_sequences(1, 1)
    h1 = [b.token_block.sequence_hash for b in b1]
    h2 = [b.token_block.sequence_hash for b in b2]
    for b in reversed(b2):
        pool.insert(b)
    for b in reversed(b1):
        pool.insert(b)
    pool.reset_all()
    assert pool.match_blocks(h2) == [] and pool.available_blocks() == 4
    assert len(pool.take_blocks(10)) == 4                      # all four now sit in the un-initialised set
    pool = AvailableBlocks()                                   # reuse.rs:977-1030: reset only sequence 2
    b1, b2 = _two_sequences(1, 1)
    for b in reversed(b2):
        pool.insert(b)
    for b in reversed(b1):
        pool.insert(b)
    pool.reset(h2)
    assert pool.match_blocks(h2) == []
    m1 = pool.match_blocks(h1)
    assert len(m1) == 2


def test_duplicate_hash_and_zero_hash_go_to_the_uninitialised_set():
    pool = AvailableBlocks()                                   # reuse.rs:317-323
    a, b = make_blocks([1, 2], 2, 50)[0], make_blocks([1, 2], 2, 50)[0]
    pool.insert(a)
    pool.insert(b)                                             # same sequence hash: second copy is "un-initialised"
    pool.insert(KvBlock())                                     # default block: hash 0
    assert len(pool.uninitialized_set) == 2 and len(pool.lookup_map) == 1
    first = pool.take_blocks(1)[0]
    assert first.block is b                                    # un-initialised blocks are handed out first


def test_reserved_blocks_share_and_release():
    pool = AvailableBlocks()                                   # reserved.rs:66-150, tests :161-242
    blocks = make_blocks([1, 2, 3, 4], 2, 300)
    for b in blocks:
        pool.insert(b)
    reserved = ReservedBlocks(2)
    items = pool.match_blocks([300, 301])
    r0 = reserved.register(items[0])
    r1 = reserved.register(items[1])
    assert pool.available_blocks() == 0 and r0.inflight_count() == 1
    again = reserved.match_sequence_hashes([300, 301, 302])
    assert len(again) == 2 and r0.inflight_count() == 2        # shared, not copied
    # a second builder registers the same block: it gets the existing one, its duplicate goes back to the pool
    pool.insert(make_blocks([1, 2], 2, 300)[0])
    dup = pool.take_blocks(1)[0]
    r0b = reserved.register(dup)
    assert r0b.inflight_count() == 3 and pool.available_blocks() == 1
    for r in (r0, r0b, again[0]):
        r.drop()
    assert reserved.match_sequence_hashes([300]) == []         # last holder gone: back in the pool, out of the map
    assert pool.available_blocks() == 2
    with pytest.raises(KvError, match="Block size mismatch"):
        pool2 = AvailableBlocks()
        pool2.insert(make_blocks([9], 1, 900)[0])
        reserved.register(pool2.take_blocks(1)[0])
    r1.drop()
    again[1].drop()
    assert pool.available_blocks() == 3


@pytest.mark.gpu
def test_manager_plans_a_prefill_and_the_kernels_move_its_blocks(gpu):
    """tokens -> GPU chained hashes -> match in-flight / freed blocks -> take the rest -> block ids -> relayout kernel."""
    import ctypes as C

    from dynamo_exp_b200 import (CopyPlan, DType, KvBlockDetails, KvBlockStorage, KvContiguousBlocks, KvLayout,
                                 KvModelDetails, StorageType, TransferEngine)
    bs, n_cache = 4, 32
    mgr = KvStorageManager(bs)
    for i in range(n_cache):
        mgr.available_blocks.insert(KvBlock(block_id=i))
    toks = list(range(100, 100 + 4 * bs + 2))                  # 4 full blocks + 2 tokens
    matched = mgr.prepare_prefill_sequence(toks)
    assert len(matched.inflight_blocks) == 0 and len(matched.remaining_blocks) == 4
    assert len(matched.tail_block.tokens()) == 2
    off = mgr.prepare_prefill_offload(matched)
    ids = off.block_ids()
    assert len(ids) == 5 and len(set(ids)) == 5
    assert mgr.available_blocks.available_blocks() == n_cache - 5
    # "prefill" fills those cache slots; then the blocks become in-flight (registered) and a second, longer request
    # that shares the 4-block prefix finds them
    bd = KvBlockDetails(KvLayout.KvFirst, bs, KvModelDetails(2, 2, 8, DType.BF16))
    cache = KvBlockStorage.allocate(n_cache, bd, StorageType.Device(0))
    chunk = bd.bytes_per_token_block_per_layer() // 2
    host = []
    for l, layer in enumerate(cache.layers):
        data = np.arange(layer.storage_size(), dtype=np.uint32).astype(np.uint8) + np.uint8(l)
        gpu.check(gpu.lib().cuda_memcpy_sync(layer.get_pointer(), data.ctypes.data, data.nbytes))
        host.append(data)
    registered = [mgr.inflight_blocks.register(b) for b in off.complete_prefill_blocks]
    toks2 = toks[:4 * bs] + [7, 7, 7, 7, 9]
    m2 = mgr.prepare_prefill_sequence(toks2)
    assert [b.block_id for b in m2.inflight_blocks] == ids[:4]            # prefix hit on the in-flight blocks
    assert len(m2.remaining_blocks) == 1 and len(m2.tail_block.tokens()) == 1
    # move the prefix-hit blocks out to a contiguous slab with one launch, ids straight from the manager
    contig = KvContiguousBlocks(8, 2, chunk, StorageType.Device(0))
    eng = TransferEngine(8)
    eng.run(CopyPlan.contig(cache, contig, True), ids[:4], [0, 1, 2, 3], num_layers=2)
    eng.sync()
    got = np.empty(contig.storage_size(), dtype=np.uint8)
    gpu.check(gpu.lib().cuda_memcpy_sync(got.ctypes.data, contig.get_pointer(), got.nbytes))
    got = got.reshape(8, 2, 2, chunk)
    for slot, bid in enumerate(ids[:4]):
        for l in range(2):
            assert np.array_equal(got[slot, l], host[l].reshape(2, n_cache, chunk)[:, bid])
    # releasing every holder returns the blocks; a third identical request now matches them in the FREED pool
    for r in registered + m2.inflight_blocks:
        r.drop()
    assert mgr.available_blocks.available_blocks() == n_cache - 1          # only the first request's tail block is out
    m3 = mgr.prepare_prefill_sequence(toks)
    assert [b.block_id for b in m3.inflight_blocks] == ids[:4] and m3.remaining_blocks == []
    with pytest.raises(KvError, match="expected"):
        tiny = KvStorageManager(bs)
        tiny.prepare_prefill_offload(tiny.prepare_prefill_sequence(toks))
