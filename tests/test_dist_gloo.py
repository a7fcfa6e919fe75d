"""CPU, world_size 2 over gloo: the host-side logic of the multi-rank path — role split, TP target maps and the
metadata wire format — agrees across processes without any GPU (the data movement itself is covered by the gpu tests)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dynamo_exp_b200.connector import KvdpMetadata, split_roles, tp_targets


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, results):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        prefill, decode, pairs = split_roles(world)
        role = "prefill" if rank in prefill else "decode"
        # every rank publishes a (fake-pointer) metadata blob the way an engine would publish to etcd
        blob = KvdpMetadata(f"{role}-engine", [f"agent-{rank}".encode()], [[[0x1000 * (rank + 1), 0x2000 * (rank + 1)]]],
                            64 + rank).encode()
        gathered = [None] * world
        dist.all_gather_object(gathered, blob)
        metas = [KvdpMetadata.decode(b) for b in gathered]
        assert [m.num_blocks for m in metas] == [64 + r for r in range(world)]
        assert metas[rank].agent_metadata == [f"agent-{rank}".encode()]
        # both ranks derive the same pair list and the same TP maps
        t = torch.tensor([hash(tuple(pairs)) % (2 ** 31), sum(sum(tp_targets(4, 2, r)) for r in range(4)),
                          sum(sum(tp_targets(2, 8, r)) for r in range(2))], dtype=torch.int64)
        lo, hi = t.clone(), t.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        assert torch.equal(lo, hi)
        # the partner relation is symmetric: my partner's partner is me
        partner = dict(pairs).get(rank, next((p for p, d in pairs if d == rank), None))
        partners = [None] * world
        dist.all_gather_object(partners, partner)
        assert partners[partner] == rank
        results[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_two_ranks_agree_on_roles_targets_and_metadata():
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    results = mgr.dict()
    mp.spawn(_worker, args=(world, port, results), nprocs=world, join=True)
    assert dict(results) == {0: "ok", 1: "ok"}


def test_tp_target_maps():
    # reference direction (patch:1427-1471): rank * tp_multiplier + i
    assert tp_targets(1, 4, 0) == [0, 1, 2, 3]
    assert tp_targets(2, 4, 1) == [2, 3]
    assert tp_targets(2, 2, 1) == [1]
    # the inverse direction, new in this build: BASELINE config 3, TP4 prefill -> TP2 decode on 6 GPUs
    assert [tp_targets(4, 2, p) for p in range(4)] == [[0], [0], [1], [1]]
    # every decode rank is covered by exactly src_tp/dst_tp (or 1) senders
    for s, d in ((1, 2), (2, 8), (8, 2), (4, 4), (4, 1)):
        hits = [0] * d
        for p in range(s):
            for q in tp_targets(s, d, p):
                hits[q] += 1
        assert set(hits) == {max(1, s // d)}
    from dynamo_exp_b200.storage import KvError
    for bad in ((3, 4, 0), (4, 3, 0), (2, 4, 2), (0, 1, 0)):
        with pytest.raises(KvError):
            tp_targets(*bad)


def test_split_roles():
    assert split_roles(2) == ([0], [1], [(0, 1)])
    assert split_roles(4) == ([0, 1], [2, 3], [(0, 2), (1, 3)])
    assert split_roles(8)[2] == [(0, 4), (1, 5), (2, 6), (3, 7)]
    assert split_roles(8, interleaved=True) == ([0, 2, 4, 6], [1, 3, 5, 7], [(0, 1), (2, 3), (4, 5), (6, 7)])
    assert split_roles(3, interleaved=True) == split_roles(3)           # odd worlds keep the plain split
    p, d, pairs = split_roles(6)
    assert (p, d) == ([0, 1, 2], [3, 4, 5])


def test_range_helpers_match_the_reference_docstrings():
    from dynamo_exp_b200.connector import get_ranges, get_same_length_ranges
    # patch:1202-1203
    assert get_ranges([0, 1, 2, 4, 5, 6]) == [[0, 2], [4, 6]]
    assert get_ranges([]) == [] and get_ranges([7]) == [[7, 7]] and get_ranges([3, 2, 1]) == [[3, 3], [2, 2], [1, 1]]
    # patch:1248-1250
    assert get_same_length_ranges([[0, 2], [4, 8]], [[1, 3], [5, 7], [9, 10]]) == \
        ([[0, 2], [4, 6], [7, 8]], [[1, 3], [5, 7], [9, 10]])
    s, d = get_same_length_ranges([[0, 9]], [[0, 3], [10, 15]])
    assert s == [[0, 3], [4, 9]] and d == [[0, 3], [10, 15]]
    assert all(a[1] - a[0] == b[1] - b[0] for a, b in zip(s, d))
