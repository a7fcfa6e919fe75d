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


def test_reshard_leg_configurations_and_canonical_data_are_deterministic():
    """Host-side pieces of the multi-GPU reshard leg (bench_reshard.py), on the CPU: the TP configurations each GPU count
    runs, the senders every decode rank waits for, and the closed-form canonical cache every rank regenerates."""
    import types

    import torch

    import bench_reshard as br
    from dynamo_exp_b200.connector import tp_targets
    assert br.configs_for(8) == [(4, 2), (4, 4)] and br.configs_for(4) == [(2, 2), (2, 1)] and br.configs_for(2) == [(1, 1)]
    for n, cfgs in ((2, br.configs_for(2)), (4, br.configs_for(4)), (8, br.configs_for(8))):
        for p_tp, d_tp in cfgs:
            assert p_tp + d_tp <= n
            cover = {}
            for p in range(p_tp):
                for q in tp_targets(p_tp, d_tp, p):
                    cover.setdefault(q, []).append(p)
            assert sorted(cover) == list(range(d_tp))                       # every decode rank hears from someone
            assert all(len(v) == max(1, p_tp // d_tp) for v in cover.values())
    cpu = types.SimpleNamespace(arange=lambda n, dtype, device: torch.arange(n, dtype=dtype), int64=torch.int64,
                                int16=torch.int16)
    saved = (br.BS, br.D70)
    try:
        br.BS, br.D70 = 4, 8                                                # small geometry, same arithmetic
        a, b, c = br.canonical_layer(cpu, 3, 5), br.canonical_layer(cpu, 3, 5), br.canonical_layer(cpu, 4, 5)
    finally:
        br.BS, br.D70 = saved
    assert a.shape == (2, 5, 4, 8, 8) and torch.equal(a, b) and not torch.equal(a, c)
    assert len(set(a.flatten().tolist())) > a.numel() // 2                  # values vary: a misrouted chunk cannot pass


def test_reference_connector_range_helpers_agree_with_the_restatement():
    """oracle/ref_triton.py replays the reference's range splitting around its Triton kernel; its helpers must agree with
    the product's get_ranges / get_same_length_ranges on the reference's docstring examples and on random id lists."""
    import numpy as np

    from dynamo_exp_b200.connector import get_ranges, get_same_length_ranges
    from oracle import ref_triton
    assert ref_triton._ranges([0, 1, 2, 4, 5, 6]) == get_ranges([0, 1, 2, 4, 5, 6]) == [[0, 2], [4, 6]]
    s, d = [[0, 2], [4, 8]], [[1, 3], [5, 7], [9, 10]]
    assert ref_triton._same_length(s, d) == tuple(get_same_length_ranges(s, d))
    rng = np.random.default_rng(3)
    for _ in range(50):
        n = int(rng.integers(1, 20))
        a = sorted(int(x) for x in rng.choice(60, size=n, replace=False))
        b = sorted(int(x) for x in rng.choice(60, size=n, replace=False))
        assert ref_triton._same_length(ref_triton._ranges(a), ref_triton._ranges(b)) == \
            tuple(get_same_length_ranges(get_ranges(a), get_ranges(b)))
