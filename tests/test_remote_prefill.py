"""Host logic of the request / completion protocol around the connector (dynamo_exp_b200/remote_prefill.py), against the
behaviour of the reference's patched vLLM engine and worker (patch:2991-2992, 3159-3261, 4199-4256, 4528-4539, 4779-4800)."""
import msgpack

from dynamo_exp_b200.remote_prefill import (MemoryOpType, MemoryTransferRequest, RemotePrefillParams,
                                            RemotePrefillRequest, TransferTracker, collect_transfer_counters,
                                            execute_memory_transfers, make_remote_prefill_request,
                                            plan_memory_transfers)


class FakeConnector:
    def __init__(self, notifs=None, done=None):
        self.calls, self._notifs, self._done = [], notifs or {}, done or []

    def read_blocks(self, local, staging, remote, engine):
        self.calls.append(("read", list(local), list(staging), list(remote), engine))

    def write_blocks(self, local, staging, remote, engine, msg):
        self.calls.append(("write", list(local), list(staging), list(remote), engine, msg))

    def get_new_notifs(self):
        out, self._notifs = self._notifs, {}
        return out

    def get_done_tranfers(self):
        out, self._done = self._done, []
        return out


def test_wire_formats_round_trip_and_match_msgspec_layouts():
    r = RemotePrefillRequest("eng-d", "req-7", [1, 2, 3], {"temperature": 0.0}, [4, 9, 2], [4])
    assert RemotePrefillRequest.decode(r.encode()) == r
    assert set(msgpack.unpackb(r.encode())) == {"engine_id", "request_id", "prompt_token_ids", "sampling_params",
                                                 "block_ids", "computed_block_ids"}                  # a map
    m = MemoryTransferRequest("req-7", [1, 2], [], [8, 9], "eng-d", "req-7", MemoryOpType.WRITE)
    assert MemoryTransferRequest.decode(m.encode()) == m
    assert msgpack.unpackb(m.encode()) == ["req-7", [1, 2], [], [8, 9], "eng-d", "req-7", "WRITE"]   # array_like


def test_no_remote_prefill_request_when_everything_is_cached():
    assert make_remote_prefill_request("e", "r", [1, 2], {}, [5, 6], [5, 6]) is None                  # patch:3165-3168
    got = make_remote_prefill_request("e", "r", [1, 2], {}, [5, 6, 7], [5])
    assert got.block_ids == [5, 6, 7] and got.computed_block_ids == [5] and got.prompt_token_ids == [1, 2]


def test_plan_memory_transfers_like_llm_engine_step():
    """patch:3188-3221: the first `len(computed)` blocks are READ from the decode engine, the rest WRITTEN to it."""
    p = RemotePrefillParams(is_remote_decode=True, decode_block_ids=[30, 31, 32, 33, 34], decode_engine_id="dec")
    reqs = plan_memory_transfers("r1", [10, 11, 12, 13, 14], [20, 21, 22, 23, 24], [10, 11], p)
    assert reqs == [
        MemoryTransferRequest("r1", [10, 11], [20, 21], [30, 31], "dec", "r1", MemoryOpType.READ),
        MemoryTransferRequest("r1", [12, 13, 14], [22, 23, 24], [32, 33, 34], "dec", "r1", MemoryOpType.WRITE)]
    # nothing computed locally: only the WRITE (which may be empty: the notification alone is sent, patch:1373-1377)
    reqs = plan_memory_transfers("r2", [1, 2], [], [], p)
    assert [r.op_type for r in reqs] == [MemoryOpType.WRITE] and reqs[0].remote_block_ids == [30, 31, 32, 33, 34]
    everything = plan_memory_transfers("r3", [1, 2], [], [1, 2], RemotePrefillParams(
        is_remote_decode=True, decode_block_ids=[7, 8], decode_engine_id="dec"))
    assert [(r.op_type, r.local_block_ids) for r in everything] == [(MemoryOpType.READ, [1, 2]), (MemoryOpType.WRITE, [])]
    assert plan_memory_transfers("r4", [1], [], [], RemotePrefillParams()) == []                      # not a remote decode


def test_worker_step_reads_before_the_model_and_writes_after():
    c = FakeConnector()
    order = []
    reqs = [MemoryTransferRequest("a", [3], [], [4], "dec", "a", MemoryOpType.WRITE),
            MemoryTransferRequest("a", [1], [], [2], "dec", "a", MemoryOpType.READ)]
    execute_memory_transfers(c, reqs, run_model=lambda: order.append(len(c.calls)))
    assert [x[0] for x in c.calls] == ["read", "write"] and order == [1]
    assert c.calls[1] == ("write", [3], [], [4], "dec", "a")


def test_counters_and_engine_tracker_need_one_count_per_tp_worker():
    """patch:2991-2992, 3251-3261, 4779-4800 with tensor_parallel_size = 2."""
    w0 = FakeConnector({"agent-p0": ["r1", b"r2"]}, ["r9"])
    w1 = FakeConnector({"agent-p1": ["r1"]}, ["ignored: only the driver worker's done list is counted"])
    notif, done = collect_transfer_counters([w0, w1])
    assert notif == {"r1": 2, "r2": 1} and done == {"r9": 1}
    t = TransferTracker(tensor_parallel_size=2)
    t.update(notif, done)
    assert t.finished_prefills == {"r1"} and t.finished_transfers == set()
    assert t.pending() == ({"r2": -1}, {"r9": -1})
    t.update({"r2": 1}, {"r9": 1})
    assert t.finished_prefills == {"r1", "r2"} and t.finished_transfers == {"r9"} and t.pending() == ({}, {})


def test_metadata_store_like_nixl_metadata_store(tmp_path):
    """examples/llm/utils/nixl.py:59-116: put under <namespace>/nixl_metadata/<engine>, cached get, a clashing put only
    warns, an unknown engine raises."""
    import warnings

    import pytest

    from dynamo_exp_b200.connector import DirKvClient, KvdpMetadata, KvdpMetadataStore
    from dynamo_exp_b200.storage import KvError
    client = DirKvClient(str(tmp_path))
    a, b = KvdpMetadataStore("dynamo", client), KvdpMetadataStore("dynamo", client)
    meta = KvdpMetadata("eng-1", [b"\x00\x01handle", b"\xff"], [[[4096, 8192]], [[12288, 16384]]], 77)
    a.put("eng-1", meta)
    a.put("eng-1", meta)                                            # same value: validated, no complaint
    got = b.get("eng-1")
    assert got == meta and b.get("eng-1") is got                     # second lookup comes from the cache
    assert [i["key"] for i in client.kv_get_prefix("dynamo/nixl_metadata/")] == ["dynamo/nixl_metadata/eng-1"]
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        a.put("eng-1", KvdpMetadata("eng-1", [b"other"], [], 1))
    assert any("different metadata" in str(x.message) for x in w)
    assert KvdpMetadataStore("dynamo", client).get("eng-1") == meta  # the first value stays
    with pytest.raises(KvError):
        b.get("eng-2")
    with pytest.raises(KvError):
        KvdpMetadataStore("dynamo", None)
