"""End-to-end walk through the whole hot path with two OS processes on one GPU (also runnable by hand):

    decode engine (parent)                           prefill engine (child)
    ----------------------                           ----------------------
    registers its KV cache, publishes metadata  -->  adds the remote agent
    plans blocks for two requests (KvStorageManager)
    sends RemotePrefillRequest (msgpack file)   -->  hashes the prompt on the GPU, plans its own blocks, "computes" KV
                                                     (synthetic payload keyed by the block's sequence hash),
                                                     plan_memory_transfers + execute_memory_transfers: NVLink/IPC push
                                                     publishes KV stored events through the engine-facing C ABI
    TransferTracker sees the notification       <--  in-kernel completion flag + mailbox
    checks the received blocks byte for byte
    request 2 shares request 1's first blocks:       its manager finds them in flight (prefix hit): only the new blocks
                                                     are computed and written
    router side: the events (RouterEvent JSON read back from the child's outbox) are applied to a prefix index;
    find_matches_for_request(prompt 2) scores the prefill worker with the shared prefix length.

Nothing here waits on another process from inside a kernel: host-side polling only."""
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from dynamo_exp_b200 import _lib  # noqa: E402
from dynamo_exp_b200.connector import (DirKvClient, DynamoKvdpConnector, KvdpMetadata, KvdpMetadataStore)  # noqa: E402
from dynamo_exp_b200.indexer import KvIndexer, RouterEvent  # noqa: E402
from dynamo_exp_b200.publisher import KvEventPublisher  # noqa: E402
from dynamo_exp_b200.remote_prefill import (RemotePrefillParams, RemotePrefillRequest, TransferTracker,  # noqa: E402
                                            collect_transfer_counters, execute_memory_transfers,
                                            make_remote_prefill_request, plan_memory_transfers)
from dynamo_exp_b200.reuse import KvBlock, KvStorageManager  # noqa: E402
from dynamo_exp_b200.tokens import TokenBlockSequence, XXH3_SEED  # noqa: E402
from tests.util import DevBuf, mix64  # noqa: E402

L, H, D, BS, E, NB = 3, 4, 16, 8, 2, 24
CHUNK = BS * H * D * E
LAYER_BYTES = 2 * NB * CHUNK
PREFILL_WORKER = 4242


def wait_for(path, timeout=120):
    t0 = time.time()
    while not os.path.exists(path):
        if time.time() - t0 > timeout:
            raise TimeoutError(path)
        time.sleep(0.005)
    time.sleep(0.01)
    with open(path, "rb") as f:
        return f.read()


def publish(path, data):
    with open(path + ".tmp", "wb") as f:
        f.write(data)
    os.rename(path + ".tmp", path)


def kv_payload(sequence_hash, layer, kv):
    """What the 'model' writes for one block: a function of the block's chained hash, so that the decode side can
    recompute the expected bytes from the prompt alone."""
    seed = (int(sequence_hash) ^ (layer * 0x9E37 + kv * 0x51)) & 0xFFFFFFFFFFFFFFFF
    idx = np.arange(CHUNK // 8, dtype=np.uint64)
    with np.errstate(over="ignore"):
        return mix64(idx + np.uint64(seed)).view(np.uint8)


def prompts():
    p1 = [int(x) for x in np.arange(1000, 1000 + 5 * BS + 3)]          # 5 full blocks + 3 tokens
    p2 = p1[:3 * BS] + [int(x) for x in np.arange(7000, 7000 + 2 * BS + 1)]   # shares 3 blocks, then 2 new ones
    return p1, p2


def decode_main(d):
    _lib.require_gpu()
    child = subprocess.Popen([sys.executable, os.path.abspath(__file__), "prefill", d])
    bufs = [DevBuf(_lib, LAYER_BYTES) for _ in range(L)]
    conn = DynamoKvdpConnector("decode", 0, 1, device=0)
    conn.register_kv_caches([b.ptr for b in bufs], shape=(2, NB, BS, H, D), element_size=E)
    store = KvdpMetadataStore("dynamo", DirKvClient(os.path.join(d, "etcd")))
    store.put("decode", KvdpMetadata("decode", [conn.get_agent_metadata()], [conn.kv_caches_base_addr["decode"]], NB))
    mgr = KvStorageManager(BS)
    for i in range(NB):
        mgr.available_blocks.insert(KvBlock(block_id=i))
    tracker = TransferTracker(tensor_parallel_size=1)
    received = {}
    for k, prompt in enumerate(prompts()):
        req_id = "req-%d" % (k + 1)
        offload = mgr.prepare_prefill_offload(mgr.prepare_prefill_sequence(prompt))
        block_table = offload.block_ids()                      # decode-side slots, sequence order (tail block last)
        req = make_remote_prefill_request("decode", req_id, prompt, {"max_tokens": 1}, block_table, [])
        publish(os.path.join(d, req_id + ".req"), req.encode())
        t0 = time.time()
        while req_id not in tracker.finished_prefills:
            tracker.update(*collect_transfer_counters([conn]))
            if time.time() - t0 > 120:
                raise TimeoutError("no notification for " + req_id)
            time.sleep(0.002)
        seq = TokenBlockSequence(prompt, BS, XXH3_SEED)
        for i, tb in enumerate(seq.blocks()):
            for layer in range(L):
                got = bufs[layer].download().reshape(2, NB, CHUNK)
                for kv in (0, 1):
                    assert np.array_equal(got[kv, block_table[i]], kv_payload(tb.sequence_hash, layer, kv)), (req_id, i)
        received[req_id] = block_table
        publish(os.path.join(d, req_id + ".ok"), b"1")
    # router: replay the prefill worker's events into a prefix index and score the second prompt
    events = json.loads(wait_for(os.path.join(d, "events.json")))
    index = KvIndexer(BS, max_nodes=1024)
    for js in events:
        index.apply_event(RouterEvent.from_json(js))
    p1, p2 = prompts()
    assert index.find_matches_for_request(p1).scores == {PREFILL_WORKER: 5}
    assert index.find_matches_for_request(p2).scores == {PREFILL_WORKER: 5}        # 3 shared + 2 of its own
    assert index.find_matches_for_request(p2[:3 * BS] + [1] * BS).scores == {PREFILL_WORKER: 3}
    rc = child.wait(timeout=120)
    assert rc == 0, rc
    conn.shutdown()
    index.tree.close()
    print("DISAGG FLOW OK")


def prefill_main(d):
    _lib.require_gpu()
    cache = [DevBuf(_lib, LAYER_BYTES) for _ in range(L)]
    conn = DynamoKvdpConnector("prefill", 0, 1, device=0)
    conn.register_kv_caches([b.ptr for b in cache], shape=(2, NB, BS, H, D), element_size=E)
    store = KvdpMetadataStore("dynamo", DirKvClient(os.path.join(d, "etcd")))
    t0 = time.time()
    while True:
        try:
            meta = store.get("decode")
            break
        except Exception:
            if time.time() - t0 > 120:
                raise
            time.sleep(0.01)
    conn.add_remote_agent("decode", meta.agent_metadata, 1, meta.kv_caches_base_addr, meta.num_blocks)
    mgr = KvStorageManager(BS)
    for i in range(NB):
        mgr.available_blocks.insert(KvBlock(block_id=i))
    pub = KvEventPublisher("dynamo/VllmWorker", PREFILL_WORKER, BS)
    tracker = TransferTracker(tensor_parallel_size=1)
    held, computed_counts, event_id = [], [], 0
    for k in range(2):
        req_id = "req-%d" % (k + 1)
        req = RemotePrefillRequest.decode(wait_for(os.path.join(d, req_id + ".req")))
        matched = mgr.prepare_prefill_sequence(req.prompt_token_ids)
        n_hit = len(matched.inflight_blocks)                   # blocks this engine already holds (prefix cache hit)
        new_token_blocks = list(matched.remaining_blocks)
        offload = mgr.prepare_prefill_offload(matched)
        local_ids = offload.block_ids()
        # "model execution": fill the blocks that were not cached
        for j, tb in enumerate(new_token_blocks):
            slot = local_ids[n_hit + j]
            for layer in range(L):
                for kv in (0, 1):
                    cache[layer].upload(kv_payload(tb.sequence_hash, layer, kv), offset=(kv * NB + slot) * CHUNK)
        computed_counts.append(len(new_token_blocks))
        # the decode engine has none of these blocks (computed_block_ids empty), so everything is WRITTEN; the partial
        # tail block is not transferred (the decode engine recomputes the last tokens itself)
        n_full = n_hit + len(new_token_blocks)
        params = RemotePrefillParams(is_remote_decode=True, decode_block_ids=req.block_ids[:n_full],
                                     decode_engine_id=req.engine_id)
        reqs = plan_memory_transfers(req.request_id, local_ids[:n_full], [], req.computed_block_ids, params)
        execute_memory_transfers(conn, reqs)
        conn.sync()
        tracker.update(*collect_transfer_counters([conn]))
        assert req.request_id in tracker.finished_transfers
        # KV events for the blocks that were newly stored, chained to the last cached block
        if new_token_blocks:
            toks = [t for tb in new_token_blocks for t in tb.tokens]
            parent = offload.inflight_blocks[-1].token_block.sequence_hash if n_hit else None
            pub.publish_stored(event_id, toks, [BS] * len(new_token_blocks),
                               [tb.sequence_hash for tb in new_token_blocks], 0, parent)
            event_id += 1
        # blocks stay registered (in flight) so that the next request can hit them
        held += [mgr.inflight_blocks.register(b) for b in offload.complete_prefill_blocks] + offload.inflight_blocks
        wait_for(os.path.join(d, req_id + ".ok"))
    assert computed_counts == [5, 2], computed_counts          # request 2 reused 3 blocks
    pub.flush()
    out = []
    while True:
        ev = pub.poll()
        if ev is None:
            break
        out.append(ev.to_json())
    publish(os.path.join(d, "events.json"), json.dumps(out).encode())
    pub.shutdown()
    conn.shutdown()


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "prefill":
        prefill_main(sys.argv[2])
    else:
        decode_main(tempfile.mkdtemp(prefix="kvdp_flow_"))
