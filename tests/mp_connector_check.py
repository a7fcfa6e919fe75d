"""Helper for tests/test_gpu_connector.py::test_two_process_ipc_transfer (also runnable by hand).

Parent = decode engine (TP1), child = prefill engine (TP1), both on cuda:0 unless KVDP_MP_DEVICES="a,b" names two
devices. They exchange KvdpMetadata through files, the child write_blocks() into the parent's cache over CUDA IPC and
then read_blocks() them back into a second local cache; the parent polls get_new_notifs() and checks the bytes."""
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from dynamo_exp_b200 import _lib  # noqa: E402
from dynamo_exp_b200.connector import DynamoKvdpConnector, KvdpMetadata  # noqa: E402
from dynamo_exp_b200.remote_prefill import (RemotePrefillParams, TransferTracker, collect_transfer_counters,  # noqa: E402
                                            execute_memory_transfers, plan_memory_transfers)
from tests.util import DevBuf, payload  # noqa: E402

L, H, D, BS, E, NB = 2, 4, 16, 8, 2, 8
LAYER_BYTES = 2 * NB * BS * H * D * E


def wait_for(path, timeout=120):
    t0 = time.time()
    while not os.path.exists(path):
        if time.time() - t0 > timeout:
            raise TimeoutError(path)
        time.sleep(0.01)
    time.sleep(0.02)
    return open(path, "rb").read()


def publish(path, data):
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(data)
    os.rename(tmp, path)


def main():
    role = sys.argv[1] if len(sys.argv) > 1 else "parent"
    devs = [int(x) for x in os.environ.get("KVDP_MP_DEVICES", "0,0").split(",")]
    if role == "parent":
        d = tempfile.mkdtemp(prefix="kvdp_mp_")
        child = subprocess.Popen([sys.executable, os.path.abspath(__file__), "child", d])
        dev = devs[0]
        _lib.require_gpu()
        bufs = [DevBuf(_lib, LAYER_BYTES, dev) for _ in range(L)]
        c = DynamoKvdpConnector("decode", 0, 1, device=dev)
        c.register_kv_caches([b.ptr for b in bufs], shape=(2, NB, BS, H, D), element_size=E)
        meta = KvdpMetadata("decode", [c.get_agent_metadata()], [c.kv_caches_base_addr["decode"]], NB)
        publish(os.path.join(d, "decode.meta"), meta.encode())
        tracker = TransferTracker(tensor_parallel_size=1)     # decode engine: one notification per TP worker
        t0 = time.time()
        while "req-1" not in tracker.finished_prefills:
            tracker.update(*collect_transfer_counters([c]))
            if time.time() - t0 > 120:
                raise TimeoutError("no notification")
            time.sleep(0.005)
        src = [payload(LAYER_BYTES, 300 + l).reshape(2, NB, -1) for l in range(L)]
        for l in range(L):
            got = bufs[l].download().reshape(2, NB, -1)
            assert np.array_equal(got[:, [1, 6, 2]], src[l][:, [5, 0, 7]]), l
        publish(os.path.join(d, "parent.ok"), b"1")
        rc = child.wait(timeout=120)
        assert rc == 0, rc
        c.shutdown()
        print("IPC OK")
    else:
        d = sys.argv[2]
        dev = devs[1]
        _lib.require_gpu()
        src_h = [payload(LAYER_BYTES, 300 + l) for l in range(L)]
        src = [DevBuf(_lib, LAYER_BYTES, dev) for _ in range(L)]
        for b, a in zip(src, src_h):
            b.upload(a)
        c = DynamoKvdpConnector("prefill", 0, 1, device=dev)
        c.register_kv_caches([b.ptr for b in src], shape=(2, NB, BS, H, D), element_size=E)
        meta = KvdpMetadata.decode(wait_for(os.path.join(d, "decode.meta")))
        c.add_remote_agent("decode", meta.agent_metadata, 1, meta.kv_caches_base_addr, meta.num_blocks)
        # the engine's plan for a remote-decode request whose blocks were all computed here (patch:3188-3221) ...
        params = RemotePrefillParams(is_remote_decode=True, decode_block_ids=[1, 6, 2], decode_engine_id="decode")
        reqs = plan_memory_transfers("req-1", [5, 0, 7], [], [], params)
        ran = []
        execute_memory_transfers(c, reqs, run_model=lambda: ran.append(1))        # ... run as a worker step
        c.sync()
        tracker = TransferTracker(tensor_parallel_size=1)
        tracker.update(*collect_transfer_counters([c]))
        assert ran == [1] and tracker.finished_transfers == {"req-1"}
        wait_for(os.path.join(d, "parent.ok"))
        # pull the same blocks back into a second local cache and compare with the originals
        back = [DevBuf(_lib, LAYER_BYTES, dev) for _ in range(L)]
        c2 = DynamoKvdpConnector("prefill2", 0, 1, device=dev)
        c2.register_kv_caches([b.ptr for b in back], shape=(2, NB, BS, H, D), element_size=E)
        c2.add_remote_agent("decode", meta.agent_metadata, 1, meta.kv_caches_base_addr, meta.num_blocks)
        c2.read_blocks([3, 4, 0], [0, 0, 0], [1, 6, 2], "decode")
        for l in range(L):
            got = back[l].download().reshape(2, NB, -1)
            want = src_h[l].reshape(2, NB, -1)
            assert np.array_equal(got[:, [3, 4, 0]], want[:, [5, 0, 7]]), l
        c.shutdown()
        c2.shutdown()


if __name__ == "__main__":
    main()
