"""KV event publishing through libkvdp.so's engine-facing ABI (include/kvdp.h §I).

`KvEventPublisher` keeps the method names and argument order of the reference's Python binding
(lib/bindings/python/rust/llm/kv.rs:123-180: `publish_stored(event_id, token_ids, num_block_tokens, block_hashes,
lora_id, parent_hash=None)`, `publish_removed(event_id, block_hashes)`) and calls the same five C symbols the patched
vLLM calls through ctypes (lib/bindings/c/src/lib.rs:50-311). Like the reference's `KV_PUB` OnceCell the publisher is
one per process: a second `KvEventPublisher` shares the first one's worker id and block size until `shutdown()`.

Short blocks follow the C path (the batch ends at the first block whose token count is not kv_block_size, every time,
lib.rs:191-210); the pyo3 path stops checking after three warnings and would then index past short blocks
(kv.rs:212-221), which is not reproduced.

The reference publishes on NATS; here events go to a Python callback, straight into a `RadixTree`, or wait in the
library's outbox for `poll()`.
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from .indexer import RadixTree, RouterEvent
from .storage import KvError

_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_char_p, C.POINTER(C.c_char), C.c_size_t)


def _names(component) -> Tuple[str, str]:
    """Accepts "namespace/component", a (namespace, component) pair, or an object with the binding's accessors."""
    if isinstance(component, str):
        ns, _, name = component.partition("/")
        if not ns or not name:
            raise KvError('component must look like "namespace/component"')
        return ns, name
    if isinstance(component, (tuple, list)) and len(component) == 2:
        return str(component[0]), str(component[1])
    return str(component.namespace), str(component.name)


class KvEventPublisher:
    def __init__(self, component, worker_id: int, kv_block_size: int):
        ns, name = _names(component)
        self._L = _lib.lib()
        self.worker_id = int(worker_id)
        self.kv_block_size = int(kv_block_size)
        self.subject = f"namespace.{ns}.component.{name}.kv_events"
        if self._L.dynamo_llm_init(ns.encode(), name.encode(), self.worker_id, self.kv_block_size) != 0:
            raise KvError("dynamo_llm_init failed")
        self._cb = None            # keeps the ctypes thunk alive while the library holds it

    # ------------------------------------------------------------------ the binding's two methods
    def publish_stored(self, event_id: int, token_ids: Sequence[int], num_block_tokens: Sequence[int],
                       block_hashes: Sequence[int], lora_id: int, parent_hash: Optional[int] = None) -> None:
        tok = np.ascontiguousarray(token_ids, dtype=np.uint32)
        n = min(len(num_block_tokens), len(block_hashes))      # the binding zips the two lists (kv.rs:212)
        counts = (C.c_size_t * max(n, 1))(*[int(x) for x in num_block_tokens[:n]])
        ids = (C.c_uint64 * max(n, 1))(*[int(x) for x in block_hashes[:n]])
        need, bs = 0, self.kv_block_size
        for c in num_block_tokens[:n]:
            if int(c) != bs:
                break
            need += bs
        if need > tok.size:
            raise KvError("token_ids is shorter than the blocks it is said to hold")
        parent = None if parent_hash is None else (C.c_uint64 * 1)(int(parent_hash))
        rc = self._L.dynamo_kv_event_publish_stored(int(event_id), tok.ctypes.data_as(C.POINTER(C.c_uint32)), counts, ids,
                                                    n, parent, int(lora_id))
        if rc != 0:
            raise KvError("dynamo_kv_event_publish_stored failed")

    def publish_removed(self, event_id: int, block_hashes: Sequence[int]) -> None:
        n = len(block_hashes)
        ids = (C.c_uint64 * max(n, 1))(*[int(x) for x in block_hashes])
        if self._L.dynamo_kv_event_publish_removed(int(event_id), ids, n) != 0:
            raise KvError("dynamo_kv_event_publish_removed failed")

    # ------------------------------------------------------------------ sinks
    def on_event(self, fn: Optional[Callable[[str, str], None]]) -> None:
        """`fn(subject, router_event_json)` runs on the publisher thread for every event; None removes it."""
        if fn is None:
            _lib.check(self._L.kvdp_event_sink_callback(None, None))
            self._cb = None
            return

        def thunk(_user, subject, data, length):
            fn(subject.decode(), C.string_at(data, length).decode())

        cb = _CB(thunk)
        _lib.check(self._L.kvdp_event_sink_callback(C.cast(cb, C.c_void_p), None))
        self._cb = cb

    def attach_index(self, tree: Optional[RadixTree]) -> None:
        """Apply every event to `tree` under this publisher's worker id (what the router's subscriber task does with
        the events it receives, kv_router.rs:80-110)."""
        if tree is None:
            _lib.check(self._L.kvdp_event_sink_index(None, 0))
            return
        _lib.check(self._L.kvdp_event_sink_index(tree._h, tree._slot_of(self.worker_id, True)))

    def flush(self) -> None:
        _lib.check(self._L.kvdp_event_flush())

    def poll(self) -> Optional[RouterEvent]:
        """Next event from the library's outbox (used when neither sink is set), or None."""
        raw = poll_raw()
        return None if raw is None else RouterEvent.from_json(raw)

    def stats(self) -> dict:
        names = ["submitted", "delivered", "hash_batches", "blocks_hashed", "short_blocks", "parent_missing",
                 "failed", "outbox_dropped", "outbox_len"]
        return {k: int(self._L.kvdp_event_stat(i)) for i, k in enumerate(names)}

    def shutdown(self) -> None:
        self._L.dynamo_llm_shutdown()


def poll_raw() -> Optional[str]:
    L = _lib.lib()
    n = C.c_size_t(0)
    buf = C.create_string_buffer(4096)
    rc = L.kvdp_event_poll(buf, len(buf), C.byref(n))
    if rc == _lib.KVDP_ERANGE:
        buf = C.create_string_buffer(n.value)
        rc = L.kvdp_event_poll(buf, len(buf), C.byref(n))
    if rc < 0:
        _lib.check(rc)
    return buf.raw[:n.value].decode() if rc == 1 else None
