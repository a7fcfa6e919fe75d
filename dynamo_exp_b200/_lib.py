"""ctypes binding of libkvdp.so — the only way Python reaches the data plane.

There is no CPU fallback: if the shared library is missing or does not load, importing this module's `lib()` raises.
The prototypes below are transcribed from include/kvdp.h; tests/test_abi.py checks that every symbol the header
declares is exported and bound here.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libkvdp.so")

KVDP_OK = 0
KVDP_EINVAL = -22
KVDP_ERANGE = -34
KVDP_ESTATE = -71
KVDP_EALIGN = -72
KVDP_EDUP = -73
KVDP_ENODEV = -19

XFER_CHECK_DST_UNIQUE = 1
HASH_TWIN_FIRST = 1
HASH_NOALLOC_MAX_REQUESTS = 1 << 22
IPC_HANDLE_BYTES = 64


class KvdpError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"kvdp error {code}: {message}")
        self.code = code
        self.message = message


class PlanDesc(C.Structure):
    _fields_ = [
        ("num_layers", C.c_int32),
        ("row_groups", C.c_int32),
        ("prefix_dim", C.c_int32),
        ("rows", C.c_int32),
        ("run_bytes", C.c_int64),
        ("src_prefix_stride", C.c_int64),
        ("src_block_stride", C.c_int64),
        ("dst_prefix_stride", C.c_int64),
        ("dst_block_stride", C.c_int64),
        ("dst_row_stride", C.c_int64),
        ("dst_col_offset", C.c_int64),
        ("src_row_stride", C.c_int64),
        ("src_col_offset", C.c_int64),
        ("src_blocks", C.c_int64),
        ("dst_blocks", C.c_int64),
        ("src_layer_ptrs", C.c_void_p),
        ("dst_layer_ptrs", C.c_void_p),
        ("device", C.c_int32),
    ]


_vp, _i, _i32, _i64, _u32, _u64, _sz = C.c_void_p, C.c_int, C.c_int32, C.c_int64, C.c_uint32, C.c_uint64, C.c_size_t
_pp = C.POINTER(C.c_void_p)

# name -> (restype, argtypes). Order follows include/kvdp.h.
PROTOTYPES = {
    "kvdp_last_error": (C.c_char_p, []),
    "kvdp_version": (_i, []),
    "kvdp_set_option": (_i, [C.c_char_p, _i64]),
    "kvdp_get_option": (_i64, [C.c_char_p]),
    # legacy
    "copy_blocks_3d": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i]),
    "copy_blocks_launcher_3d": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "copy_blocks_memcpy_3d": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "copy_stream_create": (_i, [_pp, _i, _i]),
    "copy_stream_destroy": (_i, [_vp]),
    "copy_stream_prepare_block_ids": (_i, [_vp, _vp, _vp, _i]),
    "copy_stream_launch": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i]),
    "copy_stream_memcpy": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i]),
    "copy_stream_sync": (_i, [_vp]),
    "copy_stream_scatter": (_i, [_vp, _vp, _vp, _vp, _u32, _u32, _u32, _u32, _u32]),
    "cuda_malloc_host": (_i, [_pp, _sz]),
    "cuda_free_host": (_i, [_vp]),
    "cuda_memcpy_async": (_i, [_vp, _vp, _sz, _vp]),
    "cuda_memcpy_sync": (_i, [_vp, _vp, _sz]),
    # v2
    "kvdp_device_count": (_i, [C.POINTER(_i)]),
    "kvdp_device_alloc": (_i, [_pp, _sz, _i]),
    "kvdp_device_free": (_i, [_vp]),
    "kvdp_memset": (_i, [_vp, _i, _sz]),
    "kvdp_pointer_kind": (_i, [_vp, C.POINTER(_i), C.POINTER(_i)]),
    "kvdp_plan_create": (_i, [_pp, C.POINTER(PlanDesc)]),
    "kvdp_plan_destroy": (_i, [_vp]),
    "kvdp_plan_bytes_per_pair": (_i64, [_vp]),
    "kvdp_plan_block_copy": (_i, [_pp, _vp, _vp, _i32, _i32, _i64, _i64, _i64, _i32, _i32]),
    "kvdp_plan_contig": (_i, [_pp, _vp, _i32, _i32, _i64, _i64, _vp, _i64, _i32, _i32]),
    "kvdp_plan_reshard": (_i, [_pp, _vp, _vp, _i32, _i64, _i64, _i32, _i64, _i32, _i32, _i32, _i32]),
    "kvdp_xfer_create": (_i, [_pp, _i32, _i32]),
    "kvdp_xfer_destroy": (_i, [_vp]),
    "kvdp_xfer_stream": (_vp, [_vp]),
    "kvdp_xfer_run": (_i, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _u32, _vp, _u32]),
    "kvdp_xfer_signal": (_i, [_vp, _vp, _u32]),
    "kvdp_xfer_sync": (_i, [_vp]),
    "kvdp_plan_launch": (_i, [_vp, _vp, _vp, _i32, _i32, _i32, _vp]),
    "kvdp_stream_synchronize": (_i, [_vp]),
    "kvdp_stream_wait_flag": (_i, [_vp, _vp, _u32]),
    "kvdp_flag_cas64": (_i, [_vp, _u64, _u64, C.POINTER(_u64), _i32]),
    "kvdp_host_register": (_i, [_vp, _sz]),
    "kvdp_host_unregister": (_i, [_vp]),
    "kvdp_peer_enable": (_i, [_i32, _i32]),
    "kvdp_ipc_export": (_i, [_vp, _vp]),
    "kvdp_ipc_export_range": (_i, [_vp, _vp, C.POINTER(_u64)]),
    "kvdp_ipc_import": (_i, [_vp, _i32, _pp]),
    "kvdp_ipc_close": (_i, [_vp]),
    "kvdp_host_tier_create": (_i, [_pp, _vp, _i32, _i32, _i64, _i64, _vp, _i64, _i32, _i32]),
    "kvdp_host_tier_destroy": (_i, [_vp]),
    "kvdp_host_tier_offload": (_i, [_vp, _vp, _vp, _i32]),
    "kvdp_host_tier_onboard": (_i, [_vp, _vp, _vp, _i32]),
    "kvdp_host_tier_sync": (_i, [_vp]),
    "kvdp_hash_workspace_bytes": (_sz, [_i64]),
    "kvdp_hash_token_blocks": (_i, [_vp, _vp, _i64, _i64, _i32, _u64, _vp, _vp, _vp, _vp, _sz, _u32, _vp]),
    "kvdp_hash_token_blocks_noalloc": (_i, [_vp, _vp, _i64, _i64, _i32, _u64, _vp, _vp, _vp, _u32, _vp]),
    "kvdp_hash_token_blocks_host": (_i, [_vp, _vp, _i64, _i32, _u64, _vp, _vp, _vp, _u32, _i32]),
    "kvdp_xxh3_64_batch": (_i, [_vp, _i64, _i64, _u64, _vp, _vp]),
    "kvdp_index_create": (_i, [_pp, _i64, _i32, _i32]),
    "kvdp_index_destroy": (_i, [_vp]),
    "kvdp_index_stream": (_vp, [_vp]),
    "kvdp_index_apply_stored": (_i, [_vp, _i32, _i32, _u64, _vp, _vp, _i32]),
    "kvdp_index_apply_removed": (_i, [_vp, _i32, _vp, _i32]),
    "kvdp_index_remove_worker": (_i, [_vp, _i32]),
    "kvdp_index_flush": (_i, [_vp]),
    "kvdp_index_find_matches": (_i, [_vp, _vp, _vp, _i64, _i32, _i32, _vp, _vp, _vp]),
    "kvdp_index_find_matches_host": (_i, [_vp, _vp, _vp, _i64, _i32, _i32, _vp, _vp]),
    "kvdp_index_find_matches_for_tokens_host": (_i, [_vp, _vp, _vp, _i64, _i32, _i32, _i32, _vp, _vp]),
    "kvdp_index_touch": (_i64, [_vp, _vp, _i64, _i32, _i64, _i64, _vp, _i64]),
    "kvdp_index_stat": (_i64, [_vp, _i32, _u64]),
    # §I engine-facing KV-event ABI (argtypes as in the reference's CtypesEventPublisher fixture,
    # lib/bindings/python/tests/test_kv_bindings.py:141-176)
    "dynamo_llm_init": (C.c_uint32, [C.c_char_p, C.c_char_p, _i64, C.c_uint32]),
    "dynamo_llm_shutdown": (C.c_uint32, []),
    "dynamo_llm_load_publisher_create": (C.c_uint32, []),
    "dynamo_kv_event_publish_stored": (C.c_uint32, [_u64, C.POINTER(C.c_uint32), C.POINTER(C.c_size_t),
                                                    C.POINTER(_u64), C.c_size_t, C.POINTER(_u64), _u64]),
    "dynamo_kv_event_publish_removed": (C.c_uint32, [_u64, C.POINTER(_u64), C.c_size_t]),
    "kvdp_event_sink_callback": (_i, [_vp, _vp]),
    "kvdp_event_sink_index": (_i, [_vp, _i32]),
    "kvdp_event_flush": (_i, []),
    "kvdp_event_poll": (_i, [_vp, C.c_size_t, C.POINTER(C.c_size_t)]),
    "kvdp_event_stat": (_i64, [_i32]),
}

_lib = None


def lib() -> C.CDLL:
    """Load libkvdp.so (once). Raises if it is absent: the product has no other implementation."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m dynamo_exp_b200.build` "
                "(nvcc, sm_100a). dynamo_exp_b200 has no CPU or PyTorch fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def last_error() -> str:
    msg = lib().kvdp_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(rc: int) -> int:
    if rc != 0:
        raise KvdpError(rc, last_error())
    return rc


def require_gpu() -> None:
    """Fail loudly when no CUDA device is usable (tests that need the GPU call this first)."""
    n = C.c_int(0)
    rc = lib().kvdp_device_count(C.byref(n))
    if rc != 0 or n.value < 1:
        raise KvdpError(KVDP_ENODEV, f"no CUDA device visible to libkvdp.so: {last_error()}")


def set_option(name: str, value: int) -> None:
    check(lib().kvdp_set_option(name.encode(), int(value)))


def get_option(name: str) -> int:
    return int(lib().kvdp_get_option(name.encode()))
