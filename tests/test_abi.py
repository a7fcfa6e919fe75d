"""CPU: libkvdp.so builds, loads without a GPU, and exports exactly the symbols include/kvdp.h declares
(the 14 legacy block_copy.cu symbols + the v2 API). No compute calls here."""
import ctypes as C
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "kvdp.h")

LEGACY = ["copy_blocks_3d", "copy_blocks_launcher_3d", "copy_blocks_memcpy_3d", "copy_stream_create",
          "copy_stream_destroy", "copy_stream_prepare_block_ids", "copy_stream_launch", "copy_stream_memcpy",
          "copy_stream_sync", "copy_stream_scatter", "cuda_malloc_host", "cuda_free_host", "cuda_memcpy_async",
          "cuda_memcpy_sync"]


def header_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"^\s*#.*$", "", src, flags=re.M)
    names = re.findall(r"\b([a-z_0-9]+)\s*\([^;{}]*\)\s*;", src)
    return sorted(set(n for n in names if n.startswith(("kvdp_", "copy_", "cuda_", "dynamo_"))))


def test_header_declares_the_14_legacy_symbols():
    fns = header_functions()
    for name in LEGACY:
        assert name in fns, name


def test_library_exports_every_declared_symbol(kvlib):
    L = C.CDLL(kvlib.LIB_PATH)
    for name in header_functions():
        assert hasattr(L, name), f"{name} declared in kvdp.h but not exported"


def test_python_binding_covers_the_header(kvlib):
    assert sorted(kvlib.PROTOTYPES) == header_functions()


def test_no_cpu_compute_symbols_and_sm100a_only(kvlib):
    out = subprocess.run(["cuobjdump", "--list-elf", kvlib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_tma_bulk_copies_present_in_sass(kvlib):
    """UBLKCP is the SASS of cp.async.bulk (B200_PROFILING.md): both the copy ring and the hash staging use it."""
    sass = subprocess.run(["cuobjdump", "-sass", kvlib.LIB_PATH], capture_output=True, text=True).stdout
    assert "UBLKCP" in sass
    per_fn = re.split(r"Function : ", sass)
    with_tma = [f.split("\n", 1)[0] for f in per_fn if "UBLKCP" in f]
    assert any("copy_tma_tma" in f for f in with_tma)
    assert any("copy_tma_in" in f for f in with_tma)
    assert any("hash_blocks_tma" in f for f in with_tma)


def test_loads_and_reports_without_gpu(kvlib):
    L = kvlib.lib()
    assert L.kvdp_version() == 200
    assert L.kvdp_hash_workspace_bytes(0) >= 8
    assert L.kvdp_set_option(b"no_such_option", 1) == kvlib.KVDP_EINVAL
    assert b"unknown option" in L.kvdp_last_error()
    assert L.kvdp_set_option(b"copy_path", 7) == kvlib.KVDP_EINVAL
    assert L.kvdp_get_option(b"copy_path") == 0
    # host-side tunables: defaults, round trip, range checks
    assert L.kvdp_get_option(b"host_dma_piece_kib") == 8192
    assert L.kvdp_set_option(b"host_dma_piece_kib", 0) == 0 and L.kvdp_get_option(b"host_dma_piece_kib") == 0
    assert L.kvdp_set_option(b"host_dma_piece_kib", -1) == kvlib.KVDP_EINVAL
    assert L.kvdp_set_option(b"host_dma_piece_kib", 8192) == 0
    # registering host memory needs a CUDA context: without a GPU it must fail with an error, not crash; NULL is EINVAL
    assert L.kvdp_host_register(None, 4096) == kvlib.KVDP_EINVAL
    assert L.kvdp_host_unregister(None) == kvlib.KVDP_EINVAL


def test_oracle_is_not_linked_or_imported_by_the_product():
    pkg = os.path.join(ROOT, "dynamo_exp_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "kvoracle" not in text and "oracle/" not in text and "import oracle" not in text, f
    ldd = subprocess.run(["ldd", os.path.join(pkg, "libkvdp.so")], capture_output=True, text=True).stdout
    assert "kvoracle" not in ldd


def test_rust_sys_crate_declares_every_header_function():
    """rust/kvdp-sys is shipped as source only (no Rust toolchain here); at least keep it in step with the header."""
    text = open(os.path.join(ROOT, "rust", "kvdp-sys", "src", "lib.rs")).read()
    declared = sorted(set(re.findall(r"pub fn ([a-z_0-9]+)\(", text)))
    assert declared == header_functions()
    # struct layout mirror: every kvdp_plan_desc field appears, in header order
    hdr = open(HEADER).read()
    body = hdr[hdr.index("typedef struct kvdp_plan_desc {"):hdr.index("} kvdp_plan_desc;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = []
    for stmt in body.split("{", 1)[1].split(";"):
        stmt = stmt.strip()
        if not stmt:
            continue
        names = re.sub(r"^(const\s+)?[a-z0-9_]+\s*\*?", "", stmt)
        fields += [n.strip().lstrip("*").strip() for n in names.split(",")]
    rust = re.findall(r"pub ([a-z_]+): ", text[text.index("pub struct KvdpPlanDesc"):text.index("pub const KVDP_OK")])
    assert rust == fields
