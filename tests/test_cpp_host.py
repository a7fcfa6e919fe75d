"""The C++ host layer (include/kvdp.hpp: the reference's Rust kv::storage / kv::layer / lib-tokens interface restated in
C++ over the C ABI) compiles everywhere and, on a GPU, passes the reference's own unit tests restated in
tests/cpp/host_mirror_test.cpp."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def compile_host_test(kvlib, out_dir, name="host_mirror_test"):
    exe = os.path.join(out_dir, name)
    libdir = os.path.dirname(kvlib.LIB_PATH)
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", name + ".cpp"), "-L", libdir, "-lkvdp",
           "-Wl,-rpath," + libdir, "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_cpp_host_layer_compiles_against_the_c_abi(kvlib, tmp_path):
    exe = compile_host_test(kvlib, str(tmp_path))
    assert os.path.exists(exe)


def test_c_header_is_plain_c(tmp_path):
    """kvdp.h must be consumable by a C compiler (it is what cgo / bindgen / ctypes users read)."""
    src = tmp_path / "t.c"
    src.write_text('#include "kvdp.h"\nint main(void){ kvdp_plan_desc d; (void)d; return KVDP_OK; }\n')
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(src),
                        "-o", str(tmp_path / "t.o")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


@pytest.mark.gpu
def test_cpp_host_layer_passes_the_reference_unit_tests(gpu, tmp_path):
    exe = compile_host_test(gpu, str(tmp_path))
    r = subprocess.run([exe], capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "CPP HOST MIRROR OK" in r.stdout


def test_cpp_reuse_pool_passes_the_reference_pool_tests(kvlib, tmp_path):
    """include/kvdp_router.hpp: AvailableBlocks / ReservedBlocks are plain host logic, so the reference's pool tests
    (kv/reuse.rs:684-1062, kv/reserved.rs:161-242) run without a GPU."""
    exe = compile_host_test(kvlib, str(tmp_path), "router_reuse_test")
    r = subprocess.run([exe, "pool"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "CPP POOL OK" in r.stdout


@pytest.mark.gpu
def test_cpp_router_and_reuse_layer_on_gpu(gpu, tmp_path):
    """RadixTree / KvIndexerSharded / KvEventPublisher / KvStorageManager in C++ over the C ABI."""
    exe = compile_host_test(gpu, str(tmp_path), "router_reuse_test")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "CPP ROUTER/REUSE OK" in r.stdout
