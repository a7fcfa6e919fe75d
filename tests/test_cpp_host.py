"""The C++ host layer (include/kvdp.hpp: the reference's Rust kv::storage / kv::layer / lib-tokens interface restated in
C++ over the C ABI) compiles everywhere and, on a GPU, passes the reference's own unit tests restated in
tests/cpp/host_mirror_test.cpp."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def compile_host_test(kvlib, out_dir):
    exe = os.path.join(out_dir, "host_mirror_test")
    libdir = os.path.dirname(kvlib.LIB_PATH)
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "host_mirror_test.cpp"), "-L", libdir, "-lkvdp",
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
