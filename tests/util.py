"""Shared helpers for the parity tests: synthetic data and raw device I/O through the C ABI."""
import ctypes as C

import numpy as np

MASK = (1 << 64) - 1


def mix64(x):
    """splitmix64 finaliser over numpy uint64 arrays."""
    x = (np.asarray(x, dtype=np.uint64) + np.uint64(0x9E3779B97F4A7C15))
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def payload(nbytes, seed):
    """Deterministic byte payload in which every 8-byte word is distinct with overwhelming probability."""
    with np.errstate(over="ignore"):
        n = (nbytes + 7) // 8
        w = mix64(np.arange(n, dtype=np.uint64) ^ (np.uint64(seed) << np.uint64(40)))
    return w.view(np.uint8)[:nbytes].copy()


def tokens_for(req, length, seed=0xD15A66, vocab=128256):
    with np.errstate(over="ignore"):
        pos = np.arange(length, dtype=np.uint64)
        v = mix64(np.uint64(seed) ^ ((np.uint64(req) << np.uint64(32)) | pos))
    return (v % np.uint64(vocab)).astype(np.uint32)


def permutation(n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.permutation(n).astype(np.int32)


class DevBuf:
    """Raw device allocation through kvdp_device_alloc with numpy upload/download."""

    def __init__(self, lib, nbytes, device=-1):
        self.lib = lib
        self.nbytes = int(nbytes)
        p = C.c_void_p()
        lib.check(lib.lib().kvdp_device_alloc(C.byref(p), self.nbytes, device))
        self.ptr = p.value

    def upload(self, arr, offset=0):
        a = np.ascontiguousarray(arr)
        assert offset + a.nbytes <= self.nbytes
        self.lib.check(self.lib.lib().cuda_memcpy_sync(C.c_void_p(self.ptr + offset), C.c_void_p(a.ctypes.data),
                                                       a.nbytes))

    def download(self, dtype=np.uint8, nbytes=None, offset=0):
        n = self.nbytes - offset if nbytes is None else nbytes
        out = np.empty(n, dtype=np.uint8)
        self.lib.check(self.lib.lib().cuda_memcpy_sync(C.c_void_p(out.ctypes.data), C.c_void_p(self.ptr + offset), n))
        return out.view(dtype)

    def free(self):
        if self.ptr:
            self.lib.lib().kvdp_device_free(C.c_void_p(self.ptr))
            self.ptr = 0

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class PinnedBuf:
    def __init__(self, lib, nbytes):
        self.lib = lib
        self.nbytes = int(nbytes)
        p = C.c_void_p()
        lib.check(lib.lib().cuda_malloc_host(C.byref(p), self.nbytes))
        self.ptr = p.value
        self.array = np.frombuffer((C.c_uint8 * self.nbytes).from_address(self.ptr), dtype=np.uint8)
        self.array[:] = 0

    def free(self):
        if self.ptr:
            self.array = None
            self.lib.lib().cuda_free_host(C.c_void_p(self.ptr))
            self.ptr = 0

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
