"""Host KV memory that several worker processes of one node map and pin.

The reference pins host KV storage per process (`lib/llm/src/kv/storage.rs:91-101`, `cuda_malloc_host`), so a block held
in a prefill worker's host memory can only reach a decode GPU through the prefill GPU's PCIe link. On a B200 box every
GPU has its own PCIe link to the same host memory and NVLink between the GPUs is 15x faster than either, so a host tier
that is mapped by both workers lets the decode GPU fetch part of the blocks itself while the prefill GPU relays the rest:
two links instead of one for the same onboard.

`SharedHostRegion.create` makes an anonymous memory file (memfd), maps it and registers the mapping with CUDA
(`kvdp_host_register`); `handle()` is a small tuple another process of the same user on the same node can pass to
`SharedHostRegion.open` (it reaches the file through /proc/<pid>/fd/<fd>, so the creator must stay alive until the
opener has mapped it). After that both processes use `region.ptr` like `cuda_malloc_host` memory with every libkvdp entry
point. `register=False` skips the CUDA registration (host-side logic and CPU tests)."""
from __future__ import annotations

import ctypes as C
import mmap
import os
from typing import Tuple

Handle = Tuple[int, int, int]  # (pid of the creator, fd in the creator, bytes)


class SharedHostRegion:
    def __init__(self, fd: int, nbytes: int, owner: bool, register: bool):
        self.fd, self.nbytes, self.owner = fd, int(nbytes), owner
        self._mm = mmap.mmap(fd, self.nbytes, flags=mmap.MAP_SHARED, prot=mmap.PROT_READ | mmap.PROT_WRITE)
        self._anchor = (C.c_uint8 * self.nbytes).from_buffer(self._mm)
        self.ptr = C.addressof(self._anchor)
        self.registered = False
        if register:
            from . import _lib
            try:
                _lib.check(_lib.lib().kvdp_host_register(C.c_void_p(self.ptr), self.nbytes))
            except Exception:
                self._unmap()
                raise
            self.registered = True

    # ---------------------------------------------------------------- construction
    @classmethod
    def create(cls, nbytes: int, register: bool = True, name: str = "kvdp_host") -> "SharedHostRegion":
        if nbytes <= 0:
            raise ValueError("SharedHostRegion.create: size must be positive")
        page = mmap.PAGESIZE
        nbytes = (int(nbytes) + page - 1) // page * page
        fd = os.memfd_create(name, 0)
        try:
            os.ftruncate(fd, nbytes)
            return cls(fd, nbytes, True, register)
        except Exception:
            os.close(fd)
            raise

    @classmethod
    def open(cls, handle: Handle, register: bool = True) -> "SharedHostRegion":
        pid, fd, nbytes = (int(x) for x in handle)
        path = "/proc/%d/fd/%d" % (pid, fd)
        try:
            mine = os.open(path, os.O_RDWR)
        except OSError as e:
            raise OSError(e.errno, "SharedHostRegion.open: cannot reach %s (creator gone, or another user / pid "
                                   "namespace): %s" % (path, e.strerror)) from e
        try:
            if os.fstat(mine).st_size < nbytes:
                raise ValueError("SharedHostRegion.open: the file behind %s holds %d bytes, the handle says %d"
                                 % (path, os.fstat(mine).st_size, nbytes))
            return cls(mine, nbytes, False, register)
        except Exception:
            os.close(mine)
            raise

    def handle(self) -> Handle:
        return (os.getpid(), self.fd, self.nbytes)

    # ---------------------------------------------------------------- access
    def array(self):
        """numpy uint8 view of the whole region."""
        import numpy as np
        return np.frombuffer(self._anchor, dtype=np.uint8)

    # ---------------------------------------------------------------- teardown
    def _unmap(self):
        if self._mm is not None:
            self._anchor = None
            try:
                self._mm.close()
            except BufferError:
                pass          # a numpy view handed out by array() is still alive; the mapping goes with it
            self._mm = None
        if self.fd >= 0:
            os.close(self.fd)
            self.fd = -1

    def close(self):
        if self.registered:
            from . import _lib
            _lib.check(_lib.lib().kvdp_host_unregister(C.c_void_p(self.ptr)))
            self.registered = False
        self._unmap()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def split_for_links(num_requests: int, links: int = 2):
    """Which requests of a step each link carries: contiguous, near-equal shares (link 0 = the owner's GPU)."""
    if num_requests < 0 or links < 1:
        raise ValueError("split_for_links: bad arguments")
    base, extra = divmod(num_requests, links)
    out, o = [], 0
    for i in range(links):
        n = base + (1 if i < extra else 0)
        out.append((o, o + n))
        o += n
    return out
