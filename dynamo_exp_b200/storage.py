"""Host-side mirror of the reference's `kv::storage` module (lib/llm/src/kv/storage.rs), over libkvdp.so.

Same names, argument meaning and error behaviour for the parts of the API the KV data plane uses:
`StorageType`, `DType`, `Storage`, `OwnedStorage`, `DeviceStorageOwned`, `CudaPinnedMemory`,
`DeviceStorageFromAny`, `TensorView`. Errors are raised as `KvError` where the reference returns `Err(..)`.
Device memory comes from `kvdp_device_alloc` (cudarc in the reference), pinned memory from `cuda_malloc_host`.
"""
from __future__ import annotations

import ctypes as C
import enum
from dataclasses import dataclass
from typing import Any, List, Optional, Sequence

import numpy as np

from . import _lib


class KvError(RuntimeError):
    """The Python stand-in for the reference's `anyhow`/`dynamo_runtime::Result` errors."""


# ---------------------------------------------------------------------------------------------- StorageType / DType
@dataclass(frozen=True)
class StorageType:
    """reference: kv/storage.rs:47-52 (`Device(Arc<CudaContext>) | Pinned | System`)."""
    kind: str
    device: int = -1

    @staticmethod
    def Device(device: int = 0) -> "StorageType":
        return StorageType("device", int(device))

    @property
    def is_device(self) -> bool:
        return self.kind == "device"


StorageType.Pinned = StorageType("pinned")  # type: ignore[attr-defined]
StorageType.System = StorageType("system")  # type: ignore[attr-defined]


class DType(enum.Enum):
    """reference: kv/storage.rs:55-89."""
    F32 = 4
    F16 = 2
    BF16 = 102  # values only need to be distinct; size comes from size_in_bytes
    FP8 = 1
    U8 = 201
    U16 = 202
    U32 = 204
    U64 = 8
    I8 = 301
    I16 = 302
    I32 = 304
    I64 = 308

    def size_in_bytes(self) -> int:
        return {
            DType.F32: 4, DType.F16: 2, DType.BF16: 2, DType.FP8: 1, DType.U8: 1, DType.U16: 2, DType.U32: 4,
            DType.U64: 8, DType.I8: 1, DType.I16: 2, DType.I32: 4, DType.I64: 8,
        }[self]


# ---------------------------------------------------------------------------------------------- Storage
class Storage:
    """reference: `trait Storage`, kv/storage.rs:103-124."""

    def get_pointer(self) -> int:
        raise NotImplementedError

    def storage_size(self) -> int:
        raise NotImplementedError

    def storage_type(self) -> StorageType:
        raise NotImplementedError

    def view(self, shape: Sequence[int], dtype: DType) -> "TensorView":
        return TensorView(self, shape, dtype.size_in_bytes())


class DeviceStorageOwned(Storage):
    """Zero-filled device slab (reference: kv/storage.rs:189-230)."""

    def __init__(self, nbytes: int, device: int = 0):
        _lib.require_gpu()
        p = C.c_void_p()
        rc = _lib.lib().kvdp_device_alloc(C.byref(p), int(nbytes), int(device))
        if rc != 0:
            raise KvError(f"Failed to allocate {nbytes} bytes of device memory: {_lib.last_error()}")
        self._ptr = p.value or 0
        self._bytes = int(nbytes)
        self._device = int(device)

    def get_pointer(self) -> int:
        return self._ptr

    def storage_size(self) -> int:
        return self._bytes

    def storage_type(self) -> StorageType:
        return StorageType.Device(self._device)

    def device_ptr(self) -> int:
        return self._ptr

    def __del__(self):
        if getattr(self, "_ptr", 0):
            try:
                _lib.lib().kvdp_device_free(C.c_void_p(self._ptr))
            except Exception:
                pass
            self._ptr = 0

    def __repr__(self):
        return f"DeviceStorageOwned(ptr=0x{self._ptr:x}, bytes={self._bytes}, device={self._device})"


class CudaPinnedMemory(Storage):
    """Pinned host slab, zeroed by the CPU (reference: kv/storage.rs:242-315)."""

    def __init__(self, nbytes: int):
        if nbytes == 0:
            raise KvError("Bytes must be greater than 0")
        _lib.require_gpu()
        p = C.c_void_p()
        rc = _lib.lib().cuda_malloc_host(C.byref(p), int(nbytes))
        if rc != 0 or not p.value:
            raise KvError("cudaMallocHost failed")
        self._ptr = p.value
        self._bytes = int(nbytes)
        C.memset(self._ptr, 0, self._bytes)

    def get_pointer(self) -> int:
        return self._ptr

    def storage_size(self) -> int:
        return self._bytes

    def size(self) -> int:
        return self._bytes

    def storage_type(self) -> StorageType:
        return StorageType.Pinned

    def as_numpy(self, dtype=np.uint8) -> np.ndarray:
        buf = (C.c_uint8 * self._bytes).from_address(self._ptr)
        return np.frombuffer(buf, dtype=dtype)

    def __del__(self):
        if getattr(self, "_ptr", 0):
            try:
                _lib.lib().cuda_free_host(C.c_void_p(self._ptr))
            except Exception:
                pass
            self._ptr = 0

    def __repr__(self):
        return f"CudaPinnedMemory(ptr=0x{self._ptr:x}, bytes={self._bytes})"


class SystemMemory(Storage):
    """Plain (pageable) host memory backed by a numpy buffer; the reference's `StorageType::System`."""

    def __init__(self, nbytes: int):
        self._buf = np.zeros(int(nbytes), dtype=np.uint8)

    def get_pointer(self) -> int:
        return self._buf.ctypes.data

    def storage_size(self) -> int:
        return self._buf.size

    def storage_type(self) -> StorageType:
        return StorageType.System

    def as_numpy(self, dtype=np.uint8) -> np.ndarray:
        return self._buf.view(dtype)


class DeviceStorageFromAny(Storage):
    """Bring-your-own device pointer kept alive by `source` (reference: kv/storage.rs:1171-1232)."""

    def __init__(self, source: Any, device: int, ptr: int, nbytes: int):
        self._source = source
        self._device = int(device)
        self._ptr = int(ptr)
        self._bytes = int(nbytes)

    def source(self) -> Any:
        return self._source

    def get_pointer(self) -> int:
        return self._ptr

    def storage_size(self) -> int:
        return self._bytes

    def storage_type(self) -> StorageType:
        return StorageType.Device(self._device)


class OwnedStorage(Storage):
    """reference: kv/storage.rs:126-187."""

    def __init__(self, storage: Storage):
        self.storage = storage

    @staticmethod
    def create(nbytes: int, storage_type: StorageType) -> "OwnedStorage":
        if storage_type.kind == "device":
            return OwnedStorage.create_device_array(nbytes, storage_type.device)
        if storage_type.kind == "pinned":
            return OwnedStorage.create_pinned_array(nbytes)
        # kv/storage.rs:141: `StorageType::System => raise!("System memory not yet supported")`
        raise KvError("System memory not yet supported")

    @staticmethod
    def create_device_array(nbytes: int, device: int = 0) -> "OwnedStorage":
        return OwnedStorage(DeviceStorageOwned(nbytes, device))

    @staticmethod
    def create_pinned_array(nbytes: int) -> "OwnedStorage":
        return OwnedStorage(CudaPinnedMemory(nbytes))

    @staticmethod
    def byo_device_array(device_ptr: int, nbytes: int, device: int, owner: Any) -> "OwnedStorage":
        return OwnedStorage(DeviceStorageFromAny(owner, device, device_ptr, nbytes))

    def get_pointer(self) -> int:
        return self.storage.get_pointer()

    def storage_size(self) -> int:
        return self.storage.storage_size()

    def storage_type(self) -> StorageType:
        return self.storage.storage_type()

    def __repr__(self):
        return f"OwnedStorage({self.storage_type()})"


# ---------------------------------------------------------------------------------------------- TensorView
def calculate_strides(shape: Sequence[int]) -> List[int]:
    """row-major element strides (reference: kv/storage.rs:1137-1149)."""
    strides = [0] * len(shape)
    if shape:
        strides[-1] = 1
        for i in range(len(shape) - 2, -1, -1):
            strides[i] = strides[i + 1] * shape[i + 1]
    return strides


def unflatten_index(flat_idx: int, shape: Sequence[int]) -> List[int]:
    """reference: kv/storage.rs:1113-1135."""
    idx = [0] * len(shape)
    rem = flat_idx
    for i in range(len(shape) - 1, -1, -1):
        idx[i] = rem % shape[i]
        rem //= shape[i]
    return idx


class TensorView:
    """Strided view over a `Storage` (reference: kv/storage.rs:329-1110)."""

    def __init__(self, storage: Storage, shape: Sequence[int], element_size: int):
        shape = [int(s) for s in shape]
        strides = calculate_strides(shape)
        total = int(np.prod(shape, dtype=np.int64)) if shape else 1
        if total * element_size > storage.storage_size():
            raise KvError(f"Shape {shape} requires {total * element_size} bytes, but storage only has "
                          f"{storage.storage_size()} bytes")
        self.storage = storage
        self._shape = shape
        self._strides = strides
        self._byte_strides = [s * element_size for s in strides]
        self.offset = 0
        self._element_size = int(element_size)
        self._total = total

    @classmethod
    def with_strides(cls, storage: Storage, shape: Sequence[int], strides: Sequence[int], offset: int,
                     element_size: int) -> "TensorView":
        shape = [int(s) for s in shape]
        strides = [int(s) for s in strides]
        if len(shape) != len(strides):
            raise KvError("shape and strides must have the same rank")
        byte_strides = [s * element_size for s in strides]
        max_offset = offset + sum((d - 1) * bs for d, bs in zip(shape, byte_strides) if d > 0)
        if max_offset > storage.storage_size():
            raise KvError(f"View would access up to byte offset {max_offset}, but storage size is only "
                          f"{storage.storage_size()} bytes")
        v = cls.__new__(cls)
        v.storage = storage
        v._shape = shape
        v._strides = strides
        v._byte_strides = byte_strides
        v.offset = int(offset)
        v._element_size = int(element_size)
        v._total = int(np.prod(shape, dtype=np.int64)) if shape else 1
        return v

    # -- accessors
    def shape(self) -> List[int]:
        return list(self._shape)

    def strides(self) -> List[int]:
        return list(self._strides)

    def byte_strides(self) -> List[int]:
        return list(self._byte_strides)

    def element_size(self) -> int:
        return self._element_size

    def num_elements(self) -> int:
        return self._total

    def size_in_bytes(self) -> int:
        return self._total * self._element_size

    def data(self) -> int:
        return self.storage.get_pointer()

    def storage_type(self) -> StorageType:
        return self.storage.storage_type()

    # -- index math
    def _validate(self, indices: Sequence[int]) -> None:
        if len(indices) != len(self._shape):
            raise KvError(f"Expected {len(self._shape)} indices, got {len(indices)}")
        for dim, (idx, size) in enumerate(zip(indices, self._shape)):
            if idx >= size or idx < 0:
                raise KvError(f"Index {idx} out of bounds for dimension {dim} with size {size}")

    def in_bounds(self, indices: Sequence[int]) -> bool:
        return len(indices) == len(self._shape) and all(0 <= i < s for i, s in zip(indices, self._shape))

    def flat_index(self, indices: Sequence[int]) -> int:
        self._validate(indices)
        return sum(i * s for i, s in zip(indices, self._strides))

    def byte_offset(self, indices: Sequence[int]) -> int:
        self._validate(indices)
        return self.offset + sum(i * s for i, s in zip(indices, self._byte_strides))

    def address(self, indices: Sequence[int]) -> int:
        return self.storage.get_pointer() + self.byte_offset(indices)

    def indices_iter(self):
        """All index tuples in row-major order (reference: kv/storage.rs:809-815)."""
        for flat in range(self._total):
            yield unflatten_index(flat, self._shape)

    def is_contiguous(self) -> bool:
        expected, expected_b = 1, self._element_size
        for i in range(len(self._shape) - 1, -1, -1):
            if self._strides[i] != expected or self._byte_strides[i] != expected_b:
                return False
            expected *= self._shape[i]
            expected_b *= self._shape[i]
        return True

    def slice(self, dim: int, start: int, end: Optional[int] = None) -> "TensorView":
        if dim >= len(self._shape):
            raise KvError(f"Dimension {dim} out of bounds for tensor with {len(self._shape)} dimensions")
        end_idx = self._shape[dim] if end is None else end
        if end_idx > self._shape[dim]:
            raise KvError(f"End index {end_idx} out of bounds for dimension {dim} with size {self._shape[dim]}")
        if start >= end_idx:
            raise KvError(f"Invalid slice range: start={start}, end={end_idx}")
        new_shape = list(self._shape)
        new_shape[dim] = end_idx - start
        v = TensorView.__new__(TensorView)
        v.storage = self.storage
        v._shape = new_shape
        v._strides = list(self._strides)
        v._byte_strides = list(self._byte_strides)
        v.offset = self.offset + start * self._byte_strides[dim]
        v._element_size = self._element_size
        v._total = int(np.prod(new_shape, dtype=np.int64))
        return v

    # -- host access (pinned / system storage only)
    def as_numpy(self, dtype) -> np.ndarray:
        """Strided numpy view (the reference's `as_ndarray_view`, kv/storage.rs:719-800)."""
        if self.storage_type().is_device:
            raise KvError("Cannot create a host view of device memory")
        dt = np.dtype(dtype)
        if dt.itemsize != self._element_size:
            raise KvError(f"Element size mismatch: {dt.itemsize} vs {self._element_size}")
        n = self.storage.storage_size()
        buf = (C.c_uint8 * n).from_address(self.storage.get_pointer())
        base = np.frombuffer(buf, dtype=np.uint8)[self.offset:]
        usable = (base.size // dt.itemsize) * dt.itemsize
        return np.lib.stride_tricks.as_strided(base[:usable].view(dt), shape=self._shape, strides=self._byte_strides,
                                               writeable=True)

    # -- copies
    def _pair_checks(self, other: "TensorView") -> None:
        if self._shape != other._shape:
            raise KvError(f"Shape mismatch: {self._shape} vs {other._shape}")
        if self._element_size != other._element_size:
            raise KvError(f"Element size mismatch: {self._element_size} vs {other._element_size}")

    def h2d(self, device_view: "TensorView", stream: int = 0) -> None:
        """reference: kv/storage.rs:915-982 (async on `stream`; a raw cudaStream_t value, 0 = default)."""
        if self.storage_type().is_device:
            raise KvError("Source must be a host tensor (System or Pinned)")
        if not device_view.storage_type().is_device:
            raise KvError("Destination must be a device tensor")
        self._pair_checks(device_view)
        if not self.is_contiguous():
            raise KvError("Source tensor must be contiguous")
        if not device_view.is_contiguous():
            raise KvError("Destination tensor must be contiguous")
        rc = _lib.lib().cuda_memcpy_async(C.c_void_p(device_view.data() + device_view.offset),
                                          C.c_void_p(self.data() + self.offset), self.size_in_bytes(),
                                          C.c_void_p(stream))
        if rc != 0:
            raise KvError("cudaMemcpyAsync failed during host-to-device transfer")

    def d2h(self, host_view: "TensorView", stream: int = 0) -> None:
        """reference: kv/storage.rs:984-1050."""
        if not self.storage_type().is_device:
            raise KvError("Source must be a device tensor")
        if host_view.storage_type().is_device:
            raise KvError("Destination must be a host tensor (System or Pinned)")
        self._pair_checks(host_view)
        if not self.is_contiguous():
            raise KvError("Source tensor must be contiguous")
        if not host_view.is_contiguous():
            raise KvError("Destination tensor must be contiguous")
        rc = _lib.lib().cuda_memcpy_async(C.c_void_p(host_view.data() + host_view.offset),
                                          C.c_void_p(self.data() + self.offset), self.size_in_bytes(),
                                          C.c_void_p(stream))
        if rc != 0:
            raise KvError("cudaMemcpyAsync failed during device-to-host transfer")

    def copy_to_view_blocking(self, dst_view: "TensorView") -> None:
        """reference: kv/storage.rs:637-675 (whole-storage contiguous blocking copy)."""
        if self._shape != dst_view._shape or self._strides != dst_view._strides:
            raise KvError(f"Shape or strides mismatch: {self._shape} vs {dst_view._shape}")
        if not self.is_contiguous():
            raise KvError("Source is not contiguous")
        if not dst_view.is_contiguous():
            raise KvError("Destination is not contiguous")
        rc = _lib.lib().cuda_memcpy_sync(C.c_void_p(dst_view.data()), C.c_void_p(self.data()), self.size_in_bytes())
        if rc != 0:
            raise KvError("cudaMemcpyAsync failed")

    def __repr__(self):
        return (f"TensorView(shape={self._shape}, strides={self._strides}, offset={self.offset}, "
                f"element_size={self._element_size}, storage_type={self.storage_type()})")
