"""Host-side mirror of the reference's `kv::layer` module (lib/llm/src/kv/layer.rs), over libkvdp.so.

Same names, argument meaning, state machine and error behaviour: `KvLayout`, `KvModelDetails`, `KvBlockDetails`,
`KvLayer`, `KvBlockStorage`, `CopyStreamBlockMap`, `CopyStream`. `CopyStream` drives the LEGACY C ABI exactly as the
Rust code does (copy_stream_create / _prepare_block_ids / _memcpy / _sync / _scatter), so these classes double as the
drop-in test of that ABI. The v2 additions (`KvContiguousBlocks`, `TransferEngine`) live in `transfer.py`.
"""
from __future__ import annotations

import ctypes as C
import enum
from dataclasses import dataclass, field
from typing import List, Sequence

import numpy as np

from . import _lib
from .storage import DType, KvError, OwnedStorage, Storage, StorageType, TensorView


class KvLayout(enum.Enum):
    """reference: kv/layer.rs:99-106."""
    KvFirst = 0      # [2, blocks, block_size, heads, head_size]
    BlockFirst = 1   # [blocks, 2, block_size, heads, head_size]


@dataclass(frozen=True)
class KvModelDetails:
    """reference: kv/layer.rs:108-138."""
    number_of_layers: int
    number_of_heads: int
    head_size: int
    dtype: DType

    def number_of_elements_per_token_per_layer(self) -> int:
        return 2 * self.number_of_heads * self.head_size

    def bytes_per_token_per_layer(self) -> int:
        return self.number_of_elements_per_token_per_layer() * self.dtype.size_in_bytes()


@dataclass(frozen=True)
class KvBlockDetails:
    """reference: kv/layer.rs:140-212."""
    layout: KvLayout
    block_size: int
    model_details: KvModelDetails
    tp_rank: int = 0
    tp_size: int = 1

    def validate(self) -> None:
        md = self.model_details
        if self.tp_size < 1:
            raise KvError("tp_size must be at least 1")
        if md.number_of_heads % self.tp_size != 0:
            raise KvError("tp_size must evenly divide num_heads")
        if self.tp_rank >= self.tp_size:
            raise KvError("tp_rank must be less than tp_size")
        if self.tp_size > md.number_of_heads:
            raise KvError("tp_size must be less than num_heads")

    def bytes_per_token_block_per_layer(self) -> int:
        return (self.model_details.bytes_per_token_per_layer() * self.block_size) // self.tp_size

    def is_compatible(self, other: "KvBlockDetails") -> bool:
        return (self.layout == other.layout and self.block_size == other.block_size
                and self.tp_size == other.tp_size and self.model_details == other.model_details)

    def prefix_dim(self) -> int:
        return 2 if self.layout == KvLayout.KvFirst else 1

    def suffix_dim(self) -> int:
        s = self.block_size * (self.model_details.number_of_heads // self.tp_size) * self.model_details.head_size
        return s if self.layout == KvLayout.KvFirst else 2 * s

    def elem_size(self) -> int:
        return self.model_details.dtype.size_in_bytes()


def _ids32(ids: Sequence[int]) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(ids, dtype=np.int64).astype(np.int32))


class KvLayer(Storage):
    """One layer's paged KV tensor (reference: kv/layer.rs:214-451)."""

    def __init__(self, layout: KvLayout, storage: OwnedStorage, number_of_blocks: int, block_size: int,
                 number_of_heads: int, head_size: int, dtype: DType, tp_size: int = 1, tp_rank: int = 0):
        self.layout = layout
        self.storage = storage
        self.number_of_blocks = int(number_of_blocks)
        self.block_size = int(block_size)
        self.number_of_heads = int(number_of_heads)
        self.head_size = int(head_size)
        self.dtype = dtype
        self.tp_size = int(tp_size)
        self.tp_rank = int(tp_rank)
        self.validate()

    def validate(self) -> None:
        for name in ("number_of_blocks", "block_size", "number_of_heads", "head_size"):
            if getattr(self, name) < 1:
                raise KvError(f"{name} must be at least 1")
        if self.tp_size < 1 or self.number_of_heads % self.tp_size != 0:
            raise KvError("number_of_heads must be divisible by tp_size")
        if self.tp_rank >= self.tp_size:
            raise KvError("tp_rank must be less than tp_size")
        if self.tp_size > self.number_of_heads:
            raise KvError("tp_size must be less than number_of_heads")
        nbytes = int(np.prod(self.layer_shape(), dtype=np.int64)) * self.dtype.size_in_bytes()
        if self.storage.storage_size() < nbytes:
            raise KvError("storage must be at least as large as the layer")

    @classmethod
    def from_storage(cls, block_details: KvBlockDetails, number_of_blocks: int, storage: OwnedStorage) -> "KvLayer":
        md = block_details.model_details
        return cls(block_details.layout, storage, number_of_blocks, block_details.block_size, md.number_of_heads,
                   md.head_size, md.dtype, block_details.tp_size, block_details.tp_rank)

    # Storage
    def storage_type(self) -> StorageType:
        return self.storage.storage_type()

    def get_pointer(self) -> int:
        return self.storage.get_pointer()

    def storage_size(self) -> int:
        return self.storage.storage_size()

    def layer_shape(self) -> List[int]:
        h = self.number_of_heads // self.tp_size
        if self.layout == KvLayout.KvFirst:
            return [2, self.number_of_blocks, self.block_size, h, self.head_size]
        return [self.number_of_blocks, 2, self.block_size, h, self.head_size]

    def view(self) -> TensorView:  # type: ignore[override]
        if self.number_of_heads % self.tp_size != 0:
            raise KvError(f"Number of heads ({self.number_of_heads}) is not divisible by tp_size ({self.tp_size})")
        return TensorView(self, self.layer_shape(), self.dtype.size_in_bytes())

    def copy_blocks_to(self, src_block_ids: Sequence[int], dst: "KvLayer", dst_block_ids: Sequence[int]) -> None:
        """Blocking single-layer gather/scatter via `copy_blocks_3d` (reference: kv/layer.rs:371-450)."""
        if len(src_block_ids) != len(dst_block_ids):
            raise KvError("src_block_ids and dst_block_ids must have the same length")
        if self.layout != dst.layout:
            raise KvError("src and dst must have the same layout")
        s, d = self.storage.storage_type(), dst.storage.storage_type()
        if s.kind == "pinned" and d.kind == "pinned":
            raise KvError("Pinned to Pinned copy not implemented")
        if s.kind == "device" and d.kind == "device":
            raise KvError("Device to Device copy not implemented")
        if s.kind == "system":
            raise KvError("System to Device copy not implemented")
        if d.kind == "system":
            raise KvError("Device to System copy not implemented")
        hs, hd = _ids32(src_block_ids), _ids32(dst_block_ids)
        prefix_dim = 2 if self.layout == KvLayout.KvFirst else 1
        suffix_dim = self.head_size * (self.number_of_heads // self.tp_size) * self.block_size
        if self.layout == KvLayout.BlockFirst:
            suffix_dim *= 2
        rc = _lib.lib().copy_blocks_3d(C.c_void_p(self.storage.get_pointer()), C.c_void_p(dst.storage.get_pointer()),
                                       C.c_void_p(hs.ctypes.data), C.c_void_p(hd.ctypes.data), len(hs), prefix_dim,
                                       self.number_of_blocks, dst.number_of_blocks, suffix_dim,
                                       self.dtype.size_in_bytes())
        if rc != 0:
            raise KvError("Failed to copy blocks")


class KvBlockStorage:
    """Layer-separate set of KV layers (reference: kv/layer.rs:453-596)."""

    def __init__(self, block_details: KvBlockDetails, storage_type: StorageType, number_of_blocks: int,
                 layers: List[KvLayer]):
        self.block_details = block_details
        self._storage_type = storage_type
        self._number_of_blocks = number_of_blocks
        self.layers = layers

    @classmethod
    def from_layers(cls, layers: List[KvLayer]) -> "KvBlockStorage":
        if not layers:
            raise KvError("Layers must not be empty")
        st = layers[0].storage.storage_type()
        if any(l.storage.storage_type() != st for l in layers):
            raise KvError("All layers must have the same storage type")
        nb = layers[0].number_of_blocks
        if any(l.number_of_blocks != nb for l in layers):
            raise KvError("All layers must have the same number of blocks")
        md = KvModelDetails(len(layers), layers[0].number_of_heads, layers[0].head_size, layers[0].dtype)
        bd = KvBlockDetails(layers[0].layout, layers[0].block_size, md, layers[0].tp_rank, layers[0].tp_size)
        bd.validate()
        per_block = bd.bytes_per_token_block_per_layer()
        if any(l.storage.storage_size() < per_block for l in layers):
            raise KvError("All layers must have enough capacity to store hold the block data")
        return cls(bd, st, nb, layers)

    @classmethod
    def allocate(cls, number_of_blocks: int, block_details: KvBlockDetails,
                 storage_type: StorageType) -> "KvBlockStorage":
        block_details.validate()
        nbytes = block_details.bytes_per_token_block_per_layer() * number_of_blocks
        layers = []
        for layer in range(block_details.model_details.number_of_layers):
            try:
                storage = OwnedStorage.create(nbytes, storage_type)
            except KvError as e:
                raise KvError(f"Failed to allocate memory for KV BlockStorage for layer {layer}: {e}") from e
            layers.append(KvLayer.from_storage(block_details, number_of_blocks, storage))
        return cls.from_layers(layers)

    def layer(self, layer: int) -> KvLayer:
        if layer >= len(self.layers) or layer < 0:
            raise KvError(f"Layer index {layer} out of bounds (max {len(self.layers) - 1})")
        return self.layers[layer]

    layer_mut = layer

    def number_of_blocks(self) -> int:
        return self._number_of_blocks

    def storage_type(self) -> StorageType:
        return self._storage_type

    def layer_pointers(self) -> np.ndarray:
        return np.array([l.storage.get_pointer() for l in self.layers], dtype=np.uint64)


@dataclass
class CopyStreamBlockMap:
    """Per-(src set, dst set) copy description (reference: kv/layer.rs:592-709)."""
    src_layer_ptrs: List[int] = field(default_factory=list)
    dst_layer_ptrs: List[int] = field(default_factory=list)
    prefix_dim: int = 0
    src_block_dim: int = 0
    dst_block_dim: int = 0
    suffix_dim: int = 0
    elem_size: int = 0

    @classmethod
    def new(cls, src: KvBlockStorage, dst: KvBlockStorage) -> "CopyStreamBlockMap":
        if not src.block_details.is_compatible(dst.block_details):
            raise KvError("src and dst must have compatible block details")
        sp = [l.storage.get_pointer() for l in src.layers]
        dp = [l.storage.get_pointer() for l in dst.layers]
        if len(sp) != len(dp):
            raise KvError("src and dst must have the same number of layers")
        m = cls(sp, dp, src.block_details.prefix_dim(), src.number_of_blocks(), dst.number_of_blocks(),
                src.block_details.suffix_dim(), src.block_details.elem_size())
        m.validate()
        return m

    def validate(self) -> None:
        for name in ("prefix_dim", "src_block_dim", "dst_block_dim", "suffix_dim"):
            if getattr(self, name) < 1:
                raise KvError(f"{name} must be at least 1")
        if not (1 <= self.elem_size <= 8):
            raise KvError("elem_size must be between 1 and 8")
        if not self.src_layer_ptrs:
            raise KvError("src_layer_ptrs must not be empty")
        if not self.dst_layer_ptrs:
            raise KvError("dst_layer_ptrs must not be empty")
        if len(self.src_layer_ptrs) != len(self.dst_layer_ptrs):
            raise KvError("src_layer_ptrs and dst_layer_ptrs must have the same length")


class CopyStream:
    """Stateful async block copy over the legacy C ABI (reference: kv/layer.rs:711-1143).

    State machine: prepare_block_map -> prepare_block_ids -> trigger_layer* -> sync_stream -> reset.
    """

    def __init__(self, num_layers: int, num_blocks: int, check_unique_ids: bool = True):
        _lib.require_gpu()
        h = C.c_void_p()
        rc = _lib.lib().copy_stream_create(C.byref(h), int(num_layers), int(num_blocks))
        if rc != 0 or not h.value:
            raise KvError("Failed to create copy stream")
        self._handle = h
        self.max_num_layers = int(num_layers)
        self.max_num_blocks = int(num_blocks)
        self.staged_layers = False
        self.staged_block_ids = False
        self.layer_doorbells = [False] * int(num_layers)
        self.src_block_ids = np.zeros(0, dtype=np.int32)
        self.dst_block_ids = np.zeros(0, dtype=np.int32)
        self.layer_details = CopyStreamBlockMap()
        # the reference checks id uniqueness only under debug_assertions (kv/layer.rs:848-861)
        self.check_unique_ids = check_unique_ids

    @classmethod
    def new(cls, num_layers: int, num_blocks: int) -> "CopyStream":
        return cls(num_layers, num_blocks)

    def src_block_dim(self) -> int:
        return self.layer_details.src_block_dim

    def dst_block_dim(self) -> int:
        return self.layer_details.dst_block_dim

    def prepare_block_map(self, details: CopyStreamBlockMap) -> None:
        layer_count = len(details.src_layer_ptrs)
        if self.max_num_layers < layer_count:
            raise KvError(f"Number of layers {layer_count} exceeds max number of layers {self.max_num_layers}")
        if self.staged_layers:
            raise KvError("Layers already loaded")
        self.staged_layers = True
        self.layer_details = details
        self.layer_doorbells = [False] * len(self.layer_doorbells)

    def prepare_block_ids(self, src_block_ids: Sequence[int], dst_block_ids: Sequence[int]) -> None:
        if len(src_block_ids) != len(dst_block_ids):
            raise KvError("src_block_ids and dst_block_ids must have the same length")
        if self.check_unique_ids:
            if len(set(int(x) for x in dst_block_ids)) != len(dst_block_ids):
                raise KvError("dst_block_ids must be unique")
            if len(set(int(x) for x in src_block_ids)) != len(src_block_ids):
                raise KvError("src_block_ids must be unique")
        if self.max_num_blocks < len(src_block_ids):
            raise KvError(f"Number of blocks {len(src_block_ids)} exceeds max number of blocks {self.max_num_blocks}")
        if not self.staged_layers:
            raise KvError("Layers must be loaded before preparing block ids")
        if self.staged_block_ids:
            raise KvError("Block ids already loaded")
        self.src_block_ids = _ids32(src_block_ids)
        self.dst_block_ids = _ids32(dst_block_ids)
        rc = _lib.lib().copy_stream_prepare_block_ids(self._handle, C.c_void_p(self.src_block_ids.ctypes.data),
                                                      C.c_void_p(self.dst_block_ids.ctypes.data),
                                                      len(self.src_block_ids))
        if rc != 0:
            raise KvError("Failed to prepare block ids")
        self.staged_block_ids = True

    def trigger_layer(self, layer: int) -> None:
        if not self.staged_layers:
            raise KvError("Layers must be loaded before triggering a layer")
        if not self.staged_block_ids:
            raise KvError("Block ids must be loaded before triggering a layer")
        n = len(self.layer_details.src_layer_ptrs)
        if layer >= n or layer < 0:
            raise KvError(f"layer index {layer} out of bounds (max {n - 1})")
        if self.layer_doorbells[layer]:
            return  # already triggered: no-op (kv/layer.rs:925-928)
        d = self.layer_details
        rc = _lib.lib().copy_stream_memcpy(self._handle, C.c_void_p(d.src_layer_ptrs[layer]),
                                           C.c_void_p(d.dst_layer_ptrs[layer]), d.prefix_dim, d.suffix_dim,
                                           d.elem_size, d.src_block_dim, d.dst_block_dim)
        if rc != 0:
            raise KvError(f"Failed to execute layer {layer} copy")
        self.layer_doorbells[layer] = True

    def layer_count(self) -> int:
        return len(self.layer_details.src_layer_ptrs)

    def trigger_all_layers(self) -> None:
        for layer in range(self.layer_count()):
            self.trigger_layer(layer)

    def sync_stream(self) -> None:
        rc = _lib.lib().copy_stream_sync(self._handle)
        if rc != 0:
            raise KvError("Failed to synchronize copy stream")

    def scatter_copy_layer(self, layer: int, dims: Sequence[int], elem_size: int, block_dim_index: int,
                           src_tp_size: int, dst_tp_size: int) -> None:
        """TP-up reshard of one layer (reference: kv/layer.rs:1027-1130)."""
        if len(dims) != 5:
            raise KvError("Expected 5 dimensions for src_data")
        if block_dim_index > 1:
            raise KvError("block_dim_index must be 0 or 1")
        if elem_size == 0 or elem_size > 8:
            raise KvError("elem_size must be greater than 0 and less or equal to 8 bytes")
        if src_tp_size >= dst_tp_size:
            raise KvError("src_tp_size must be less than dst_tp_size")
        if src_tp_size & (src_tp_size - 1) != 0:
            raise KvError("src_tp_size must be a power of 2")
        if dst_tp_size & (dst_tp_size - 1) != 0:
            raise KvError("dst_tp_size must be a power of 2")
        scatter_factor = dst_tp_size // src_tp_size
        if not self.staged_layers:
            raise KvError("Layers must be loaded before performing permutation")
        if not self.staged_block_ids:
            raise KvError("Block IDs must be loaded before performing permutation")
        n = len(self.layer_details.src_layer_ptrs)
        if layer >= n:
            raise KvError(f"layer index {layer} out of bounds (max {n - 1})")
        dims6 = np.array([dims[0], dims[1], dims[2], scatter_factor, dims[3] // scatter_factor, dims[4]],
                         dtype=np.uint32)
        d = self.layer_details
        rc = _lib.lib().copy_stream_scatter(self._handle, C.c_void_p(d.src_layer_ptrs[layer]),
                                            C.c_void_p(d.dst_layer_ptrs[layer]), C.c_void_p(dims6.ctypes.data), 6,
                                            int(elem_size), int(block_dim_index), self.src_block_dim(),
                                            self.dst_block_dim())
        if rc != 0:
            raise KvError("Failed to execute tensor permutation")

    def reset(self) -> None:
        self.staged_block_ids = False
        self.staged_layers = False

    def on_return(self) -> None:
        """`Returnable::on_return` (kv/layer.rs:1138-1142): pools call this when the stream is handed back."""
        self.reset()

    def close(self) -> None:
        # the reference has no Drop and leaks the C handle; the mirror releases it
        if getattr(self, "_handle", None) is not None and self._handle.value:
            _lib.lib().copy_stream_destroy(self._handle)
            self._handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
