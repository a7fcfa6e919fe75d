"""CPU: host-side mirror of kv::layer / kv::storage math (no device calls). Values follow the reference's unit tests
(lib/llm/src/kv/storage.rs:1260-1516) and layout definitions (lib/llm/src/kv/layer.rs:140-212,322-339)."""
import numpy as np
import pytest

from dynamo_exp_b200.layer import CopyStreamBlockMap, KvBlockDetails, KvLayout, KvModelDetails
from dynamo_exp_b200.storage import DType, KvError, Storage, StorageType, TensorView, calculate_strides, unflatten_index


class MockTensor(Storage):
    """kv/storage.rs:1238-1258: data_ptr 0x1000, caller-chosen size."""

    def __init__(self, size):
        self._size = size

    def get_pointer(self):
        return 0x1000

    def storage_size(self):
        return self._size

    def storage_type(self):
        return StorageType.System


def test_dtype_sizes():
    want = {"F32": 4, "F16": 2, "BF16": 2, "FP8": 1, "U8": 1, "U16": 2, "U32": 4, "U64": 8, "I8": 1, "I16": 2,
            "I32": 4, "I64": 8}
    assert {d.name: d.size_in_bytes() for d in DType} == want


def test_tensor_view_strides_offsets_addresses():
    t = MockTensor(2 * 3 * 4 * 4)
    v = TensorView(t, [2, 3, 4], 4)
    assert v.strides() == [12, 4, 1]
    assert v.byte_strides() == [48, 16, 4]
    assert v.byte_offset([0, 0, 0]) == 0
    assert v.byte_offset([1, 2, 3]) == 48 + 32 + 12
    assert v.address([1, 2, 3]) == 0x1000 + 92
    assert v.flat_index([1, 1, 1]) == 17
    assert v.num_elements() == 24 and v.size_in_bytes() == 96
    assert v.is_contiguous()
    assert v.in_bounds([1, 2, 3]) and not v.in_bounds([2, 0, 0])
    with pytest.raises(KvError):
        v.byte_offset([2, 0, 0])


def test_tensor_view_too_large_for_storage():
    with pytest.raises(KvError, match="requires 96 bytes"):
        TensorView(MockTensor(95), [2, 3, 4], 4)


def test_slice_semantics():
    v = TensorView(MockTensor(2 * 3 * 4 * 4), [2, 3, 4], 4)
    s = v.slice(1, 1, 3)
    assert s.shape() == [2, 2, 4] and s.offset == 16 and s.strides() == [12, 4, 1]
    assert not s.is_contiguous()
    assert s.address([0, 0, 0]) == 0x1000 + 16
    s0 = v.slice(0, 1)
    assert s0.shape() == [1, 3, 4] and s0.offset == 48 and s0.is_contiguous()
    for bad in ((3, 0, 1), (1, 0, 4), (1, 2, 2)):
        with pytest.raises(KvError):
            v.slice(*bad)


def test_with_strides_bounds():
    t = MockTensor(100)
    v = TensorView.with_strides(t, [2, 2], [8, 2], 4, 4)
    assert v.byte_strides() == [32, 8] and v.byte_offset([1, 1]) == 4 + 40
    with pytest.raises(KvError, match="storage size is only"):
        TensorView.with_strides(t, [4, 4], [8, 2], 0, 4)


def test_stride_helpers():
    assert calculate_strides([2, 3, 4]) == [12, 4, 1]
    assert unflatten_index(17, [2, 3, 4]) == [1, 1, 1]
    assert unflatten_index(23, [2, 3, 4]) == [1, 2, 3]


def test_h2d_d2h_reject_wrong_sides_before_any_cuda_call():
    host = TensorView(MockTensor(64), [4, 4], 4)
    with pytest.raises(KvError, match="Destination must be a device tensor"):
        host.h2d(host)
    with pytest.raises(KvError, match="Source must be a device tensor"):
        host.d2h(host)


def llama8b(layout, tp=1, rank=0):
    md = KvModelDetails(number_of_layers=32, number_of_heads=8, head_size=128, dtype=DType.BF16)
    return KvBlockDetails(layout=layout, block_size=64, model_details=md, tp_rank=rank, tp_size=tp)


def test_block_details_llama8b_shapes():
    kv = llama8b(KvLayout.KvFirst)
    assert kv.prefix_dim() == 2 and kv.suffix_dim() == 65536 and kv.elem_size() == 2
    assert kv.bytes_per_token_block_per_layer() == 256 * 1024
    bf = llama8b(KvLayout.BlockFirst)
    assert bf.prefix_dim() == 1 and bf.suffix_dim() == 131072
    assert kv.model_details.bytes_per_token_per_layer() == 4096
    tp4 = llama8b(KvLayout.KvFirst, tp=4, rank=3)
    tp4.validate()
    assert tp4.suffix_dim() == 16384 and tp4.bytes_per_token_block_per_layer() == 64 * 1024
    assert not kv.is_compatible(bf) and not kv.is_compatible(tp4) and kv.is_compatible(llama8b(KvLayout.KvFirst))


def test_block_details_validation():
    with pytest.raises(KvError, match="evenly divide"):
        llama8b(KvLayout.KvFirst, tp=3).validate()
    with pytest.raises(KvError, match="tp_rank"):
        llama8b(KvLayout.KvFirst, tp=2, rank=2).validate()
    with pytest.raises(KvError):
        llama8b(KvLayout.KvFirst, tp=16).validate()


def test_block_map_validation():
    with pytest.raises(KvError, match="src_layer_ptrs must not be empty"):
        CopyStreamBlockMap([], [], 1, 1, 1, 1, 2).validate()
    with pytest.raises(KvError, match="elem_size"):
        CopyStreamBlockMap([1], [2], 1, 1, 1, 1, 9).validate()
    with pytest.raises(KvError, match="same length"):
        CopyStreamBlockMap([1, 2], [2], 1, 1, 1, 1, 2).validate()
    CopyStreamBlockMap([1], [2], 2, 4, 4, 8, 2).validate()


# ---- the reference's own TensorView unit tests, value for value (kv/storage.rs:1260-1516) ----
def test_ref_tensor_view_creation_and_indexing():
    v = TensorView(MockTensor(96), [2, 3, 4], 4)                      # :1260-1279
    assert v.shape() == [2, 3, 4] and v.strides() == [12, 4, 1] and v.byte_strides() == [48, 16, 4]
    assert v.num_elements() == 24 and v.size_in_bytes() == 96 and v.is_contiguous()
    for idx, flat in (([0, 0, 0], 0), ([0, 0, 1], 1), ([0, 1, 0], 4), ([1, 0, 0], 12), ([1, 2, 3], 23)):   # :1295-1299
        assert v.flat_index(idx) == flat
    for idx, off in (([0, 0, 0], 0), ([0, 0, 1], 4), ([0, 1, 0], 16), ([1, 0, 0], 48)):                    # :1302-1305
        assert v.byte_offset(idx) == off
    assert v.address([0, 0, 0]) == 0x1000 and v.address([0, 0, 1]) == 0x1004                               # :1308-1310
    assert v.address([1, 2, 3]) == 0x1000 + 23 * 4


def test_ref_tensor_view_slicing():
    v = TensorView(MockTensor(96), [2, 3, 4], 4)                      # :1313-1343
    s = v.slice(1, 1, 3)
    assert s.shape() == [2, 2, 4] and s.strides() == [12, 4, 1] and s.byte_strides() == [48, 16, 4] and s.offset == 16
    assert s.address([0, 0, 0]) == 0x1000 + 16
    assert s.address([1, 1, 3]) == 0x1000 + 16 + 48 + 16 + 12


def test_ref_custom_strides_and_offsets():
    t = MockTensor(96)                                                # :1345-1406
    nc = TensorView.with_strides(t, [2, 2, 4], [12, 4, 1], 0, 4)
    assert not nc.is_contiguous() and nc.strides() == [12, 4, 1] and nc.byte_strides() == [48, 16, 4]
    assert nc.byte_offset([1, 1, 3]) == (12 + 4 + 3) * 4 < t.storage_size()
    with pytest.raises(KvError, match="would access up to byte offset 108"):
        TensorView.with_strides(t, [2, 3, 4], [16, 4, 1], 0, 4)
    t2 = MockTensor(120)                                              # :1408-1454
    ov = TensorView.with_strides(t2, [2, 3, 4], [12, 4, 1], 16, 4)
    assert ov.is_contiguous() and ov.offset == 16
    assert ov.byte_offset([0, 0, 0]) == 16 and ov.byte_offset([1, 2, 3]) == 16 + (12 + 2 * 4 + 3) * 4
    with pytest.raises(KvError, match="would access up to byte offset"):
        TensorView.with_strides(t2, [2, 3, 4], [12, 4, 1], 80, 4)


def test_ref_in_bounds_validate_and_indices_iter():
    v = TensorView(MockTensor(96), [2, 3, 4], 4)                      # :1456-1493
    assert v.in_bounds([0, 0, 0]) and v.in_bounds([1, 2, 3])
    for bad in ([2, 0, 0], [0, 3, 0], [0, 0, 4], [2, 3, 4]):
        assert not v.in_bounds(bad)
    for bad in ([2, 0, 0], [0, 3, 0], [0, 0, 4]):
        with pytest.raises(KvError):
            v.byte_offset(bad)
    v2 = TensorView(MockTensor(24), [2, 3], 4)                        # :1495-1514
    assert list(v2.indices_iter()) == [[0, 0], [0, 1], [0, 2], [1, 0], [1, 1], [1, 2]]
