// kvdp.hpp — header-only C++17 host layer over the C ABI in kvdp.h.
//
// The reference's host code for this path is Rust (lib/llm/src/kv/storage.rs, lib/llm/src/kv/layer.rs,
// lib/tokens/src/lib.rs). No Rust toolchain exists in the build image, so this header restates that interface in C++
// with the same type and method names, argument meaning, state machine and error behaviour; `Result<T>`/`Err(..)`
// becomes a thrown kvdp::Error. It owns no algorithm: every byte is moved and every hash computed by libkvdp.so.
#pragma once
#include <algorithm>
#include <array>
#include <cstdint>
#include <cstring>
#include <memory>
#include <numeric>
#include <optional>
#include <set>
#include <stdexcept>
#include <string>
#include <vector>

#include "kvdp.h"

namespace kvdp {

struct Error : std::runtime_error {
  using std::runtime_error::runtime_error;
};
[[noreturn]] inline void raise(const std::string& m) { throw Error(m); }

// ------------------------------------------------------------------------------------------------ kv::storage
// reference: kv/storage.rs:47-52
struct StorageType {
  enum Kind { Device, Pinned, System } kind;
  int device = -1;
  static StorageType device_(int d) { return {Device, d}; }
  static StorageType pinned() { return {Pinned, -1}; }
  static StorageType system() { return {System, -1}; }
  bool operator==(const StorageType& o) const { return kind == o.kind && (kind != Device || device == o.device); }
  bool operator!=(const StorageType& o) const { return !(*this == o); }
};

// reference: kv/storage.rs:55-89
enum class DType { F32, F16, BF16, FP8, U8, U16, U32, U64, I8, I16, I32, I64 };
inline size_t size_in_bytes(DType t) {
  switch (t) {
    case DType::F32: case DType::U32: case DType::I32: return 4;
    case DType::F16: case DType::BF16: case DType::U16: case DType::I16: return 2;
    case DType::FP8: case DType::U8: case DType::I8: return 1;
    default: return 8;
  }
}

// reference: trait Storage, kv/storage.rs:103-124
struct Storage {
  virtual ~Storage() = default;
  virtual uint64_t get_pointer() const = 0;
  virtual size_t storage_size() const = 0;
  virtual StorageType storage_type() const = 0;
};

// reference: kv/storage.rs:189-230 (zero-filled device slab)
class DeviceStorageOwned : public Storage {
 public:
  DeviceStorageOwned(size_t bytes, int device) : bytes_(bytes), device_(device) {
    if (kvdp_device_alloc(&ptr_, bytes, device) != 0) raise(std::string("device allocation failed: ") + kvdp_last_error());
  }
  ~DeviceStorageOwned() override { kvdp_device_free(ptr_); }
  uint64_t get_pointer() const override { return (uint64_t)(uintptr_t)ptr_; }
  size_t storage_size() const override { return bytes_; }
  StorageType storage_type() const override { return StorageType::device_(device_); }

 private:
  void* ptr_ = nullptr;
  size_t bytes_;
  int device_;
};

// reference: kv/storage.rs:242-315 (pinned slab, zeroed by the CPU)
class CudaPinnedMemory : public Storage {
 public:
  explicit CudaPinnedMemory(size_t bytes) : bytes_(bytes) {
    if (bytes == 0) raise("Bytes must be greater than 0");
    if (cuda_malloc_host(&ptr_, bytes) != 0 || !ptr_) raise("cudaMallocHost failed");
    std::memset(ptr_, 0, bytes);
  }
  ~CudaPinnedMemory() override { cuda_free_host(ptr_); }
  void* as_mut_ptr() { return ptr_; }
  const void* as_ptr() const { return ptr_; }
  size_t size() const { return bytes_; }
  uint64_t get_pointer() const override { return (uint64_t)(uintptr_t)ptr_; }
  size_t storage_size() const override { return bytes_; }
  StorageType storage_type() const override { return StorageType::pinned(); }

 private:
  void* ptr_ = nullptr;
  size_t bytes_;
};

// reference: kv/storage.rs:1171-1232 (bring-your-own device pointer kept alive by `source`)
class DeviceStorageFromAny : public Storage {
 public:
  DeviceStorageFromAny(std::shared_ptr<void> source, uint64_t ptr, size_t bytes, int device)
      : source_(std::move(source)), ptr_(ptr), bytes_(bytes), device_(device) {}
  uint64_t get_pointer() const override { return ptr_; }
  size_t storage_size() const override { return bytes_; }
  StorageType storage_type() const override { return StorageType::device_(device_); }

 private:
  std::shared_ptr<void> source_;
  uint64_t ptr_;
  size_t bytes_;
  int device_;
};

// reference: kv/storage.rs:126-187
class OwnedStorage : public Storage {
 public:
  explicit OwnedStorage(std::shared_ptr<Storage> s) : storage_(std::move(s)) {}
  static OwnedStorage create(size_t bytes, StorageType t) {
    switch (t.kind) {
      case StorageType::Device: return create_device_array(bytes, t.device);
      case StorageType::Pinned: return create_pinned_array(bytes);
      default: raise("System memory not yet supported");
    }
  }
  static OwnedStorage create_device_array(size_t bytes, int device) {
    return OwnedStorage(std::make_shared<DeviceStorageOwned>(bytes, device));
  }
  static OwnedStorage create_pinned_array(size_t bytes) { return OwnedStorage(std::make_shared<CudaPinnedMemory>(bytes)); }
  static OwnedStorage byo_device_array(uint64_t ptr, size_t bytes, int device, std::shared_ptr<void> owner) {
    return OwnedStorage(std::make_shared<DeviceStorageFromAny>(std::move(owner), ptr, bytes, device));
  }
  uint64_t get_pointer() const override { return storage_->get_pointer(); }
  size_t storage_size() const override { return storage_->storage_size(); }
  StorageType storage_type() const override { return storage_->storage_type(); }

 private:
  std::shared_ptr<Storage> storage_;
};

// reference: kv/storage.rs:329-1110
template <size_t D>
class TensorView {
 public:
  TensorView(const Storage& storage, std::array<size_t, D> shape, size_t element_size)
      : storage_(&storage), shape_(shape), element_size_(element_size) {
    if (D > 0) {
      strides_[D - 1] = 1;
      for (size_t i = D - 1; i-- > 0;) strides_[i] = strides_[i + 1] * shape_[i + 1];
    }
    for (size_t i = 0; i < D; ++i) byte_strides_[i] = strides_[i] * element_size;
    total_ = std::accumulate(shape_.begin(), shape_.end(), (size_t)1, std::multiplies<size_t>());
    if (total_ * element_size > storage.storage_size())
      raise("Shape requires " + std::to_string(total_ * element_size) + " bytes, but storage only has " +
            std::to_string(storage.storage_size()) + " bytes");
  }
  const std::array<size_t, D>& shape() const { return shape_; }
  const std::array<size_t, D>& strides() const { return strides_; }
  const std::array<size_t, D>& byte_strides() const { return byte_strides_; }
  size_t element_size() const { return element_size_; }
  size_t num_elements() const { return total_; }
  size_t size_in_bytes() const { return total_ * element_size_; }
  size_t offset() const { return offset_; }
  uint64_t data() const { return storage_->get_pointer(); }
  StorageType storage_type() const { return storage_->storage_type(); }
  size_t byte_offset(const std::array<size_t, D>& idx) const {
    size_t off = offset_;
    for (size_t i = 0; i < D; ++i) {
      if (idx[i] >= shape_[i])
        raise("Index " + std::to_string(idx[i]) + " out of bounds for dimension " + std::to_string(i) + " with size " +
              std::to_string(shape_[i]));
      off += idx[i] * byte_strides_[i];
    }
    return off;
  }
  uint64_t address(const std::array<size_t, D>& idx) const { return data() + byte_offset(idx); }
  bool is_contiguous() const {
    size_t es = 1, eb = element_size_;
    for (size_t i = D; i-- > 0;) {
      if (strides_[i] != es || byte_strides_[i] != eb) return false;
      es *= shape_[i];
      eb *= shape_[i];
    }
    return true;
  }
  TensorView slice(size_t dim, size_t start, std::optional<size_t> end = std::nullopt) const {
    if (dim >= D) raise("Dimension out of bounds");
    size_t e = end.value_or(shape_[dim]);
    if (e > shape_[dim]) raise("End index out of bounds");
    if (start >= e) raise("Invalid slice range");
    TensorView v = *this;
    v.shape_[dim] = e - start;
    v.offset_ = offset_ + start * byte_strides_[dim];
    v.total_ = std::accumulate(v.shape_.begin(), v.shape_.end(), (size_t)1, std::multiplies<size_t>());
    return v;
  }
  // kv/storage.rs:915-982
  void h2d(TensorView& device_view, kvdp_stream_t stream) const {
    if (storage_type().kind == StorageType::Device) raise("Source must be a host tensor (System or Pinned)");
    if (device_view.storage_type().kind != StorageType::Device) raise("Destination must be a device tensor");
    pair_checks(device_view);
    if (cuda_memcpy_async((void*)(uintptr_t)(device_view.data() + device_view.offset_),
                          (const void*)(uintptr_t)(data() + offset_), size_in_bytes(), stream) != 0)
      raise("cudaMemcpyAsync failed during host-to-device transfer");
  }
  // kv/storage.rs:984-1050
  void d2h(TensorView& host_view, kvdp_stream_t stream) const {
    if (storage_type().kind != StorageType::Device) raise("Source must be a device tensor");
    if (host_view.storage_type().kind == StorageType::Device) raise("Destination must be a host tensor (System or Pinned)");
    pair_checks(host_view);
    if (cuda_memcpy_async((void*)(uintptr_t)(host_view.data() + host_view.offset_),
                          (const void*)(uintptr_t)(data() + offset_), size_in_bytes(), stream) != 0)
      raise("cudaMemcpyAsync failed during device-to-host transfer");
  }
  // kv/storage.rs:637-675
  void copy_to_view_blocking(TensorView& dst) const {
    if (shape_ != dst.shape_ || strides_ != dst.strides_) raise("Shape or strides mismatch");
    if (!is_contiguous()) raise("Source is not contiguous");
    if (!dst.is_contiguous()) raise("Destination is not contiguous");
    if (cuda_memcpy_sync((void*)(uintptr_t)dst.data(), (const void*)(uintptr_t)data(), size_in_bytes()) != 0)
      raise("cudaMemcpyAsync failed");
  }

 private:
  void pair_checks(const TensorView& o) const {
    if (shape_ != o.shape_) raise("Shape mismatch");
    if (element_size_ != o.element_size_) raise("Element size mismatch");
    if (!is_contiguous()) raise("Source tensor must be contiguous");
    if (!o.is_contiguous()) raise("Destination tensor must be contiguous");
  }
  const Storage* storage_;
  std::array<size_t, D> shape_{}, strides_{}, byte_strides_{};
  size_t offset_ = 0, element_size_ = 0, total_ = 0;
};

// ------------------------------------------------------------------------------------------------ kv::layer
enum class KvLayout { KvFirst, BlockFirst };  // kv/layer.rs:99-106

struct KvModelDetails {  // kv/layer.rs:108-138
  size_t number_of_layers, number_of_heads, head_size;
  DType dtype;
  size_t number_of_elements_per_token_per_layer() const { return 2 * number_of_heads * head_size; }
  size_t bytes_per_token_per_layer() const { return number_of_elements_per_token_per_layer() * size_in_bytes(dtype); }
  bool operator==(const KvModelDetails& o) const {
    return number_of_layers == o.number_of_layers && number_of_heads == o.number_of_heads && head_size == o.head_size &&
           dtype == o.dtype;
  }
};

struct KvBlockDetails {  // kv/layer.rs:140-212
  KvLayout layout;
  size_t block_size;
  KvModelDetails model_details;
  size_t tp_rank = 0, tp_size = 1;
  void validate() const {
    if (tp_size == 0 || model_details.number_of_heads % tp_size != 0) raise("tp_size must evenly divide num_heads");
    if (tp_rank >= tp_size) raise("tp_rank must be less than tp_size");
    if (tp_size > model_details.number_of_heads) raise("tp_size must be less than num_heads");
  }
  size_t bytes_per_token_block_per_layer() const { return model_details.bytes_per_token_per_layer() * block_size / tp_size; }
  bool is_compatible(const KvBlockDetails& o) const {
    return layout == o.layout && block_size == o.block_size && tp_size == o.tp_size && model_details == o.model_details;
  }
  size_t prefix_dim() const { return layout == KvLayout::KvFirst ? 2 : 1; }
  size_t suffix_dim() const {
    size_t s = block_size * (model_details.number_of_heads / tp_size) * model_details.head_size;
    return layout == KvLayout::KvFirst ? s : 2 * s;
  }
  size_t elem_size() const { return size_in_bytes(model_details.dtype); }
};

// kv/layer.rs:214-451
class KvLayer : public Storage {
 public:
  KvLayer(const KvBlockDetails& bd, size_t number_of_blocks, OwnedStorage storage)
      : layout(bd.layout), storage(std::move(storage)), number_of_blocks(number_of_blocks), block_size(bd.block_size),
        number_of_heads(bd.model_details.number_of_heads), head_size(bd.model_details.head_size),
        dtype(bd.model_details.dtype), tp_size(bd.tp_size), tp_rank(bd.tp_rank) {
    if (!number_of_blocks || !block_size || !number_of_heads || !head_size) raise("layer dimensions must be at least 1");
    if (number_of_heads % tp_size != 0) raise("number_of_heads must be divisible by tp_size");
    if (tp_rank >= tp_size) raise("tp_rank must be less than tp_size");
    auto s = layer_shape();
    size_t bytes = s[0] * s[1] * s[2] * s[3] * s[4] * size_in_bytes(dtype);
    if (this->storage.storage_size() < bytes) raise("storage must be at least as large as the layer");
  }
  std::array<size_t, 5> layer_shape() const {
    size_t h = number_of_heads / tp_size;
    if (layout == KvLayout::KvFirst) return {2, number_of_blocks, block_size, h, head_size};
    return {number_of_blocks, 2, block_size, h, head_size};
  }
  TensorView<5> view() const { return TensorView<5>(*this, layer_shape(), size_in_bytes(dtype)); }
  uint64_t get_pointer() const override { return storage.get_pointer(); }
  size_t storage_size() const override { return storage.storage_size(); }
  StorageType storage_type() const override { return storage.storage_type(); }
  // kv/layer.rs:371-450 -> copy_blocks_3d
  void copy_blocks_to(const std::vector<size_t>& src_block_ids, KvLayer& dst, const std::vector<size_t>& dst_block_ids) const {
    if (src_block_ids.size() != dst_block_ids.size()) raise("src_block_ids and dst_block_ids must have the same length");
    if (layout != dst.layout) raise("src and dst must have the same layout");
    auto s = storage_type().kind, d = dst.storage_type().kind;
    if (s == StorageType::Pinned && d == StorageType::Pinned) raise("Pinned to Pinned copy not implemented");
    if (s == StorageType::Device && d == StorageType::Device) raise("Device to Device copy not implemented");
    if (s == StorageType::System) raise("System to Device copy not implemented");
    if (d == StorageType::System) raise("Device to System copy not implemented");
    std::vector<int> hs(src_block_ids.begin(), src_block_ids.end()), hd(dst_block_ids.begin(), dst_block_ids.end());
    int prefix = layout == KvLayout::KvFirst ? 2 : 1;
    size_t suffix = head_size * (number_of_heads / tp_size) * block_size * (layout == KvLayout::KvFirst ? 1 : 2);
    int rc = copy_blocks_3d((const void*)(uintptr_t)get_pointer(), (void*)(uintptr_t)dst.get_pointer(), hs.data(), hd.data(),
                            (int)hs.size(), prefix, (int)number_of_blocks, (int)dst.number_of_blocks, (int)suffix,
                            (int)size_in_bytes(dtype));
    if (rc != 0) raise("Failed to copy blocks");
  }
  KvLayout layout;
  OwnedStorage storage;
  size_t number_of_blocks, block_size, number_of_heads, head_size;
  DType dtype;
  size_t tp_size, tp_rank;
};

// kv/layer.rs:453-596
class KvBlockStorage {
 public:
  static KvBlockStorage from_layers(std::vector<KvLayer> layers) {
    if (layers.empty()) raise("Layers must not be empty");
    auto st = layers[0].storage_type();
    size_t nb = layers[0].number_of_blocks;
    for (auto& l : layers) {
      if (l.storage_type() != st) raise("All layers must have the same storage type");
      if (l.number_of_blocks != nb) raise("All layers must have the same number of blocks");
    }
    KvBlockDetails bd{layers[0].layout, layers[0].block_size,
                      KvModelDetails{layers.size(), layers[0].number_of_heads, layers[0].head_size, layers[0].dtype},
                      layers[0].tp_rank, layers[0].tp_size};
    bd.validate();
    for (auto& l : layers)
      if (l.storage_size() < bd.bytes_per_token_block_per_layer())
        raise("All layers must have enough capacity to store hold the block data");
    return KvBlockStorage(bd, st, nb, std::move(layers));
  }
  static KvBlockStorage allocate(size_t number_of_blocks, const KvBlockDetails& bd, StorageType st) {
    bd.validate();
    size_t bytes = bd.bytes_per_token_block_per_layer() * number_of_blocks;
    std::vector<KvLayer> layers;
    for (size_t l = 0; l < bd.model_details.number_of_layers; ++l)
      layers.emplace_back(bd, number_of_blocks, OwnedStorage::create(bytes, st));
    return from_layers(std::move(layers));
  }
  KvLayer& layer(size_t i) {
    if (i >= layers.size()) raise("Layer index " + std::to_string(i) + " out of bounds (max " + std::to_string(layers.size() - 1) + ")");
    return layers[i];
  }
  size_t number_of_blocks() const { return number_of_blocks_; }
  StorageType storage_type() const { return storage_type_; }
  KvBlockDetails block_details;
  std::vector<KvLayer> layers;

 private:
  KvBlockStorage(KvBlockDetails bd, StorageType st, size_t nb, std::vector<KvLayer> l)
      : block_details(bd), layers(std::move(l)), storage_type_(st), number_of_blocks_(nb) {}
  StorageType storage_type_;
  size_t number_of_blocks_;
};

// kv/layer.rs:592-709
struct CopyStreamBlockMap {
  std::vector<uint64_t> src_layer_ptrs, dst_layer_ptrs;
  int prefix_dim = 0, src_block_dim = 0, dst_block_dim = 0, suffix_dim = 0, elem_size = 0;
  static std::shared_ptr<CopyStreamBlockMap> make(const KvBlockStorage& src, const KvBlockStorage& dst) {
    if (!src.block_details.is_compatible(dst.block_details)) raise("src and dst must have compatible block details");
    auto m = std::make_shared<CopyStreamBlockMap>();
    for (auto& l : src.layers) m->src_layer_ptrs.push_back(l.get_pointer());
    for (auto& l : dst.layers) m->dst_layer_ptrs.push_back(l.get_pointer());
    if (m->src_layer_ptrs.size() != m->dst_layer_ptrs.size()) raise("src and dst must have the same number of layers");
    m->prefix_dim = (int)src.block_details.prefix_dim();
    m->suffix_dim = (int)src.block_details.suffix_dim();
    m->elem_size = (int)src.block_details.elem_size();
    m->src_block_dim = (int)src.number_of_blocks();
    m->dst_block_dim = (int)dst.number_of_blocks();
    m->validate();
    return m;
  }
  void validate() const {
    if (prefix_dim < 1 || src_block_dim < 1 || dst_block_dim < 1 || suffix_dim < 1) raise("block map dims must be at least 1");
    if (elem_size < 1 || elem_size > 8) raise("elem_size must be between 1 and 8");
    if (src_layer_ptrs.empty()) raise("src_layer_ptrs must not be empty");
    if (dst_layer_ptrs.empty()) raise("dst_layer_ptrs must not be empty");
    if (src_layer_ptrs.size() != dst_layer_ptrs.size()) raise("src_layer_ptrs and dst_layer_ptrs must have the same length");
  }
};

// kv/layer.rs:711-1143: prepare_block_map -> prepare_block_ids -> trigger_layer* -> sync_stream -> reset
class CopyStream {
 public:
  CopyStream(size_t num_layers, size_t num_blocks) : max_num_layers(num_layers), max_num_blocks(num_blocks) {
    if (copy_stream_create(&handle_, (int)num_layers, (int)num_blocks) != 0 || !handle_) raise("Failed to create copy stream");
    layer_doorbells.assign(num_layers, false);
    layer_details = std::make_shared<CopyStreamBlockMap>();
  }
  ~CopyStream() { copy_stream_destroy(handle_); }  // the reference has no Drop
  CopyStream(const CopyStream&) = delete;
  CopyStream& operator=(const CopyStream&) = delete;
  void prepare_block_map(std::shared_ptr<CopyStreamBlockMap> details) {
    size_t n = details->src_layer_ptrs.size();
    if (max_num_layers < n)
      raise("Number of layers " + std::to_string(n) + " exceeds max number of layers " + std::to_string(max_num_layers));
    if (staged_layers) raise("Layers already loaded");
    staged_layers = true;
    layer_details = std::move(details);
    std::fill(layer_doorbells.begin(), layer_doorbells.end(), false);
  }
  void prepare_block_ids(std::vector<int> src, std::vector<int> dst) {
    if (src.size() != dst.size()) raise("src_block_ids and dst_block_ids must have the same length");
    if (check_unique_ids) {  // debug_assertions in the reference (kv/layer.rs:848-861)
      if (std::set<int>(dst.begin(), dst.end()).size() != dst.size()) raise("dst_block_ids must be unique");
      if (std::set<int>(src.begin(), src.end()).size() != src.size()) raise("src_block_ids must be unique");
    }
    if (max_num_blocks < src.size())
      raise("Number of blocks " + std::to_string(src.size()) + " exceeds max number of blocks " + std::to_string(max_num_blocks));
    if (!staged_layers) raise("Layers must be loaded before preparing block ids");
    if (staged_block_ids) raise("Block ids already loaded");
    src_block_ids = std::move(src);
    dst_block_ids = std::move(dst);
    if (copy_stream_prepare_block_ids(handle_, src_block_ids.data(), dst_block_ids.data(), (int)src_block_ids.size()) != 0)
      raise("Failed to prepare block ids");
    staged_block_ids = true;
  }
  void trigger_layer(size_t layer) {
    if (!staged_layers) raise("Layers must be loaded before triggering a layer");
    if (!staged_block_ids) raise("Block ids must be loaded before triggering a layer");
    size_t n = layer_details->src_layer_ptrs.size();
    if (layer >= n) raise("layer index " + std::to_string(layer) + " out of bounds (max " + std::to_string(n - 1) + ")");
    if (layer_doorbells[layer]) return;
    const auto& d = *layer_details;
    if (copy_stream_memcpy(handle_, (const void*)(uintptr_t)d.src_layer_ptrs[layer], (void*)(uintptr_t)d.dst_layer_ptrs[layer],
                           d.prefix_dim, d.suffix_dim, d.elem_size, d.src_block_dim, d.dst_block_dim) != 0)
      raise("Failed to execute layer " + std::to_string(layer) + " copy");
    layer_doorbells[layer] = true;
  }
  size_t layer_count() const { return layer_details->src_layer_ptrs.size(); }
  void trigger_all_layers() {
    for (size_t l = 0; l < layer_count(); ++l) trigger_layer(l);
  }
  void sync_stream() {
    if (copy_stream_sync(handle_) != 0) raise("Failed to synchronize copy stream");
  }
  // kv/layer.rs:1027-1130
  void scatter_copy_layer(size_t layer, const std::vector<size_t>& dims, size_t elem_size, size_t block_dim_index,
                          size_t src_tp_size, size_t dst_tp_size) {
    if (dims.size() != 5) raise("Expected 5 dimensions for src_data");
    if (block_dim_index > 1) raise("block_dim_index must be 0 or 1");
    if (elem_size == 0 || elem_size > 8) raise("elem_size must be greater than 0 and less or equal to 8 bytes");
    if (src_tp_size >= dst_tp_size) raise("src_tp_size must be less than dst_tp_size");
    if (src_tp_size & (src_tp_size - 1)) raise("src_tp_size must be a power of 2");
    if (dst_tp_size & (dst_tp_size - 1)) raise("dst_tp_size must be a power of 2");
    size_t sf = dst_tp_size / src_tp_size;
    if (!staged_layers) raise("Layers must be loaded before performing permutation");
    if (!staged_block_ids) raise("Block IDs must be loaded before performing permutation");
    size_t n = layer_details->src_layer_ptrs.size();
    if (layer >= n) raise("layer index out of bounds");
    uint32_t d6[6] = {(uint32_t)dims[0], (uint32_t)dims[1], (uint32_t)dims[2], (uint32_t)sf, (uint32_t)(dims[3] / sf),
                      (uint32_t)dims[4]};
    const auto& d = *layer_details;
    if (copy_stream_scatter(handle_, (const void*)(uintptr_t)d.src_layer_ptrs[layer], (void*)(uintptr_t)d.dst_layer_ptrs[layer],
                            d6, 6, (uint32_t)elem_size, (uint32_t)block_dim_index, (uint32_t)d.src_block_dim,
                            (uint32_t)d.dst_block_dim) != 0)
      raise("Failed to execute tensor permutation");
  }
  void reset() { staged_block_ids = staged_layers = false; }
  void on_return() { reset(); }  // Returnable (kv/layer.rs:1138-1142)

  size_t max_num_layers, max_num_blocks;
  bool staged_layers = false, staged_block_ids = false, check_unique_ids = true;
  std::vector<bool> layer_doorbells;
  std::vector<int> src_block_ids, dst_block_ids;
  std::shared_ptr<CopyStreamBlockMap> layer_details;

 private:
  ::CopyStream* handle_ = nullptr;
};

// ------------------------------------------------------------------------------------------------ lib/tokens
using Token = uint32_t;
using SaltHash = uint64_t;
using BlockHash = uint64_t;
using SequenceHash = uint64_t;
constexpr uint64_t XXH3_SEED = 1337;  // kv_router/indexer.rs:64

// lib/tokens/src/lib.rs:44-46, on the device
inline uint64_t compute_hash(const void* data, size_t len, uint64_t seed) {
  void *d_in = nullptr, *d_out = nullptr;
  uint64_t out = 0;
  if (kvdp_device_alloc(&d_in, len ? len : 1, -1) || kvdp_device_alloc(&d_out, 8, -1)) raise(kvdp_last_error());
  int rc = len ? cuda_memcpy_sync(d_in, data, len) : 0;
  if (!rc) rc = kvdp_xxh3_64_batch((const uint8_t*)d_in, 1, (int64_t)len, seed, (uint64_t*)d_out, nullptr);
  if (!rc) rc = cuda_memcpy_sync(&out, d_out, 8);
  kvdp_device_free(d_in);
  kvdp_device_free(d_out);
  if (rc) raise(kvdp_last_error());
  return out;
}

struct TokenBlock {  // lib/tokens/src/lib.rs:217-262
  std::vector<Token> tokens;
  uint64_t salt_hash = 0, block_hash = 0, sequence_hash = 0;
  std::optional<uint64_t> parent_sequence_hash;
};

// batched split_tokens for one request: (block_hash, sequence_hash) per full block
inline void hash_blocks(const std::vector<Token>& t, size_t block_size, uint64_t salt, std::vector<uint64_t>& bh,
                        std::vector<uint64_t>& sh) {
  int64_t offs[2] = {0, (int64_t)t.size()}, boff[2] = {0, 0};
  size_t nb = block_size ? t.size() / block_size : 0;
  bh.assign(nb, 0);
  sh.assign(nb, 0);
  if (kvdp_hash_token_blocks_host(t.data(), offs, 1, (int32_t)block_size, salt, boff, bh.data(), sh.data(), 0, -1) != 0)
    raise(kvdp_last_error());
}

class PartialTokenBlock {  // lib/tokens/src/lib.rs:139-187
 public:
  PartialTokenBlock(std::vector<Token> tokens, size_t block_size, uint64_t salt, std::optional<uint64_t> parent)
      : tokens_(std::move(tokens)), block_size_(block_size), salt_hash_(salt), parent_sequence_hash_(parent) {}
  std::optional<TokenBlock> push_token(Token token) {
    if (tokens_.size() >= block_size_) raise("assertion failed: tokens.len() < block_size");
    tokens_.push_back(token);
    if (tokens_.size() != block_size_) return std::nullopt;
    std::vector<uint64_t> bh, sh;
    hash_blocks(tokens_, block_size_, salt_hash_, bh, sh);
    TokenBlock b;
    b.tokens = std::move(tokens_);
    tokens_.clear();
    b.salt_hash = salt_hash_;
    b.block_hash = bh[0];
    b.parent_sequence_hash = parent_sequence_hash_;
    if (parent_sequence_hash_) {
      uint64_t in[2] = {*parent_sequence_hash_, bh[0]};
      b.sequence_hash = compute_hash(in, 16, salt_hash_);
    } else {
      b.sequence_hash = bh[0];
    }
    parent_sequence_hash_ = b.sequence_hash;
    return b;
  }
  const std::vector<Token>& tokens() const { return tokens_; }
  size_t remaining_tokens() const { return block_size_ - tokens_.size(); }
  std::optional<uint64_t> parent_sequence_hash() const { return parent_sequence_hash_; }

 private:
  std::vector<Token> tokens_;
  size_t block_size_;
  uint64_t salt_hash_;
  std::optional<uint64_t> parent_sequence_hash_;
};

class TokenBlockSequence {  // lib/tokens/src/lib.rs:264-405
 public:
  TokenBlockSequence(const std::vector<Token>& tokens, size_t block_size, std::optional<uint64_t> salt_hash = std::nullopt)
      : salt_hash_(salt_hash.value_or(0)), current_({}, block_size, salt_hash_, std::nullopt) {
    std::vector<uint64_t> bh, sh;
    hash_blocks(tokens, block_size, salt_hash_, bh, sh);
    for (size_t b = 0; b < bh.size(); ++b) {
      TokenBlock tb;
      tb.tokens.assign(tokens.begin() + b * block_size, tokens.begin() + (b + 1) * block_size);
      tb.salt_hash = salt_hash_;
      tb.block_hash = bh[b];
      tb.sequence_hash = sh[b];
      if (b) tb.parent_sequence_hash = sh[b - 1];
      blocks_.push_back(std::move(tb));
    }
    std::vector<Token> rem(tokens.begin() + bh.size() * block_size, tokens.end());
    current_ = PartialTokenBlock(std::move(rem), block_size, salt_hash_,
                                 blocks_.empty() ? std::nullopt : std::optional<uint64_t>(blocks_.back().sequence_hash));
  }
  const TokenBlock* push_token(Token t) {
    auto b = current_.push_token(t);
    if (!b) return nullptr;
    blocks_.push_back(std::move(*b));
    return &blocks_.back();
  }
  const TokenBlock* last() const { return blocks_.empty() ? nullptr : &blocks_.back(); }
  const std::vector<TokenBlock>& blocks() const { return blocks_; }
  const PartialTokenBlock& current_block() const { return current_; }
  uint64_t salt_hash() const { return salt_hash_; }

 private:
  uint64_t salt_hash_;
  std::vector<TokenBlock> blocks_;
  PartialTokenBlock current_;
};

// kv_router/indexer.rs:123-135
inline std::vector<uint64_t> compute_block_hash_for_seq(const std::vector<Token>& tokens, size_t kv_block_size) {
  std::vector<uint64_t> bh, sh;
  hash_blocks(tokens, kv_block_size, XXH3_SEED, bh, sh);
  return bh;
}

}  // namespace kvdp
