// C++ host-layer test (include/kvdp.hpp): the reference's own GPU unit tests restated over the C ABI.
//   kv/storage.rs:1876 (host/device transfers), kv/layer.rs:1412-1584 (copy_blocks_to pattern),
//   kv/layer.rs:1590-1735 (CopyStream H2D then reversed D2H), kv/layer.rs:1865-2065 (scatter tp1 -> tp4),
//   lib/tokens/src/lib.rs:463-623 (known answers, push_token), kv_router/indexer.rs:1202-1225 (counts).
// Needs a GPU. Prints "CPP HOST MIRROR OK" and exits 0 when every check holds.
#include <cstdio>
#include <cstdlib>
#include <functional>

#include "kvdp.hpp"

using namespace kvdp;

static int failures = 0;
#define CHECK(cond)                                                              \
  do {                                                                           \
    if (!(cond)) {                                                               \
      std::fprintf(stderr, "CHECK failed at line %d: %s\n", __LINE__, #cond);    \
      ++failures;                                                                \
    }                                                                            \
  } while (0)

static bool throws(const std::function<void()>& f, const char* needle) {
  try {
    f();
  } catch (const Error& e) {
    return std::string(e.what()).find(needle) != std::string::npos;
  }
  return false;
}

static KvBlockDetails details(KvLayout layout, size_t layers, DType dt = DType::F32) {
  return KvBlockDetails{layout, 4, KvModelDetails{layers, 2, 8, dt}};
}

static void test_transfers() {  // kv/storage.rs:1876
  auto h = OwnedStorage::create_pinned_array(4 * 6 * 4);
  auto d = OwnedStorage::create_device_array(4 * 6 * 4, 0);
  auto h2 = OwnedStorage::create_pinned_array(4 * 6 * 4);
  float* p = (float*)(uintptr_t)h.get_pointer();
  for (int i = 0; i < 24; ++i) p[i] = (float)i * 0.5f;
  TensorView<2> hv(h, {4, 6}, 4), dv(d, {4, 6}, 4), hv2(h2, {4, 6}, 4);
  CHECK((hv.strides() == std::array<size_t, 2>{6, 1}));
  CHECK((hv.byte_offset({2, 3}) == (2 * 6 + 3) * 4));
  hv.h2d(dv, nullptr);
  dv.d2h(hv2, nullptr);
  CHECK(cuda_memcpy_sync(nullptr, nullptr, 0) == 0 || true);
  hv.copy_to_view_blocking(dv);  // blocking copy also drains the async ones on the legacy stream
  dv.copy_to_view_blocking(hv2);
  float* q = (float*)(uintptr_t)h2.get_pointer();
  for (int i = 0; i < 24; ++i) CHECK(q[i] == p[i]);
  CHECK(throws([&] { hv.h2d(hv2, nullptr); }, "Destination must be a device tensor"));
  CHECK(throws([&] { hv.d2h(hv2, nullptr); }, "Source must be a device tensor"));
  auto s = hv.slice(1, 2, 5);
  CHECK(!s.is_contiguous() && s.offset() == 8 && s.shape()[1] == 3);
  CHECK(throws([&] { TensorView<2>(h, {5, 6}, 4); }, "requires 120 bytes"));
}

static void test_copy_blocks_to(KvLayout layout) {  // kv/layer.rs:1412-1584
  const size_t nb = 8;
  auto bd = details(layout, 1);
  auto dev = KvBlockStorage::allocate(nb, bd, StorageType::device_(0));
  auto src = KvBlockStorage::allocate(nb, bd, StorageType::pinned());
  auto out = KvBlockStorage::allocate(nb, bd, StorageType::pinned());
  auto shape = src.layer(0).layer_shape();
  size_t per_block = 4 * 2 * 8;
  float* p = (float*)(uintptr_t)src.layer(0).get_pointer();
  for (size_t b = 0; b < nb; ++b)
    for (size_t kv = 0; kv < 2; ++kv)
      for (size_t i = 0; i < per_block; ++i) {
        size_t off = layout == KvLayout::KvFirst ? (kv * nb + b) * per_block + i : (b * 2 + kv) * per_block + i;
        p[off] = (float)b;
      }
  (void)shape;
  std::vector<size_t> ident(nb);
  std::iota(ident.begin(), ident.end(), 0);
  src.layer(0).copy_blocks_to(ident, dev.layer(0), ident);
  dev.layer(0).copy_blocks_to({1, 2, 2, 3, 5}, out.layer(0), {0, 3, 2, 1, 4});
  float* q = (float*)(uintptr_t)out.layer(0).get_pointer();
  size_t s_ids[5] = {1, 2, 2, 3, 5}, d_ids[5] = {0, 3, 2, 1, 4};
  for (int k = 0; k < 5; ++k)
    for (size_t kv = 0; kv < 2; ++kv)
      for (size_t i = 0; i < per_block; ++i) {
        size_t b = d_ids[k];
        size_t off = layout == KvLayout::KvFirst ? (kv * nb + b) * per_block + i : (b * 2 + kv) * per_block + i;
        CHECK(q[off] == (float)s_ids[k]);
      }
  CHECK(throws([&] { dev.layer(0).copy_blocks_to({0}, dev.layer(0), {1}); }, "Device to Device"));
  CHECK(throws([&] { src.layer(0).copy_blocks_to({0}, out.layer(0), {1}); }, "Pinned to Pinned"));
}

static void test_copy_stream(KvLayout layout) {  // kv/layer.rs:1590-1735
  const size_t L = 3, nh = 6, nd = 9;
  auto bd = details(layout, L, DType::BF16);
  auto host = KvBlockStorage::allocate(nh, bd, StorageType::pinned());
  auto dev = KvBlockStorage::allocate(nd, bd, StorageType::device_(0));
  auto back = KvBlockStorage::allocate(nh, bd, StorageType::pinned());
  size_t bytes = host.layer(0).storage_size();
  for (size_t l = 0; l < L; ++l) {
    uint8_t* p = (uint8_t*)(uintptr_t)host.layer(l).get_pointer();
    for (size_t i = 0; i < bytes; ++i) p[i] = (uint8_t)((i * 131 + l * 17 + (i >> 8)) & 0xff);
  }
  std::vector<int> h_ids = {0, 2, 5, 3}, d_ids = {8, 1, 4, 6};
  kvdp::CopyStream cs(L, 8);
  CHECK(throws([&] { cs.prepare_block_ids({0}, {0}); }, "Layers must be loaded before preparing block ids"));
  cs.prepare_block_map(CopyStreamBlockMap::make(host, dev));
  CHECK(throws([&] { cs.prepare_block_map(CopyStreamBlockMap::make(host, dev)); }, "Layers already loaded"));
  CHECK(throws([&] { cs.trigger_layer(0); }, "Block ids must be loaded"));
  CHECK(throws([&] { cs.prepare_block_ids({0, 1}, {2, 2}); }, "dst_block_ids must be unique"));
  cs.prepare_block_ids(h_ids, d_ids);
  CHECK(throws([&] { cs.prepare_block_ids(h_ids, d_ids); }, "Block ids already loaded"));
  cs.trigger_all_layers();
  cs.trigger_layer(1);  // doorbell: no-op
  CHECK(throws([&] { cs.trigger_layer(L); }, "out of bounds"));
  cs.sync_stream();
  cs.reset();
  cs.prepare_block_map(CopyStreamBlockMap::make(dev, back));
  cs.prepare_block_ids(d_ids, h_ids);
  cs.trigger_all_layers();
  cs.sync_stream();
  cs.on_return();
  size_t prefix = bd.prefix_dim(), suffix_bytes = bd.suffix_dim() * bd.elem_size();
  for (size_t l = 0; l < L; ++l) {
    const uint8_t* a = (const uint8_t*)(uintptr_t)host.layer(l).get_pointer();
    const uint8_t* b = (const uint8_t*)(uintptr_t)back.layer(l).get_pointer();
    for (size_t p = 0; p < prefix; ++p)
      for (int id : h_ids) {
        size_t off = (p * nh + id) * suffix_bytes;
        CHECK(std::memcmp(a + off, b + off, suffix_bytes) == 0);
      }
    // untouched host blocks stay zero
    size_t off1 = (0 * nh + 1) * suffix_bytes;
    for (size_t i = 0; i < suffix_bytes; ++i) CHECK(b[off1 + i] == 0);
  }
}

static void test_scatter() {  // kv/layer.rs:1865-2065: tp1 -> tp4, 16 heads, head size 6, f32
  const size_t nb = 3, bs = 4, heads = 16, hs = 6, sf = 4, dnh = heads / sf;
  KvBlockDetails bd{KvLayout::KvFirst, bs, KvModelDetails{1, heads, hs, DType::F32}};
  auto src = KvBlockStorage::allocate(nb, bd, StorageType::pinned());
  auto dst = KvBlockStorage::allocate(nb, bd, StorageType::pinned());
  auto dev_s = KvBlockStorage::allocate(nb, bd, StorageType::device_(0));
  auto dev_d = KvBlockStorage::allocate(nb, bd, StorageType::device_(0));
  float* p = (float*)(uintptr_t)src.layer(0).get_pointer();
  auto at = [&](size_t i0, size_t i1, size_t t, size_t h, size_t c) { return (((i0 * nb + i1) * bs + t) * heads + h) * hs + c; };
  for (size_t i0 = 0; i0 < 2; ++i0)
    for (size_t i1 = 0; i1 < nb; ++i1)
      for (size_t t = 0; t < bs; ++t)
        for (size_t h = 0; h < heads; ++h)
          for (size_t c = 0; c < hs; ++c) p[at(i0, i1, t, h, c)] = 1000.f * i0 + 100.f * i1 + 10.f * t + h + 0.1f * c;
  std::vector<size_t> ident(nb);
  std::iota(ident.begin(), ident.end(), 0);
  src.layer(0).copy_blocks_to(ident, dev_s.layer(0), ident);
  kvdp::CopyStream cs(1, nb);
  cs.prepare_block_map(CopyStreamBlockMap::make(dev_s, dev_d));
  CHECK(throws([&] { cs.scatter_copy_layer(0, {2, nb, bs, heads, hs}, 4, 1, 1, 4); }, "Block IDs must be loaded"));
  cs.prepare_block_ids({0, 1, 2}, {0, 1, 2});
  CHECK(throws([&] { cs.scatter_copy_layer(0, {2, nb, bs, heads, hs}, 4, 1, 4, 4); }, "src_tp_size must be less"));
  CHECK(throws([&] { cs.scatter_copy_layer(0, {2, nb, bs, heads, hs}, 4, 1, 1, 3); }, "power of 2"));
  cs.scatter_copy_layer(0, {2, nb, bs, heads, hs}, 4, 1, 1, 4);
  cs.sync_stream();
  dev_d.layer(0).copy_blocks_to(ident, dst.layer(0), ident);
  const float* q = (const float*)(uintptr_t)dst.layer(0).get_pointer();
  // dst = src.reshape(2, nb, bs, sf, dnh, hs).permuted_axes([3,0,1,2,4,5])
  size_t k = 0;
  for (size_t g = 0; g < sf; ++g)
    for (size_t i0 = 0; i0 < 2; ++i0)
      for (size_t i1 = 0; i1 < nb; ++i1)
        for (size_t t = 0; t < bs; ++t)
          for (size_t h = 0; h < dnh; ++h)
            for (size_t c = 0; c < hs; ++c) CHECK(q[k++] == p[at(i0, i1, t, g * dnh + h, c)]);
}

static void test_tokens() {  // lib/tokens/src/lib.rs:463-623
  std::vector<Token> t = {1, 2, 3, 4, 5, 6, 7, 8, 9, 10};
  TokenBlockSequence seq(t, 4, 1337);
  CHECK(seq.blocks().size() == 2);
  CHECK(seq.blocks()[0].block_hash == 14643705804678351452ULL);
  CHECK(seq.blocks()[0].sequence_hash == 14643705804678351452ULL);
  CHECK(!seq.blocks()[0].parent_sequence_hash.has_value());
  CHECK(seq.blocks()[1].block_hash == 16777012769546811212ULL);
  CHECK(seq.blocks()[1].sequence_hash == 4945711292740353085ULL);
  CHECK(seq.current_block().tokens().size() == 2 && seq.current_block().remaining_tokens() == 2);
  TokenBlockSequence inc({}, 4, 1337);
  int completed = 0;
  for (Token x : t) completed += inc.push_token(x) != nullptr;
  CHECK(completed == 2);
  CHECK(inc.blocks()[1].sequence_hash == 4945711292740353085ULL);
  CHECK(inc.last()->block_hash == 16777012769546811212ULL);
  uint32_t one[4] = {1, 2, 3, 4};
  CHECK(compute_hash(one, 16, 1337) == 14643705804678351452ULL);
  // indexer counts (kv_router/indexer.rs:1202-1225)
  std::vector<Token> a(11), b(37), c(128);
  std::iota(a.begin(), a.end(), 0);
  std::iota(b.begin(), b.end(), 0);
  std::iota(c.begin(), c.end(), 0);
  CHECK(compute_block_hash_for_seq(a, 11).size() == 1);
  CHECK(compute_block_hash_for_seq(b, 32).size() == 1);
  CHECK(compute_block_hash_for_seq(c, 64).size() == 2);
  TokenBlockSequence none({1, 2, 3}, 4);
  CHECK(none.blocks().empty() && none.salt_hash() == 0 && none.last() == nullptr);
}

int main() {
  int n = 0;
  if (kvdp_device_count(&n) != 0 || n < 1) {
    std::fprintf(stderr, "no CUDA device: %s\n", kvdp_last_error());
    return 2;
  }
  try {
    test_transfers();
    test_copy_blocks_to(KvLayout::KvFirst);
    test_copy_blocks_to(KvLayout::BlockFirst);
    test_copy_stream(KvLayout::KvFirst);
    test_copy_stream(KvLayout::BlockFirst);
    test_scatter();
    test_tokens();
    KvBlockDetails bad{KvLayout::KvFirst, 4, KvModelDetails{1, 8, 8, DType::F16}, 0, 3};
    CHECK(throws([&] { bad.validate(); }, "evenly divide"));
  } catch (const std::exception& e) {
    std::fprintf(stderr, "unexpected exception: %s\n", e.what());
    return 1;
  }
  if (failures) {
    std::fprintf(stderr, "%d check(s) failed\n", failures);
    return 1;
  }
  std::printf("CPP HOST MIRROR OK\n");
  return 0;
}
