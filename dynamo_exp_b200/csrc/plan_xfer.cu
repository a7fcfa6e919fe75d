// plan_xfer.cu — copy plans (device-resident pointer tables + strides) and the transfer engine that runs them.
//
// kvdp_plan  ~ CopyStreamBlockMap  (reference: lib/llm/src/kv/layer.rs:614-690): built once per block-set pair.
// kvdp_xfer  ~ CopyStream          (reference: lib/llm/src/kv/layer.rs:785-979, block_copy.cu:371-527): one stream
//              plus id staging; but one launch covers all layers of a request instead of 2*N*L DMAs.
#include "plan_xfer.cuh"

#include <string.h>

#include <algorithm>
#include <vector>

namespace kvdp {

// ------------------------------------------------------------------------------------------------ IdStage
int IdStage::init(int max_pairs, int dev) {
  device = dev;
  DeviceGuard g(device);
  capacity = std::max(max_pairs, 1);
  for (int s = 0; s < kSlots; ++s) {
    KVDP_CUDA(cudaMalloc(&d_slot[s], sizeof(int32_t) * 2 * (size_t)capacity));
    KVDP_CUDA(cudaHostAlloc(&h_slot[s], sizeof(int32_t) * 2 * (size_t)capacity, cudaHostAllocDefault));
    KVDP_CUDA(cudaEventCreateWithFlags(&ev[s], cudaEventDisableTiming));
    slot_cap[s] = capacity;
    used[s] = false;
  }
  KVDP_CUDA(cudaMalloc(&d_counter, sizeof(unsigned int)));
  KVDP_CUDA(cudaMemset(d_counter, 0, sizeof(unsigned int)));
  cur = -1;
  count = 0;
  ok = true;
  return 0;
}

void IdStage::destroy() {
  DeviceGuard g(device);
  for (int s = 0; s < kSlots; ++s) {
    if (used[s] && ev[s]) cudaEventSynchronize(ev[s]);
    if (d_slot[s]) cudaFree(d_slot[s]);
    if (h_slot[s]) cudaFreeHost(h_slot[s]);
    if (ev[s]) cudaEventDestroy(ev[s]);
    d_slot[s] = nullptr;
    h_slot[s] = nullptr;
    ev[s] = nullptr;
    used[s] = false;
  }
  if (d_counter) cudaFree(d_counter);
  d_counter = nullptr;
  ok = false;
  cur = -1;
}

int IdStage::upload(const int32_t* h_src, const int32_t* h_dst, int n, cudaStream_t stream) {
  if (!ok) return fail(KVDP_ESTATE, "id staging was not initialised (allocation failed at create time)");
  const int s = (cur + 1) % kSlots;
  if (used[s]) KVDP_CUDA(cudaEventSynchronize(ev[s]));  // the kernel that last read this slot has finished
  if (n > slot_cap[s]) {
    // rare: the caller exceeds the size it announced. This slot grows (its last reader is done, see above) and so does
    // every other slot that is idle right now, so that the allocation stalls are paid once and not again on each of the
    // next calls; slots that are still in flight are left alone and grow when their turn comes.
    DeviceGuard g(device);
    capacity = std::max(capacity, std::max(n, slot_cap[s] * 2));
    for (int t = 0; t < kSlots; ++t) {
      if (slot_cap[t] >= capacity) continue;
      if (t != s && used[t] && cudaEventQuery(ev[t]) != cudaSuccess) {
        cudaGetLastError();  // cudaErrorNotReady is not sticky, but keep the error state clean
        continue;
      }
      KVDP_CUDA(cudaFree(d_slot[t]));
      d_slot[t] = nullptr;
      KVDP_CUDA(cudaFreeHost(h_slot[t]));
      h_slot[t] = nullptr;
      slot_cap[t] = 0;
      KVDP_CUDA(cudaMalloc(&d_slot[t], sizeof(int32_t) * 2 * (size_t)capacity));
      KVDP_CUDA(cudaHostAlloc(&h_slot[t], sizeof(int32_t) * 2 * (size_t)capacity, cudaHostAllocDefault));
      slot_cap[t] = capacity;
    }
  }
  memcpy(h_slot[s], h_src, sizeof(int32_t) * (size_t)n);
  memcpy(h_slot[s] + n, h_dst, sizeof(int32_t) * (size_t)n);
  KVDP_CUDA(cudaMemcpyAsync(d_slot[s], h_slot[s], sizeof(int32_t) * 2 * (size_t)n, cudaMemcpyHostToDevice, stream));
  cur = s;
  count = n;
  return mark_in_use(stream);
}

int IdStage::mark_in_use(cudaStream_t stream) {
  if (cur < 0) return 0;
  KVDP_CUDA(cudaEventRecord(ev[cur], stream));
  used[cur] = true;
  return 0;
}

// ------------------------------------------------------------------------------------------------ plan
static int pointer_is_host(uint64_t p, bool* is_host, bool* is_pageable) {
  cudaPointerAttributes attr;
  cudaError_t e = cudaPointerGetAttributes(&attr, (const void*)(uintptr_t)p);
  if (e != cudaSuccess) {
    cudaGetLastError();
    *is_host = false;
    *is_pageable = true;
    return 0;
  }
  *is_host = attr.type == cudaMemoryTypeHost;
  *is_pageable = attr.type == cudaMemoryTypeUnregistered;
  return 0;
}

int plan_build(kvdp_plan** out, const kvdp_plan_desc& d) {
  if (!out) return fail(KVDP_EINVAL, "plan: NULL out");
  *out = nullptr;
  if (d.num_layers <= 0 || d.row_groups <= 0 || d.prefix_dim <= 0 || d.rows <= 0 || d.run_bytes <= 0)
    return fail(KVDP_EINVAL, "plan: num_layers/row_groups/prefix_dim/rows/run_bytes must be positive");
  if (!d.src_layer_ptrs || !d.dst_layer_ptrs) return fail(KVDP_EINVAL, "plan: NULL layer pointer table");
  const int64_t chunk = (int64_t)d.rows * d.row_groups * d.run_bytes;
  if (chunk >= ((int64_t)1 << 32) || d.run_bytes >= ((int64_t)1 << 32))
    return fail(KVDP_ERANGE, "plan: chunk of %lld bytes exceeds the 4 GiB per-chunk limit", (long long)chunk);
  const int n_dst = d.row_groups * d.num_layers;
  for (int i = 0; i < d.num_layers; ++i)
    if (!d.src_layer_ptrs[i]) return fail(KVDP_EINVAL, "plan: NULL source pointer for layer %d", i);
  for (int i = 0; i < n_dst; ++i)
    if (!d.dst_layer_ptrs[i]) return fail(KVDP_EINVAL, "plan: NULL destination pointer (entry %d)", i);

  kvdp_plan* p = new kvdp_plan();
  p->desc = d;
  p->desc.src_layer_ptrs = nullptr;
  p->desc.dst_layer_ptrs = nullptr;
  p->device = d.device;
  if (p->device < 0) cudaGetDevice(&p->device);
  p->h_src.assign(d.src_layer_ptrs, d.src_layer_ptrs + d.num_layers);
  p->h_dst.assign(d.dst_layer_ptrs, d.dst_layer_ptrs + n_dst);

  bool host = false, pageable = false, h, pg;
  for (uint64_t ptr : p->h_src) {  // every slab is looked at: layers may come from different allocations
    pointer_is_host(ptr, &h, &pg);
    host |= h;
    pageable |= pg;
  }
  for (uint64_t ptr : p->h_dst) {
    pointer_is_host(ptr, &h, &pg);
    host |= h;
    pageable |= pg;
  }
  p->host_memory = host;
  if (pageable) {
    delete p;
    return fail(KVDP_EINVAL, "plan: pageable host memory cannot be reached by the device; use pinned/registered memory");
  }

  CopyArgs a;
  memset(&a, 0, sizeof a);
  a.num_layers_total = d.num_layers;
  a.prefix_dim = d.prefix_dim;
  a.rows = d.rows;
  a.row_groups = d.row_groups;
  a.run_bytes = (uint32_t)d.run_bytes;
  a.chunk_bytes = (uint32_t)chunk;
  a.src_prefix_stride = d.src_prefix_stride;
  a.src_block_stride = d.src_block_stride;
  a.dst_prefix_stride = d.dst_prefix_stride;
  a.dst_block_stride = d.dst_block_stride;
  a.dst_row_stride = d.dst_row_stride;
  a.dst_col_offset = d.dst_col_offset;
  a.src_row_stride = d.src_row_stride;
  a.src_col_offset = d.src_col_offset;
  if (d.src_row_stride < 0 || d.src_col_offset < 0 || (d.src_row_stride == 0 && d.src_col_offset != 0)) {
    delete p;
    return fail(KVDP_EINVAL, "plan: src_row_stride/src_col_offset must be >= 0 and used together");
  }
  p->align = copy_alignment(a, p->h_src.data(), d.num_layers, p->h_dst.data(), n_dst);

  DeviceGuard g(p->device);
  cudaError_t e = cudaMalloc(&p->d_tab, sizeof(uint64_t) * (size_t)(d.num_layers + n_dst));
  if (e == cudaSuccess) {
    std::vector<uint64_t> tab(p->h_src);
    tab.insert(tab.end(), p->h_dst.begin(), p->h_dst.end());
    e = cudaMemcpy(p->d_tab, tab.data(), sizeof(uint64_t) * tab.size(), cudaMemcpyHostToDevice);
  }
  if (e != cudaSuccess) {
    if (p->d_tab) cudaFree(p->d_tab);
    delete p;
    return fail((int)e, "plan: device table upload failed: %s", cudaGetErrorString(e));
  }
  a.src_tab = p->d_tab;
  a.dst_tab = p->d_tab + d.num_layers;
  p->args = a;
  *out = p;
  return 0;
}

int plan_launch(const kvdp_plan* p, const int32_t* d_src_ids, const int32_t* d_dst_ids, int num_pairs,
                int layer_begin, int layer_end, cudaStream_t stream, unsigned int* d_counter, uint32_t* done_flag,
                uint32_t done_value) {
  if (!p) return fail(KVDP_EINVAL, "launch: NULL plan");
  if (num_pairs < 0) return fail(KVDP_EINVAL, "launch: negative num_pairs");
  if (layer_begin < 0 || layer_end > p->desc.num_layers || layer_begin > layer_end)
    return fail(KVDP_EINVAL, "launch: layer range [%d,%d) outside [0,%d)", layer_begin, layer_end,
                p->desc.num_layers);
  if (num_pairs > 0 && (!d_src_ids || !d_dst_ids)) return fail(KVDP_EINVAL, "launch: NULL id array");
  if (done_flag && !d_counter) return fail(KVDP_EINVAL, "launch: completion flag needs a counter");
  CopyArgs a = p->args;
  a.src_ids = d_src_ids;
  a.dst_ids = d_dst_ids;
  a.num_pairs = num_pairs;
  a.layer_begin = layer_begin;
  a.layers_run = layer_end - layer_begin;
  CopyLaunchHints hints;
  hints.host_memory = p->host_memory;
  hints.device = p->device;
  if (num_pairs == 0 || a.layers_run == 0) {
    // nothing to move; an attached completion flag is still published in stream order
    if (done_flag) return enqueue_signal(done_flag, done_value, stream);
    return 0;
  }
  a.done_counter = d_counter;
  a.done_flag = done_flag;
  a.done_value = done_value;
  return launch_strided_copy(a, p->align, hints, stream);
}

static int check_ids(const kvdp_plan* p, const int32_t* s, const int32_t* d, int n, uint32_t flags) {
  const int64_t sb = p->desc.src_blocks, db = p->desc.dst_blocks;
  for (int i = 0; i < n; ++i) {
    if (s[i] < 0 || (sb > 0 && s[i] >= sb))
      return fail(KVDP_EINVAL, "source block id %d (pair %d) outside [0,%lld)", s[i], i, (long long)sb);
    if (d[i] < 0 || (db > 0 && d[i] >= db))
      return fail(KVDP_EINVAL, "destination block id %d (pair %d) outside [0,%lld)", d[i], i, (long long)db);
  }
  if (flags & KVDP_XFER_CHECK_DST_UNIQUE) {
    if (db > 0 && (db >> 6) <= 8 * (int64_t)n + 1024) {  // clearing the bitmap must not cost more than the pass
      // ids are bounded (checked above): one pass over a bitmap, which matters for token-granular plans where n is
      // tens of thousands per call
      static thread_local std::vector<uint64_t> seen;
      seen.assign((size_t)((db + 63) / 64), 0);
      for (int i = 0; i < n; ++i) {
        uint64_t& w = seen[(size_t)d[i] >> 6];
        const uint64_t bit = 1ULL << (d[i] & 63);
        if (w & bit) return fail(KVDP_EDUP, "duplicate destination block id: two pairs would write the same block");
        w |= bit;
      }
    } else {
      std::vector<int32_t> v(d, d + n);
      std::sort(v.begin(), v.end());
      if (std::adjacent_find(v.begin(), v.end()) != v.end())
        return fail(KVDP_EDUP, "duplicate destination block id: two pairs would write the same block");
    }
  }
  return 0;
}

}  // namespace kvdp

using namespace kvdp;

extern "C" {

int kvdp_plan_create(kvdp_plan** out, const kvdp_plan_desc* desc) {
  if (!desc) return fail(KVDP_EINVAL, "kvdp_plan_create: NULL desc");
  return plan_build(out, *desc);
}

int kvdp_plan_destroy(kvdp_plan* plan) {
  if (!plan) return 0;
  DeviceGuard g(plan->device);
  if (plan->d_tab) cudaFree(plan->d_tab);
  delete plan;
  return 0;
}

int64_t kvdp_plan_bytes_per_pair(const kvdp_plan* plan) {
  if (!plan) return 0;
  return (int64_t)plan->args.chunk_bytes * plan->desc.prefix_dim * plan->desc.num_layers;
}

int kvdp_plan_block_copy(kvdp_plan** out, const uint64_t* src_layer_ptrs, const uint64_t* dst_layer_ptrs,
                         int32_t num_layers, int32_t prefix_dim, int64_t src_block_dim, int64_t dst_block_dim,
                         int64_t suffix_dim, int32_t elem_size, int32_t device) {
  // validation mirrors CopyStreamBlockMap's validator (kv/layer.rs:592-612,692-709): dims >= 1, elem_size in [1,8]
  if (prefix_dim < 1 || src_block_dim < 1 || dst_block_dim < 1 || suffix_dim < 1)
    return fail(KVDP_EINVAL, "block copy plan: prefix/src_block/dst_block/suffix dims must be >= 1");
  if (elem_size < 1 || elem_size > 8) return fail(KVDP_EINVAL, "block copy plan: elem_size must be in [1,8]");
  kvdp_plan_desc d;
  memset(&d, 0, sizeof d);
  const int64_t suffix_bytes = suffix_dim * elem_size;
  d.num_layers = num_layers;
  d.row_groups = 1;
  d.prefix_dim = prefix_dim;
  d.rows = 1;
  d.run_bytes = suffix_bytes;
  d.src_prefix_stride = src_block_dim * suffix_bytes;
  d.src_block_stride = suffix_bytes;
  d.dst_prefix_stride = dst_block_dim * suffix_bytes;
  d.dst_block_stride = suffix_bytes;
  d.dst_row_stride = suffix_bytes;
  d.src_blocks = src_block_dim;
  d.dst_blocks = dst_block_dim;
  d.src_layer_ptrs = src_layer_ptrs;
  d.dst_layer_ptrs = dst_layer_ptrs;
  d.device = device;
  return plan_build(out, d);
}

int kvdp_plan_contig(kvdp_plan** out, const uint64_t* layer_ptrs, int32_t num_layers, int32_t layout,
                     int64_t layer_blocks, int64_t chunk_bytes, void* contig_base, int64_t contig_blocks,
                     int32_t to_contig, int32_t device) {
  if (!layer_ptrs || !contig_base) return fail(KVDP_EINVAL, "contig plan: NULL pointer");
  if (num_layers < 1 || layer_blocks < 1 || chunk_bytes < 1 || contig_blocks < 1)
    return fail(KVDP_EINVAL, "contig plan: num_layers/layer_blocks/chunk_bytes/contig_blocks must be >= 1");
  if (layout != 0 && layout != 1) return fail(KVDP_EINVAL, "contig plan: layout must be 0 (KvFirst) or 1 (BlockFirst)");
  std::vector<uint64_t> cptrs((size_t)num_layers);
  for (int l = 0; l < num_layers; ++l) cptrs[l] = (uint64_t)(uintptr_t)contig_base + (uint64_t)l * 2 * chunk_bytes;
  const int64_t ls_prefix = layout == 0 ? layer_blocks * chunk_bytes : chunk_bytes;
  const int64_t ls_block = layout == 0 ? chunk_bytes : 2 * chunk_bytes;
  const int64_t c_prefix = chunk_bytes;
  const int64_t c_block = (int64_t)num_layers * 2 * chunk_bytes;
  kvdp_plan_desc d;
  memset(&d, 0, sizeof d);
  d.num_layers = num_layers;
  d.row_groups = 1;
  d.prefix_dim = 2;
  d.rows = 1;
  d.run_bytes = chunk_bytes;
  d.dst_row_stride = chunk_bytes;
  d.device = device;
  if (to_contig) {
    d.src_prefix_stride = ls_prefix;
    d.src_block_stride = ls_block;
    d.dst_prefix_stride = c_prefix;
    d.dst_block_stride = c_block;
    d.src_blocks = layer_blocks;
    d.dst_blocks = contig_blocks;
    d.src_layer_ptrs = layer_ptrs;
    d.dst_layer_ptrs = cptrs.data();
  } else {
    d.src_prefix_stride = c_prefix;
    d.src_block_stride = c_block;
    d.dst_prefix_stride = ls_prefix;
    d.dst_block_stride = ls_block;
    d.src_blocks = contig_blocks;
    d.dst_blocks = layer_blocks;
    d.src_layer_ptrs = cptrs.data();
    d.dst_layer_ptrs = layer_ptrs;
  }
  return plan_build(out, d);
}

int kvdp_plan_reshard(kvdp_plan** out, const uint64_t* src_layer_ptrs, const uint64_t* dst_layer_ptrs,
                      int32_t num_layers, int64_t src_blocks, int64_t dst_blocks, int32_t block_size,
                      int64_t src_row_bytes, int32_t src_tp, int32_t dst_tp, int32_t src_rank, int32_t device) {
  if (src_tp < 1 || dst_tp < 1 || src_rank < 0 || src_rank >= src_tp)
    return fail(KVDP_EINVAL, "reshard plan: need src_tp, dst_tp >= 1 and 0 <= src_rank < src_tp");
  if (block_size < 1 || src_row_bytes < 1 || src_blocks < 1 || dst_blocks < 1 || num_layers < 1)
    return fail(KVDP_EINVAL, "reshard plan: block_size/src_row_bytes/src_blocks/dst_blocks/num_layers must be >= 1");
  kvdp_plan_desc d;
  memset(&d, 0, sizeof d);
  d.num_layers = num_layers;
  d.prefix_dim = 2;
  d.rows = block_size;
  d.src_blocks = src_blocks;
  d.dst_blocks = dst_blocks;
  d.src_layer_ptrs = src_layer_ptrs;
  d.dst_layer_ptrs = dst_layer_ptrs;
  d.device = device;
  const int64_t src_chunk = (int64_t)block_size * src_row_bytes;
  d.src_prefix_stride = src_blocks * src_chunk;
  d.src_block_stride = src_chunk;
  if (dst_tp >= src_tp) {
    // scatter (the direction the reference supports: patch:1437; kv/layer.rs:1053-1056)
    if (dst_tp % src_tp) return fail(KVDP_EINVAL, "reshard plan: dst_tp must be a multiple of src_tp");
    const int m = dst_tp / src_tp;
    if (src_row_bytes % m) return fail(KVDP_EINVAL, "reshard plan: heads per source rank not divisible by %d", m);
    d.row_groups = m;
    d.run_bytes = src_row_bytes / m;
    const int64_t dst_chunk = (int64_t)block_size * d.run_bytes;
    d.dst_prefix_stride = dst_blocks * dst_chunk;
    d.dst_block_stride = dst_chunk;
    d.dst_row_stride = d.run_bytes;
    d.dst_col_offset = 0;
  } else {
    // gather (new: the reference asserts against it)
    if (src_tp % dst_tp) return fail(KVDP_EINVAL, "reshard plan: src_tp must be a multiple of dst_tp");
    const int m = src_tp / dst_tp;
    d.row_groups = 1;
    d.run_bytes = src_row_bytes;
    const int64_t dst_row = src_row_bytes * m;
    const int64_t dst_chunk = (int64_t)block_size * dst_row;
    d.dst_prefix_stride = dst_blocks * dst_chunk;
    d.dst_block_stride = dst_chunk;
    d.dst_row_stride = dst_row;
    d.dst_col_offset = (int64_t)(src_rank % m) * src_row_bytes;
  }
  return plan_build(out, d);
}

int kvdp_xfer_create(kvdp_xfer** out, int32_t max_pairs, int32_t device) {
  if (!out) return fail(KVDP_EINVAL, "kvdp_xfer_create: NULL out");
  *out = nullptr;
  if (max_pairs < 0) return fail(KVDP_EINVAL, "kvdp_xfer_create: negative max_pairs");
  kvdp_xfer* x = new kvdp_xfer();
  x->device = device;
  if (x->device < 0) cudaGetDevice(&x->device);
  DeviceGuard g(x->device);
  cudaError_t e = cudaStreamCreateWithFlags(&x->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) {
    delete x;
    return fail((int)e, "kvdp_xfer_create: stream: %s", cudaGetErrorString(e));
  }
  if (int rc = x->ids.init(max_pairs, x->device)) {
    x->ids.destroy();
    cudaStreamDestroy(x->stream);
    delete x;
    return rc;
  }
  *out = x;
  return 0;
}

int kvdp_xfer_destroy(kvdp_xfer* x) {
  if (!x) return 0;
  DeviceGuard g(x->device);
  cudaStreamSynchronize(x->stream);
  x->ids.destroy();
  cudaStreamDestroy(x->stream);
  delete x;
  return 0;
}

kvdp_stream_t kvdp_xfer_stream(kvdp_xfer* x) { return x ? (kvdp_stream_t)x->stream : nullptr; }

int kvdp_xfer_run(kvdp_xfer* x, const kvdp_plan* plan, const int32_t* h_src_ids, const int32_t* h_dst_ids,
                  int32_t num_pairs, int32_t layer_begin, int32_t layer_end, uint32_t flags, uint32_t* done_flag,
                  uint32_t done_value) {
  if (!x || !plan) return fail(KVDP_EINVAL, "kvdp_xfer_run: NULL handle");
  if (num_pairs < 0) return fail(KVDP_EINVAL, "kvdp_xfer_run: negative num_pairs");
  if (num_pairs > 0 && (!h_src_ids || !h_dst_ids)) return fail(KVDP_EINVAL, "kvdp_xfer_run: NULL id array");
  if (plan->device != x->device)
    return fail(KVDP_EINVAL, "kvdp_xfer_run: the plan was built for device %d, the engine runs on device %d",
                plan->device, x->device);
  if (int rc = check_ids(plan, h_src_ids, h_dst_ids, num_pairs, flags)) return rc;
  DeviceGuard g(x->device);
  if (num_pairs > 0)
    if (int rc = x->ids.upload(h_src_ids, h_dst_ids, num_pairs, x->stream)) return rc;
  if (int rc = plan_launch(plan, x->ids.d_src(), x->ids.d_dst(), num_pairs, layer_begin, layer_end, x->stream,
                           x->ids.d_counter, done_flag, done_value))
    return rc;
  return num_pairs > 0 ? x->ids.mark_in_use(x->stream) : 0;
}

int kvdp_xfer_signal(kvdp_xfer* x, uint32_t* flag, uint32_t value) {
  if (!x || !flag) return fail(KVDP_EINVAL, "kvdp_xfer_signal: NULL argument");
  DeviceGuard g(x->device);
  return enqueue_signal(flag, value, x->stream);
}

int kvdp_xfer_sync(kvdp_xfer* x) {
  if (!x) return fail(KVDP_EINVAL, "kvdp_xfer_sync: NULL handle");
  DeviceGuard g(x->device);
  KVDP_CUDA(cudaStreamSynchronize(x->stream));
  return 0;
}

int kvdp_plan_launch(const kvdp_plan* plan, const int32_t* d_src_ids, const int32_t* d_dst_ids, int32_t num_pairs,
                     int32_t layer_begin, int32_t layer_end, kvdp_stream_t stream) {
  if (!plan) return fail(KVDP_EINVAL, "kvdp_plan_launch: NULL plan");
  DeviceGuard g(plan->device);
  return plan_launch(plan, d_src_ids, d_dst_ids, num_pairs, layer_begin, layer_end, (cudaStream_t)stream, nullptr,
                     nullptr, 0);
}

}  // extern "C"
