// host_tier.cu — offload / onboard of KV blocks between HBM (layer-separate) and pinned host DRAM (contiguous-per-block).
//
// The reference's host tier is the same CopyStream path as everything else: per request it issues prefix*N*L
// cudaMemcpyAsync calls of one 128 KiB chunk each (lib/llm/src/kv/layer.rs:906-968 -> block_copy.cu:246-305; 4096 DMAs
// for a 4k-token Llama-3-8B request). Here the PCIe side sees one DMA per run of consecutive host block ids (>= 8 MiB
// each) and the scattered side is handled by one gather/scatter kernel per staging slot at HBM speed:
//
//   offload:  [compute stream] gather  layers[:, :, device_id] -> staging slot (contiguous-per-block)
//             [copy stream]    cudaMemcpyAsync staging slot -> pinned slab           (two slots, double-buffered)
//   onboard:  [copy stream]    cudaMemcpyAsync pinned slab -> staging slot
//             [compute stream] scatter staging slot -> layers[:, :, device_id]
#include <string.h>

#include <algorithm>
#include <vector>

#include "plan_xfer.cuh"

using namespace kvdp;

struct kvdp_host_tier {
  static constexpr int kSlots = 2;
  int device = -1;
  int num_layers = 0;
  int64_t block_bytes = 0;  // L * 2 * chunk
  int64_t host_blocks = 0;
  int64_t layer_blocks = 0;
  int staging_blocks = 0;
  char* pinned = nullptr;
  char* staging[kSlots] = {nullptr, nullptr};
  kvdp_plan* to_stage[kSlots] = {nullptr, nullptr};    // layers -> staging slot
  kvdp_plan* from_stage[kSlots] = {nullptr, nullptr};  // staging slot -> layers
  kvdp_xfer* xfer = nullptr;                           // compute stream + id staging
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t kernel_done[kSlots] = {nullptr, nullptr};
  cudaEvent_t dma_done[kSlots] = {nullptr, nullptr};
  bool slot_used[kSlots] = {false, false};
  std::vector<int32_t> iota;
};

namespace {

int dma_ranges(kvdp_host_tier* t, const int32_t* host_ids, int n, char* slot, bool to_host) {
  // one DMA per maximal run of consecutive host ids (slot positions are consecutive by construction), issued in pieces of
  // one size: an offload and an onboard running at once share the link evenly only when their DMAs are equally large
  // (see host_dma_layer in legacy_abi.cu)
  const size_t piece_opt = (size_t)options().host_dma_piece_kib << 10;
  const size_t piece = piece_opt ? piece_opt : ~(size_t)0;
  int i = 0;
  while (i < n) {
    int j = i + 1;
    while (j < n && host_ids[j] == host_ids[j - 1] + 1) ++j;
    char* h = t->pinned + (int64_t)host_ids[i] * t->block_bytes;
    char* s = slot + (int64_t)i * t->block_bytes;
    const size_t bytes = (size_t)(j - i) * (size_t)t->block_bytes;
    for (size_t o = 0; o < bytes; o += piece) {
      const size_t len = std::min(piece, bytes - o);
      if (to_host)
        KVDP_CUDA(cudaMemcpyAsync(h + o, s + o, len, cudaMemcpyDeviceToHost, t->copy_stream));
      else
        KVDP_CUDA(cudaMemcpyAsync(s + o, h + o, len, cudaMemcpyHostToDevice, t->copy_stream));
    }
    i = j;
  }
  return 0;
}

int check_tier_ids(kvdp_host_tier* t, const int32_t* dev_ids, const int32_t* host_ids, int n) {
  if (n < 0) return fail(KVDP_EINVAL, "host tier: negative block count");
  if (n > 0 && (!dev_ids || !host_ids)) return fail(KVDP_EINVAL, "host tier: NULL id array");
  for (int i = 0; i < n; ++i) {
    if (dev_ids[i] < 0 || dev_ids[i] >= t->layer_blocks)
      return fail(KVDP_EINVAL, "host tier: device block id %d outside [0,%lld)", dev_ids[i], (long long)t->layer_blocks);
    if (host_ids[i] < 0 || host_ids[i] >= t->host_blocks)
      return fail(KVDP_EINVAL, "host tier: host block id %d outside [0,%lld)", host_ids[i], (long long)t->host_blocks);
  }
  return 0;
}

}  // namespace

extern "C" {

int kvdp_host_tier_create(kvdp_host_tier** out, const uint64_t* layer_ptrs, int32_t num_layers, int32_t layout,
                          int64_t layer_blocks, int64_t chunk_bytes, void* pinned_slab, int64_t host_blocks,
                          int32_t staging_blocks, int32_t device) {
  if (!out) return fail(KVDP_EINVAL, "host tier: NULL out");
  *out = nullptr;
  if (!layer_ptrs || !pinned_slab) return fail(KVDP_EINVAL, "host tier: NULL pointer");
  if (num_layers < 1 || layer_blocks < 1 || chunk_bytes < 1 || host_blocks < 1 || staging_blocks < 1)
    return fail(KVDP_EINVAL, "host tier: sizes must be >= 1");
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, pinned_slab) != cudaSuccess || attr.type != cudaMemoryTypeHost) {
    cudaGetLastError();
    return fail(KVDP_EINVAL, "host tier: the host slab must be pinned (cuda_malloc_host) or registered memory");
  }
  kvdp_host_tier* t = new kvdp_host_tier();
  t->device = device;
  if (t->device < 0) cudaGetDevice(&t->device);
  t->num_layers = num_layers;
  t->block_bytes = (int64_t)num_layers * 2 * chunk_bytes;
  t->host_blocks = host_blocks;
  t->layer_blocks = layer_blocks;
  t->staging_blocks = staging_blocks;
  t->pinned = (char*)pinned_slab;
  t->iota.resize((size_t)staging_blocks);
  for (int i = 0; i < staging_blocks; ++i) t->iota[i] = i;
  DeviceGuard g(t->device);
  int rc = 0;
  cudaError_t e = cudaStreamCreateWithFlags(&t->copy_stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) rc = fail((int)e, "host tier: stream: %s", cudaGetErrorString(e));
  if (!rc) rc = kvdp_xfer_create(&t->xfer, staging_blocks, t->device);
  for (int s = 0; s < kvdp_host_tier::kSlots && !rc; ++s) {
    e = cudaMalloc(&t->staging[s], (size_t)staging_blocks * (size_t)t->block_bytes);
    if (e != cudaSuccess) {
      rc = fail((int)e, "host tier: staging slab of %lld bytes: %s", (long long)staging_blocks * t->block_bytes,
                cudaGetErrorString(e));
      break;
    }
    if (!rc) rc = kvdp_plan_contig(&t->to_stage[s], layer_ptrs, num_layers, layout, layer_blocks, chunk_bytes,
                                   t->staging[s], staging_blocks, 1, t->device);
    if (!rc) rc = kvdp_plan_contig(&t->from_stage[s], layer_ptrs, num_layers, layout, layer_blocks, chunk_bytes,
                                   t->staging[s], staging_blocks, 0, t->device);
    if (!rc && cudaEventCreateWithFlags(&t->kernel_done[s], cudaEventDisableTiming) != cudaSuccess)
      rc = fail(KVDP_EINVAL, "host tier: event");
    if (!rc && cudaEventCreateWithFlags(&t->dma_done[s], cudaEventDisableTiming) != cudaSuccess)
      rc = fail(KVDP_EINVAL, "host tier: event");
  }
  if (rc) {
    kvdp_host_tier_destroy(t);
    return rc;
  }
  *out = t;
  return 0;
}

int kvdp_host_tier_destroy(kvdp_host_tier* t) {
  if (!t) return 0;
  DeviceGuard g(t->device);
  if (t->xfer) kvdp_xfer_sync(t->xfer);
  if (t->copy_stream) cudaStreamSynchronize(t->copy_stream);
  for (int s = 0; s < kvdp_host_tier::kSlots; ++s) {
    kvdp_plan_destroy(t->to_stage[s]);
    kvdp_plan_destroy(t->from_stage[s]);
    if (t->staging[s]) cudaFree(t->staging[s]);
    if (t->kernel_done[s]) cudaEventDestroy(t->kernel_done[s]);
    if (t->dma_done[s]) cudaEventDestroy(t->dma_done[s]);
  }
  if (t->xfer) kvdp_xfer_destroy(t->xfer);
  if (t->copy_stream) cudaStreamDestroy(t->copy_stream);
  delete t;
  return 0;
}

int kvdp_host_tier_offload(kvdp_host_tier* t, const int32_t* h_device_ids, const int32_t* h_host_ids, int32_t n) {
  if (!t) return fail(KVDP_EINVAL, "host tier: NULL handle");
  if (int rc = check_tier_ids(t, h_device_ids, h_host_ids, n)) return rc;
  DeviceGuard g(t->device);
  cudaStream_t ks = (cudaStream_t)kvdp_xfer_stream(t->xfer);
  int batch = 0;
  for (int off = 0; off < n; off += t->staging_blocks, ++batch) {
    const int s = batch % kvdp_host_tier::kSlots;
    const int b = std::min(t->staging_blocks, n - off);
    if (t->slot_used[s]) KVDP_CUDA(cudaStreamWaitEvent(ks, t->dma_done[s], 0));  // slot drained by its last DMA
    if (int rc = kvdp_xfer_run(t->xfer, t->to_stage[s], h_device_ids + off, t->iota.data(), b, 0, t->num_layers, 0,
                               nullptr, 0))
      return rc;
    KVDP_CUDA(cudaEventRecord(t->kernel_done[s], ks));
    KVDP_CUDA(cudaStreamWaitEvent(t->copy_stream, t->kernel_done[s], 0));
    if (int rc = dma_ranges(t, h_host_ids + off, b, t->staging[s], true)) return rc;
    KVDP_CUDA(cudaEventRecord(t->dma_done[s], t->copy_stream));
    t->slot_used[s] = true;
  }
  return 0;
}

int kvdp_host_tier_onboard(kvdp_host_tier* t, const int32_t* h_host_ids, const int32_t* h_device_ids, int32_t n) {
  if (!t) return fail(KVDP_EINVAL, "host tier: NULL handle");
  if (int rc = check_tier_ids(t, h_device_ids, h_host_ids, n)) return rc;
  DeviceGuard g(t->device);
  cudaStream_t ks = (cudaStream_t)kvdp_xfer_stream(t->xfer);
  int batch = 0;
  for (int off = 0; off < n; off += t->staging_blocks, ++batch) {
    const int s = batch % kvdp_host_tier::kSlots;
    const int b = std::min(t->staging_blocks, n - off);
    // the slot may still be read by the scatter kernel (or written by the DMA) of its previous use
    if (t->slot_used[s]) {
      KVDP_CUDA(cudaStreamWaitEvent(t->copy_stream, t->kernel_done[s], 0));
      KVDP_CUDA(cudaStreamWaitEvent(t->copy_stream, t->dma_done[s], 0));
    }
    if (int rc = dma_ranges(t, h_host_ids + off, b, t->staging[s], false)) return rc;
    KVDP_CUDA(cudaEventRecord(t->dma_done[s], t->copy_stream));
    KVDP_CUDA(cudaStreamWaitEvent(ks, t->dma_done[s], 0));
    if (int rc = kvdp_xfer_run(t->xfer, t->from_stage[s], t->iota.data(), h_device_ids + off, b, 0, t->num_layers,
                               KVDP_XFER_CHECK_DST_UNIQUE, nullptr, 0))
      return rc;
    KVDP_CUDA(cudaEventRecord(t->kernel_done[s], ks));
    t->slot_used[s] = true;
  }
  return 0;
}

int kvdp_host_tier_sync(kvdp_host_tier* t) {
  if (!t) return fail(KVDP_EINVAL, "host tier: NULL handle");
  DeviceGuard g(t->device);
  KVDP_CUDA(cudaStreamSynchronize((cudaStream_t)kvdp_xfer_stream(t->xfer)));
  KVDP_CUDA(cudaStreamSynchronize(t->copy_stream));
  return 0;
}

}  // extern "C"
