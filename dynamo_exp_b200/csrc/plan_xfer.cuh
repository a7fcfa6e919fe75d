// plan_xfer.cuh — internal structs behind the opaque kvdp_plan / kvdp_xfer handles.
#pragma once
#include <vector>

#include "strided_copy.cuh"

namespace kvdp {

// Block-id staging: a ring of (pinned host, device) id buffers. The caller's arrays are copied into the pinned slot
// before the call returns; the slot is recycled only after the kernel that consumed it has finished (event).
struct IdStage {
  static constexpr int kSlots = 8;
  int32_t* h_slot[kSlots] = {};
  int32_t* d_slot[kSlots] = {};
  cudaEvent_t ev[kSlots] = {};
  int slot_cap[kSlots] = {};  // pairs each slot can hold (slots grow one at a time, when their last reader is done)
  bool used[kSlots] = {};
  unsigned int* d_counter = nullptr;  // CTA arrival counter for the in-kernel completion flag
  int capacity = 0;
  int count = 0;    // ids in the current slot
  int cur = -1;     // slot of the last upload
  int device = -1;
  bool ok = false;

  int init(int max_pairs, int dev);
  void destroy();
  // copies the ids into the next slot and enqueues the upload on `stream`
  int upload(const int32_t* h_src, const int32_t* h_dst, int n, cudaStream_t stream);
  // call after enqueuing a consumer of the current slot on `stream`
  int mark_in_use(cudaStream_t stream);
  const int32_t* d_src() const { return cur >= 0 ? d_slot[cur] : nullptr; }
  const int32_t* d_dst() const { return cur >= 0 ? d_slot[cur] + count : nullptr; }
};

int enqueue_signal(uint32_t* flag, uint32_t value, cudaStream_t stream);

}  // namespace kvdp

struct kvdp_plan {
  kvdp_plan_desc desc;
  kvdp::CopyArgs args;            // everything except ids / layer range / completion
  std::vector<uint64_t> h_src;    // host copies of the pointer tables
  std::vector<uint64_t> h_dst;
  uint64_t* d_tab = nullptr;      // device: [num_layers] src then [row_groups*num_layers] dst
  uint32_t align = 1;
  bool host_memory = false;
  int device = -1;
};

struct kvdp_xfer {
  cudaStream_t stream = nullptr;
  kvdp::IdStage ids;
  int device = -1;
};

namespace kvdp {
int plan_build(kvdp_plan** out, const kvdp_plan_desc& d);
int plan_launch(const kvdp_plan* p, const int32_t* d_src_ids, const int32_t* d_dst_ids, int num_pairs,
                int layer_begin, int layer_end, cudaStream_t stream, unsigned int* d_counter, uint32_t* done_flag,
                uint32_t done_value);
}  // namespace kvdp
