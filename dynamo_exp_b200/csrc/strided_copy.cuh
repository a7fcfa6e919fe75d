// strided_copy.cuh — launch interface of the one copy kernel family every byte-moving path maps onto.
#pragma once
#include "common.cuh"

namespace kvdp {

// Device-visible description of one launch. See kvdp.h (kvdp_plan_desc) for the address algebra.
struct CopyArgs {
  const uint64_t* src_tab;  // device [num_layers_total]            (nullptr: use src_single)
  const uint64_t* dst_tab;  // device [row_groups][num_layers_total] (nullptr: dst_single + g*dst_group_stride)
  uint64_t src_single;
  uint64_t dst_single;
  int64_t dst_group_stride;
  const int32_t* src_ids;  // device; nullptr = identity (legacy scatter ignores the staged ids)
  const int32_t* dst_ids;
  int32_t num_pairs;
  int32_t layer_begin;
  int32_t layers_run;
  int32_t num_layers_total;
  int32_t prefix_dim;
  int32_t rows;
  int32_t row_groups;
  uint32_t run_bytes;
  uint32_t chunk_bytes;  // rows * row_groups * run_bytes  (< 4 GiB on the kernel path)
  int64_t src_prefix_stride, src_block_stride;
  int64_t dst_prefix_stride, dst_block_stride, dst_row_stride, dst_col_offset;
  // strided source rows (only for pulling a column range out of wider remote rows): when src_row_stride != 0 the
  // chunk's row r starts at chunk_base + r*src_row_stride + src_col_offset instead of r*row_groups*run_bytes
  int64_t src_row_stride, src_col_offset;
  uint32_t tile_bytes;
  uint32_t l2_hint;  // path 3: bit 0 = evict_first on loads, bit 1 = on stores
  uint32_t group_lanes;  // path 4: lanes that share one chunk (4, 8, 16 or 32)
  uint32_t tiles_per_chunk;
  int64_t total_tiles;
  // optional completion signal (last CTA to finish publishes *done_flag = done_value at system scope)
  unsigned int* done_counter;
  uint32_t* done_flag;
  uint32_t done_value;
};

struct CopyLaunchHints {
  bool host_memory = false;  // one side lives in pinned host memory: few CTAs, vector path
  int device = -1;
};

// Fills tile geometry, picks path/vector width from `align` (see copy_alignment) and options, launches on `stream`.
// Returns 0 / cudaError_t / KVDP_E*.
int launch_strided_copy(CopyArgs args, uint32_t align, const CopyLaunchHints& hints, cudaStream_t stream);

// Largest power-of-two (<=16) dividing every address/stride of the job; used for the vector width.
uint32_t copy_alignment(const CopyArgs& a, const uint64_t* h_src_ptrs, int n_src, const uint64_t* h_dst_ptrs,
                        int n_dst);

}  // namespace kvdp
