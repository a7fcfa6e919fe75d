// common.cuh — shared helpers for libkvdp.so (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/kvdp.h"

namespace kvdp {

// ---------------------------------------------------------------- error plumbing
void set_error(const char* fmt, ...);
int fail(int code, const char* fmt, ...);  // records the message, returns code

#define KVDP_CUDA(call)                                                                                   \
  do {                                                                                                    \
    cudaError_t e__ = (call);                                                                             \
    if (e__ != cudaSuccess) {                                                                             \
      return ::kvdp::fail((int)e__, "%s failed at %s:%d: %s", #call, __FILE__, __LINE__,                  \
                          cudaGetErrorString(e__));                                                       \
    }                                                                                                     \
  } while (0)

// ---------------------------------------------------------------- options (kvdp_set_option)
struct Options {
  int64_t copy_path = 0;    // 0 auto, 1 vector ld/st, 2 TMA load + vector store, 3 TMA load + TMA store, 4 small chunks
  int64_t tile_bytes = 0;   // 0 = default per path
  int64_t stages = 0;       // 0 = default per path
  int64_t ctas_per_sm = 0;  // 0 = default per path
  int64_t host_path = 0;    // pinned host memory: 0 auto (CopyStream: coalesced DMA when the host-side ids form long
                            // runs, else the vector kernel), 1 per-chunk DMA loop, 2 TMA kernel, 3 vector kernel
  int64_t host_ctas = 0;    // CTAs used when one side is host memory (0 = default)
  int64_t host_dma_piece_kib = 8192;  // coalesced DMA path: runs are issued in pieces of at most this size (0 = whole runs)
  int64_t l2_hint = 0;      // path 3: 1 = evict_first on loads, 2 = on stores, 3 = both
  int64_t vec32 = 0;        // path 1: 1 = 256-bit loads/stores when everything is 32-byte aligned
  int64_t hash_stages = 0;  // chained hashing of a big batch: request ranges whose chains overlap the next range's
                            // block hashing (0 = default 4, 1 = everything on the caller's stream)
};
Options& options();

// device properties cached per device
struct DevInfo {
  int sm_count = 0;
  int max_smem_optin = 0;
};
int device_info(int device, DevInfo* out);

// token_hash.cu: kvdp_hash_token_blocks restricted to the caller's stream (no side streams are created or used)
int hash_token_blocks_single_stream(const uint32_t* d_tokens, const int64_t* d_offsets, int64_t n_req,
                                    int64_t total_tokens, int32_t block_size, uint64_t salt, int64_t* d_block_offsets,
                                    uint64_t* d_block_hash, uint64_t* d_seq_hash, void* d_workspace,
                                    size_t workspace_bytes, uint32_t flags, cudaStream_t stream);

// the event publisher stops feeding an index that is about to be destroyed (event_capi.cu)
void event_sink_forget_index(const void* index);

// RAII device switch
struct DeviceGuard {
  int prev = -1;
  bool changed = false;
  explicit DeviceGuard(int dev) {
    if (dev >= 0 && cudaGetDevice(&prev) == cudaSuccess && prev != dev) {
      changed = cudaSetDevice(dev) == cudaSuccess;
    }
  }
  ~DeviceGuard() {
    if (changed) cudaSetDevice(prev);
  }
};

// ---------------------------------------------------------------- device-side PTX helpers
#if defined(__CUDACC__)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// 16-byte streaming load/store (no L1 allocation: every byte is touched once)
__device__ __forceinline__ int4 ld_stream16(const void* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::256B.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream16(void* p, const int4& v) {
  asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}
__device__ __forceinline__ int4 lds16(uint32_t saddr) {
  int4 r;
  asm volatile("ld.shared.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(saddr));
  return r;
}

// mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// TMA 1-D bulk copies (SASS: UBLKCP). Addresses 16-B aligned, size a multiple of 16.
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void* src_gmem, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
               "l"(src_gmem), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, uint32_t src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(src_smem), "r"(bytes)
               : "memory");
}
// Same copies with an L2 eviction-priority hint (streaming data that will not be re-read soon)
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_g2s_hint(uint32_t dst_smem, const void* src_gmem, uint32_t bytes, uint32_t bar,
                                              uint64_t pol) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          dst_smem),
      "l"(src_gmem), "r"(bytes), "r"(bar), "l"(pol)
      : "memory");
}
__device__ __forceinline__ void bulk_s2g_hint(void* dst_gmem, uint32_t src_smem, uint32_t bytes, uint64_t pol) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(dst_gmem),
               "r"(src_smem), "r"(bytes), "l"(pol)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
#endif  // __CUDACC__

}  // namespace kvdp
