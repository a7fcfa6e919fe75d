// strided_copy.cu — the block gather/scatter/relayout/reshard kernel family for sm_100a.
//
// Every byte-moving path of the data plane is the same problem: for each (layer, K|V prefix, block pair) a
// CONTIGUOUS source chunk is read once and written either contiguously (block copy, layer-separate <-> contiguous
// relayout) or as equal-length runs to a row-strided destination (TP head reshard, possibly into a peer GPU).
// It replaces the reference's per-chunk cudaMemcpyAsync loops (block_copy.cu:246-305, :577-728), its dead
// copy_blocks_kernel (:41-164) and the Triton rearrange kernels + NIXL descriptor transfers of the patched vLLM
// connector (vllm_v0.8.4-dynamo-kv-disagg-patch.patch:934-1031, :1361-1421) with ONE launch per request.
//
// Three code paths (kvdp_set_option("copy_path")):
//   1  vector  : 16-byte ld.global.nc / st.global, 8 independent loads in flight per thread
//   2  tma-in  : cp.async.bulk global->shared ring filled by one thread, all threads store 16-byte vectors
//                (the permute happens on the store side; stores may target peer-mapped memory)
//   3  tma-tma : cp.async.bulk in and cp.async.bulk out, one elected thread per CTA (contiguous destinations)
// All are HBM- (or NVLink-/PCIe-) bound byte movers: no tensor cores, no arithmetic on the payload.
#include "strided_copy.cuh"

#include <algorithm>

namespace kvdp {

namespace {

constexpr int kVecThreads = 256;
constexpr int kVecUnroll = 8;
constexpr int kTmaStages = 8;      // path 3 ring depth
constexpr int kTmaLoadsAhead = 4;  // path 3: loads in flight; the other stages hold stores in flight
constexpr int kMaxStages2 = 8;     // path 2 ring depth upper bound

struct TileCoord {
  const char* src;  // start of this tile's source bytes
  uint64_t dst_chunk_rel;  // p*dst_prefix_stride + dst_id*dst_block_stride
  uint32_t layer;          // absolute layer index
  uint32_t off;            // byte offset of the tile inside the chunk
  uint32_t bytes;          // bytes in this tile
};

// total_tiles < 2^31 is enforced by the launcher, so the decode stays in 32-bit arithmetic
__device__ __forceinline__ TileCoord decode_tile(const CopyArgs& a, int64_t tile64) {
  const uint32_t tile = (uint32_t)tile64;
  uint32_t c = tile / a.tiles_per_chunk;
  const uint32_t k = tile - c * a.tiles_per_chunk;
  const uint32_t np = (uint32_t)a.num_pairs;
  const uint32_t c2 = c / np;
  const int32_t i = (int32_t)(c - c2 * np);
  const uint32_t c3 = c2 / (uint32_t)a.prefix_dim;
  const int32_t p = (int32_t)(c2 - c3 * (uint32_t)a.prefix_dim);
  const int32_t l = a.layer_begin + (int32_t)c3;
  const int64_t sid = a.src_ids ? (int64_t)__ldg(a.src_ids + i) : (int64_t)i;
  const int64_t did = a.dst_ids ? (int64_t)__ldg(a.dst_ids + i) : (int64_t)i;
  const uint64_t sbase = a.src_tab ? __ldg(a.src_tab + l) : a.src_single;
  TileCoord t;
  t.layer = (uint32_t)l;
  t.off = k * a.tile_bytes;
  t.bytes = min(a.tile_bytes, a.chunk_bytes - t.off);
  t.src = (const char*)sbase + (int64_t)p * a.src_prefix_stride + sid * a.src_block_stride + t.off;
  t.dst_chunk_rel = (uint64_t)((int64_t)p * a.dst_prefix_stride + did * a.dst_block_stride);
  return t;
}

__device__ __forceinline__ uint64_t dst_base(const CopyArgs& a, uint32_t layer, uint32_t g) {
  return a.dst_tab ? __ldg(a.dst_tab + (size_t)g * a.num_layers_total + layer)
                   : a.dst_single + (uint64_t)((int64_t)g * a.dst_group_stride);
}

// destination address of chunk byte offset `off` (start of a VEC-aligned piece that never straddles a run)
template <bool CONTIG>
__device__ __forceinline__ char* dst_addr(const CopyArgs& a, const TileCoord& t, uint64_t base0, uint32_t off) {
  if (CONTIG) return (char*)(base0 + t.dst_chunk_rel + off);
  const uint32_t q = off / a.run_bytes;
  const uint32_t x = off - q * a.run_bytes;
  uint32_t r = q, g = 0;
  if (a.row_groups > 1) {
    r = q / (uint32_t)a.row_groups;
    g = q - r * (uint32_t)a.row_groups;
  }
  const uint64_t base = (g == 0) ? base0 : dst_base(a, t.layer, g);
  return (char*)(base + t.dst_chunk_rel + (uint64_t)r * (uint64_t)a.dst_row_stride + (uint64_t)a.dst_col_offset + x);
}

// source address of chunk byte offset `off` when source rows are strided (t.src then holds the chunk base + tile.off;
// the tile offset is removed again here)
__device__ __forceinline__ const char* src_addr_strided(const CopyArgs& a, const TileCoord& t, uint32_t off) {
  const uint32_t rowlen = a.run_bytes * (uint32_t)a.row_groups;
  const uint32_t r = off / rowlen;
  const uint32_t x = off - r * rowlen;
  return t.src - t.off + (uint64_t)r * (uint64_t)a.src_row_stride + (uint64_t)a.src_col_offset + x;
}

__device__ __forceinline__ void publish_done(const CopyArgs& a) {
  if (a.done_flag == nullptr) return;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(a.done_counter, 1u);
    if (prev == gridDim.x - 1) {
      *a.done_counter = 0;  // ready for the next launch on this stream
      __threadfence_system();
      st_release_sys_u32(a.done_flag, a.done_value);
    }
  }
}

template <int VEC>
struct VecT;
struct alignas(32) Vec32 {
  uint32_t w[8];
};
template <>
struct VecT<32> {
  using T = Vec32;
};
template <>
struct VecT<16> {
  using T = int4;
};
template <>
struct VecT<8> {
  using T = int2;
};
template <>
struct VecT<4> {
  using T = int;
};
template <>
struct VecT<2> {
  using T = short;
};
template <>
struct VecT<1> {
  using T = char;
};

template <int VEC>
__device__ __forceinline__ typename VecT<VEC>::T ld_vec(const char* p) {
  if constexpr (VEC == 32) {  // 256-bit accesses (sm_100: LDG.E.256 / STG.E.256)
    Vec32 r;
    asm volatile("ld.global.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]),
                   "=r"(r.w[7])
                 : "l"(p));
    return r;
  } else if constexpr (VEC == 16) {
    return ld_stream16(p);
  } else {
    return __ldg(reinterpret_cast<const typename VecT<VEC>::T*>(p));
  }
}
template <int VEC>
__device__ __forceinline__ void st_vec(char* p, const typename VecT<VEC>::T& v) {
  if constexpr (VEC == 32) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v.w[0]), "r"(v.w[1]), "r"(v.w[2]),
                 "r"(v.w[3]), "r"(v.w[4]), "r"(v.w[5]), "r"(v.w[6]), "r"(v.w[7])
                 : "memory");
  } else if constexpr (VEC == 16) {
    st_stream16(p, v);
  } else {
    *reinterpret_cast<typename VecT<VEC>::T*>(p) = v;
  }
}

// ------------------------------------------------------------------------------------------------ path 1
template <int VEC, bool CONTIG, bool SRC_STRIDED = false>
__global__ void __launch_bounds__(kVecThreads) kvdp_copy_vec_kernel(const __grid_constant__ CopyArgs a) {
  using V = typename VecT<VEC>::T;
  for (int64_t tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x) {
    const TileCoord t = decode_tile(a, tile);
    const uint64_t base0 = dst_base(a, t.layer, 0);
    const uint32_t nvec = t.bytes / VEC;
    for (uint32_t v0 = threadIdx.x; v0 < nvec; v0 += kVecThreads * kVecUnroll) {
      V regs[kVecUnroll];
#pragma unroll
      for (int u = 0; u < kVecUnroll; ++u) {
        const uint32_t v = v0 + u * kVecThreads;
        if (v < nvec)
          regs[u] = ld_vec<VEC>(SRC_STRIDED ? src_addr_strided(a, t, t.off + v * VEC) : t.src + (size_t)v * VEC);
      }
#pragma unroll
      for (int u = 0; u < kVecUnroll; ++u) {
        const uint32_t v = v0 + u * kVecThreads;
        if (v < nvec) st_vec<VEC>(dst_addr<CONTIG>(a, t, base0, t.off + v * VEC), regs[u]);
      }
    }
  }
  publish_done(a);
}

// ------------------------------------------------------------------------------------------------ path 4
// Small chunks (token-granular rows, <= 4 KiB per (layer, K|V, pair)): a CTA per chunk would leave most threads idle
// and pay the index decode and two dependent id / table loads once per chunk. Here a GROUP of lanes (a whole warp for
// 2-4 KiB, down to 4 lanes for <= 256 B, so that a lane has about kSmallVecs vectors per chunk) owns kSmallChunks
// chunks at a time: it decodes them together, then every lane has up to kSmallChunks * kSmallVecs loads in flight
// before the first store. Lanes of a group touch consecutive vectors: full-line accesses.
constexpr int kSmallChunks = 2;  // measured: 2 chunks x 3 CTAs/SM (77 registers) beats 4 chunks x 2 CTAs/SM (128) by 13 %
constexpr int kSmallVecs = 4;
constexpr uint32_t kSmallChunkMax = 4096;
template <int VEC, bool CONTIG, bool SRC_STRIDED>
__global__ void __launch_bounds__(kVecThreads, 3) kvdp_copy_small_kernel(const __grid_constant__ CopyArgs a) {
  using V = typename VecT<VEC>::T;
  const uint32_t G = a.group_lanes;
  const uint32_t gl = threadIdx.x & (G - 1);  // lane inside the group
  const int64_t groups_total = (int64_t)gridDim.x * (kVecThreads / G);
  const int64_t group = (int64_t)blockIdx.x * (kVecThreads / G) + threadIdx.x / G;
  const uint32_t nvec = a.chunk_bytes / VEC;  // tiles_per_chunk == 1: every tile is a whole chunk
  for (int64_t c0 = group * kSmallChunks; c0 < a.total_tiles; c0 += groups_total * kSmallChunks) {
    TileCoord t[kSmallChunks];
    uint64_t base0[kSmallChunks];
#pragma unroll
    for (int u = 0; u < kSmallChunks; ++u) {
      if (c0 + u < a.total_tiles) {
        t[u] = decode_tile(a, c0 + u);
        base0[u] = dst_base(a, t[u].layer, 0);
      }
    }
    for (uint32_t v0 = gl; v0 < nvec; v0 += G * kSmallVecs) {
      V regs[kSmallChunks][kSmallVecs];
#pragma unroll
      for (int u = 0; u < kSmallChunks; ++u) {
#pragma unroll
        for (int k = 0; k < kSmallVecs; ++k) {
          const uint32_t v = v0 + k * G;
          if (c0 + u < a.total_tiles && v < nvec)
            regs[u][k] = ld_vec<VEC>(SRC_STRIDED ? src_addr_strided(a, t[u], v * VEC) : t[u].src + (size_t)v * VEC);
        }
      }
#pragma unroll
      for (int u = 0; u < kSmallChunks; ++u) {
#pragma unroll
        for (int k = 0; k < kSmallVecs; ++k) {
          const uint32_t v = v0 + k * G;
          if (c0 + u < a.total_tiles && v < nvec) st_vec<VEC>(dst_addr<CONTIG>(a, t[u], base0[u], v * VEC), regs[u][k]);
        }
      }
    }
  }
  publish_done(a);
}

// ------------------------------------------------------------------------------------------------ path 2
// smem: [stages][tile_bytes] ring, then `stages` full-barriers. Thread 0 keeps `stages` bulk loads in flight.
template <bool CONTIG>
__global__ void __launch_bounds__(kVecThreads) kvdp_copy_tma_in_kernel(const __grid_constant__ CopyArgs a,
                                                                       const int stages) {
  extern __shared__ __align__(128) unsigned char smem[];
  const uint32_t ring = smem_u32(smem);
  const uint32_t bars = ring + (uint32_t)stages * a.tile_bytes;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(bars + 8 * s, 1);
    mbar_fence_init();
  }
  __syncthreads();

  const int64_t first = blockIdx.x;
  const int64_t step = gridDim.x;
  const int64_t n = first < a.total_tiles ? (a.total_tiles - first + step - 1) / step : 0;
  const int ahead = stages;  // every stage holds a bulk load in flight

  if (threadIdx.x == 0) {
    for (int64_t j = 0; j < ahead && j < n; ++j) {
      const TileCoord t = decode_tile(a, first + j * step);
      const uint32_t s = (uint32_t)(j % stages);
      mbar_expect_tx(bars + 8 * s, t.bytes);
      bulk_g2s(ring + s * a.tile_bytes, t.src, t.bytes, bars + 8 * s);
    }
  }
  for (int64_t j = 0; j < n; ++j) {
    const uint32_t s = (uint32_t)(j % stages);
    const TileCoord t = decode_tile(a, first + j * step);
    const uint64_t base0 = dst_base(a, t.layer, 0);
    mbar_wait(bars + 8 * s, (uint32_t)((j / stages) & 1));
    const uint32_t nvec = t.bytes / 16;
    const uint32_t sbase = ring + s * a.tile_bytes;
#pragma unroll 4
    for (uint32_t v = threadIdx.x; v < nvec; v += kVecThreads) {
      const int4 val = lds16(sbase + v * 16);
      st_stream16(dst_addr<CONTIG>(a, t, base0, t.off + v * 16), val);
    }
    __syncthreads();  // every thread is done reading stage s of tile j (and all older stages)
    if (threadIdx.x == 0) {
      const int64_t jn = j + ahead;  // refills the stage tile j just vacated (all threads passed the barrier)
      if (jn < n) {
        const TileCoord tn = decode_tile(a, first + jn * step);
        const uint32_t sn = (uint32_t)(jn % stages);
        fence_proxy_async_smem();
        mbar_expect_tx(bars + 8 * sn, tn.bytes);
        bulk_g2s(ring + sn * a.tile_bytes, tn.src, tn.bytes, bars + 8 * sn);
      }
    }
  }
  publish_done(a);
}

// ------------------------------------------------------------------------------------------------ path 3
// One warp per CTA, lane 0 drives an 8-stage ring: 4 bulk loads and up to 4 bulk stores in flight.
__device__ __forceinline__ void bulk_wait_read_dyn(int pending) {
  switch (pending) {
    case 0: bulk_wait_read<0>(); break;
    case 1: bulk_wait_read<1>(); break;
    case 2: bulk_wait_read<2>(); break;
    case 3: bulk_wait_read<3>(); break;
    default: bulk_wait_read<4>(); break;
  }
}

__global__ void __launch_bounds__(32) kvdp_copy_tma_tma_kernel(const __grid_constant__ CopyArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const uint32_t ring = smem_u32(smem);
  const uint32_t bars = ring + (uint32_t)kTmaStages * a.tile_bytes;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kTmaStages; ++s) mbar_init(bars + 8 * s, 1);
    mbar_fence_init();
    const int64_t first = blockIdx.x;
    const int64_t step = gridDim.x;
    const int64_t n = first < a.total_tiles ? (a.total_tiles - first + step - 1) / step : 0;
    const uint64_t pol = a.l2_hint ? l2_policy_evict_first() : 0;
    for (int64_t it = 0; it < n + kTmaLoadsAhead; ++it) {
      const int64_t j = it - kTmaLoadsAhead;  // tile whose store is issued now
      if (j >= 0 && j < n) {
        const TileCoord t = decode_tile(a, first + j * step);
        const uint32_t s = (uint32_t)(j % kTmaStages);
        mbar_wait(bars + 8 * s, (uint32_t)((j / kTmaStages) & 1));
        fence_proxy_async_smem();
        char* dptr = (char*)(dst_base(a, t.layer, 0) + t.dst_chunk_rel + t.off);
        if (a.l2_hint & 2)
          bulk_s2g_hint(dptr, ring + s * a.tile_bytes, t.bytes, pol);
        else
          bulk_s2g(dptr, ring + s * a.tile_bytes, t.bytes);
        bulk_commit();
      }
      if (it < n) {
        const TileCoord t = decode_tile(a, first + it * step);
        const uint32_t s = (uint32_t)(it % kTmaStages);
        // the store that last read this stage (tile it-8) must have drained its shared-memory reads
        if (it >= kTmaStages) bulk_wait_read_dyn(kTmaStages - kTmaLoadsAhead);
        mbar_expect_tx(bars + 8 * s, t.bytes);
        if (a.l2_hint & 1)
          bulk_g2s_hint(ring + s * a.tile_bytes, t.src, t.bytes, bars + 8 * s, pol);
        else
          bulk_g2s(ring + s * a.tile_bytes, t.src, t.bytes, bars + 8 * s);
      }
    }
    bulk_wait_all<0>();  // all global writes of this CTA are complete
    asm volatile("fence.proxy.async;" ::: "memory");
  }
  __syncwarp();
  publish_done(a);
}

template <typename K>
int set_smem(K kernel, size_t bytes) {
  if (bytes > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute(smem=%zu): %s", bytes, cudaGetErrorString(e));
  }
  return 0;
}

template <int VEC>
int launch_small(const CopyArgs& a, bool contig, int grid, cudaStream_t stream) {
  if (a.src_row_stride != 0)
    kvdp_copy_small_kernel<VEC, false, true><<<grid, kVecThreads, 0, stream>>>(a);
  else if (contig)
    kvdp_copy_small_kernel<VEC, true, false><<<grid, kVecThreads, 0, stream>>>(a);
  else
    kvdp_copy_small_kernel<VEC, false, false><<<grid, kVecThreads, 0, stream>>>(a);
  return 0;
}

template <int VEC>
int launch_vec(const CopyArgs& a, bool contig, int grid, cudaStream_t stream) {
  if (a.src_row_stride != 0)
    kvdp_copy_vec_kernel<VEC, false, true><<<grid, kVecThreads, 0, stream>>>(a);
  else if (contig)
    kvdp_copy_vec_kernel<VEC, true><<<grid, kVecThreads, 0, stream>>>(a);
  else
    kvdp_copy_vec_kernel<VEC, false><<<grid, kVecThreads, 0, stream>>>(a);
  return 0;
}

}  // namespace

uint32_t copy_alignment(const CopyArgs& a, const uint64_t* h_src_ptrs, int n_src, const uint64_t* h_dst_ptrs,
                        int n_dst) {
  uint64_t m = 32;
  auto fold = [&m](uint64_t v) {
    while (m > 1 && (v % m) != 0) m >>= 1;
  };
  for (int i = 0; i < n_src; ++i) fold(h_src_ptrs[i]);
  for (int i = 0; i < n_dst; ++i) fold(h_dst_ptrs[i]);
  fold((uint64_t)a.run_bytes);
  fold((uint64_t)a.src_prefix_stride);
  fold((uint64_t)a.src_block_stride);
  fold((uint64_t)a.dst_prefix_stride);
  fold((uint64_t)a.dst_block_stride);
  fold((uint64_t)a.dst_row_stride);
  fold((uint64_t)a.dst_col_offset);
  fold((uint64_t)a.dst_group_stride);
  fold((uint64_t)a.src_row_stride);
  fold((uint64_t)a.src_col_offset);
  return (uint32_t)m;
}

int launch_strided_copy(CopyArgs a, uint32_t align, const CopyLaunchHints& hints, cudaStream_t stream) {
  if (a.num_pairs <= 0 || a.layers_run <= 0) return 0;
  if (a.run_bytes == 0 || a.rows <= 0 || a.row_groups <= 0 || a.prefix_dim <= 0)
    return fail(KVDP_EINVAL, "strided copy: empty geometry (run=%u rows=%d groups=%d prefix=%d)", a.run_bytes, a.rows,
                a.row_groups, a.prefix_dim);

  DevInfo di;
  int dev = hints.device;
  if (dev < 0) KVDP_CUDA(cudaGetDevice(&dev));
  if (int rc = device_info(dev, &di)) return rc;

  const Options& opt = options();
  const bool vec32 = align >= 32 && opt.vec32 != 0;  // 256-bit accesses: path 1 only, opt-in
  if (align > 16) align = 16;
  // a row-strided destination that is in fact dense collapses to the contiguous case
  const bool contig =
      a.row_groups == 1 && a.dst_col_offset == 0 && (a.rows == 1 || a.dst_row_stride == (int64_t)a.run_bytes);
  int path = (int)opt.copy_path;
  const bool host_tma = hints.host_memory && opt.host_path == 2;
  const bool must_vec = align < 16 || (hints.host_memory && !host_tma) || a.src_row_stride != 0;
  const bool small_ok = a.chunk_bytes <= kSmallChunkMax && !hints.host_memory;
  a.l2_hint = (uint32_t)opt.l2_hint;
  if (path == 4 && !small_ok) path = 0;
  if (must_vec && path != 4) path = (small_ok && opt.copy_path == 0) ? 4 : 1;
  if (path == 0) path = small_ok ? 4 : (contig ? 3 : 2);
  if (path == 3 && !contig) path = 2;

  // tile geometry: tiles are contiguous source slices; a tile is a whole number of vectors and, on the
  // strided paths, starts on a vector boundary of a run (guaranteed because run_bytes % align == 0).
  uint32_t tile = (uint32_t)opt.tile_bytes;
  if (tile == 0) tile = (path == 1) ? 32768u : 16384u;
  tile = std::max(tile, 1024u) & ~1023u;
  if (tile > a.chunk_bytes) tile = a.chunk_bytes;  // small chunks: one tile each
  if (path == 4) tile = a.chunk_bytes;
  if ((path == 2 || path == 3) && (tile % 16) != 0) path = 1;
  a.tile_bytes = tile;
  a.tiles_per_chunk = (a.chunk_bytes + tile - 1) / tile;
  a.total_tiles = (int64_t)a.layers_run * a.prefix_dim * a.num_pairs * a.tiles_per_chunk;
  if (a.total_tiles >= (int64_t)1 << 31)
    return fail(KVDP_ERANGE, "strided copy: %lld tiles in one launch (max 2^31-1)", (long long)a.total_tiles);

  int ctas_per_sm = (int)opt.ctas_per_sm;
  int64_t grid64;
  cudaError_t e = cudaSuccess;
  if (path == 1) {
    if (ctas_per_sm <= 0) ctas_per_sm = 4;
    grid64 = (int64_t)di.sm_count * ctas_per_sm;
    if (hints.host_memory) grid64 = opt.host_ctas > 0 ? opt.host_ctas : 32;
    grid64 = std::min<int64_t>(grid64, a.total_tiles);
    const int grid = (int)std::max<int64_t>(grid64, 1);
    switch (vec32 ? 32u : align) {
      case 32: launch_vec<32>(a, contig, grid, stream); break;
      case 16: launch_vec<16>(a, contig, grid, stream); break;
      case 8: launch_vec<8>(a, contig, grid, stream); break;
      case 4: launch_vec<4>(a, contig, grid, stream); break;
      case 2: launch_vec<2>(a, contig, grid, stream); break;
      default: launch_vec<1>(a, contig, grid, stream); break;
    }
  } else if (path == 4) {
    if (ctas_per_sm <= 0) ctas_per_sm = 3;
    const uint32_t nvec = a.chunk_bytes / align;
    uint32_t G = 32;
    while (G > 4 && nvec <= (G / 2) * kSmallVecs) G >>= 1;
    a.group_lanes = G;
    const int64_t groups_needed = (a.total_tiles + kSmallChunks - 1) / kSmallChunks;
    const int64_t per_cta = kVecThreads / G;
    grid64 = std::min<int64_t>((int64_t)di.sm_count * ctas_per_sm, (groups_needed + per_cta - 1) / per_cta);
    const int grid = (int)std::max<int64_t>(grid64, 1);
    switch (align) {
      case 16: launch_small<16>(a, contig, grid, stream); break;
      case 8: launch_small<8>(a, contig, grid, stream); break;
      case 4: launch_small<4>(a, contig, grid, stream); break;
      case 2: launch_small<2>(a, contig, grid, stream); break;
      default: launch_small<1>(a, contig, grid, stream); break;
    }
  } else if (path == 2) {
    int stages = (int)opt.stages;
    if (stages <= 1) stages = 4;
    stages = std::min(stages, kMaxStages2);
    size_t smem = (size_t)stages * tile + 8 * stages;
    while (smem > (size_t)di.max_smem_optin && stages > 2) {
      --stages;
      smem = (size_t)stages * tile + 8 * stages;
    }
    if (ctas_per_sm <= 0) ctas_per_sm = std::max(1, std::min(4, (int)((size_t)di.max_smem_optin / (smem + 1024))));
    grid64 = std::min<int64_t>((int64_t)di.sm_count * ctas_per_sm, a.total_tiles);
    const int grid = (int)std::max<int64_t>(grid64, 1);
    if (contig) {
      if (int rc = set_smem(kvdp_copy_tma_in_kernel<true>, smem)) return rc;
      kvdp_copy_tma_in_kernel<true><<<grid, kVecThreads, smem, stream>>>(a, stages);
    } else {
      if (int rc = set_smem(kvdp_copy_tma_in_kernel<false>, smem)) return rc;
      kvdp_copy_tma_in_kernel<false><<<grid, kVecThreads, smem, stream>>>(a, stages);
    }
  } else {
    const size_t smem = (size_t)kTmaStages * tile + 8 * kTmaStages;
    if (smem > (size_t)di.max_smem_optin)
      return fail(KVDP_EINVAL, "tile_bytes=%u needs %zu B shared memory (> %d)", tile, smem, di.max_smem_optin);
    if (ctas_per_sm <= 0) ctas_per_sm = std::max(1, std::min(8, (int)((size_t)di.max_smem_optin / (smem + 1024))));
    grid64 = std::min<int64_t>((int64_t)di.sm_count * ctas_per_sm, a.total_tiles);
    if (hints.host_memory) grid64 = std::min<int64_t>(grid64, opt.host_ctas > 0 ? opt.host_ctas : 32);
    const int grid = (int)std::max<int64_t>(grid64, 1);
    if (int rc = set_smem(kvdp_copy_tma_tma_kernel, smem)) return rc;
    kvdp_copy_tma_tma_kernel<<<grid, 32, smem, stream>>>(a);
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail((int)e, "strided copy launch (path %d): %s", path, cudaGetErrorString(e));
  return 0;
}

}  // namespace kvdp
