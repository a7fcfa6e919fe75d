// token_hash.cu — batched XXH3-64 token-block hashing on the GPU.
//
// Replaces the CPU hashing of the reference's token library and router indexer:
//   block hash     lib/tokens/src/lib.rs:197-215  (TokenBlockChunk::new / from_tokens -> compute_hash :44-46)
//   sequence hash  lib/tokens/src/lib.rs:238-261  (TokenBlock::from_chunk), chain loop :382-391
//   split_tokens   lib/tokens/src/lib.rs:369-404  (rayon par_chunks_exact over blocks, serial chain)
//   router hash    lib/llm/src/kv_router/indexer.rs:64,87-89,123-135 (seed 1337, un-chained)
//   twin quirk     lib/llm/src/tokens.rs:126-144
// The arithmetic is XXH3-64-with-seed (xxhash-rust 0.8.15 in the reference, a frozen public algorithm); it is
// written here from the specification, not from any implementation's source.
//
// Launch structure for one ragged batch (CSR offsets):
//   1. kvdp_block_count_scan_*  : block_offsets[r] = sum_{q<r} len_q / block_size       (1 CTA up to 2048 requests,
//                                 else partials / spine / apply)
//   2. kvdp_hash_blocks_*       : one thread per token block. Fast path (256/512-byte blocks): each lane pulls its
//                                 own block into a padded shared-memory row with one cp.async.bulk (TMA), so global
//                                 reads are full-line and shared reads are conflict-free 16-byte loads; the seed-derived
//                                 secret words arrive through the kernel-parameter constant bank.
//   3. kvdp_hash_chain          : one thread per request walks its blocks (16-byte XXH3 per step, serial by nature;
//                                 a register ring prefetches the next 8 block hashes to hide the loads).
// HBM-bound by design: 4*block_size bytes read + 16 bytes written per block; no tensor cores (nothing contracts).
#include <string.h>

#include <algorithm>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace kvdp {
namespace {

// ------------------------------------------------------------------------------------------------ constants
constexpr uint64_t P32_1 = 0x9E3779B1ULL, P32_2 = 0x85EBCA77ULL, P32_3 = 0xC2B2AE3DULL;
constexpr uint64_t P64_1 = 0x9E3779B185EBCA87ULL, P64_2 = 0xC2B2AE3D27D4EB4FULL, P64_3 = 0x165667B19E3779F9ULL,
                   P64_4 = 0x85EBCA77C2B2AE63ULL, P64_5 = 0x27D4EB2F165667C5ULL;
constexpr uint64_t MX1 = 0x165667919E3779F9ULL, MX2 = 0x9FB21C651E98DF25ULL;

// default 192-byte secret as 24 little-endian words
#define KVDP_SECRET_WORDS                                                                                         \
  {0xbe4ba423396cfeb8ULL, 0x1cad21f72c81017cULL, 0xdb979083e96dd4deULL, 0x1f67b3b7a4a44072ULL, 0x78e5c0cc4ee679cbULL, \
   0x2172ffcc7dd05a82ULL, 0x8e2443f7744608b8ULL, 0x4c263a81e69035e0ULL, 0xcb00c391bb52283cULL, 0xa32e531b8b65d088ULL, \
   0x4ef90da297486471ULL, 0xd8acdea946ef1938ULL, 0x3f349ce33f76faa8ULL, 0x1d4f0bc7c7bbdcf9ULL, 0x3159b4cd4be0518aULL, \
   0x647378d9c97e9fc8ULL, 0xc3ebd33483acc5eaULL, 0xeb6313faffa081c5ULL, 0x49daf0b751dd0d17ULL, 0x9e68d429265516d3ULL, \
   0xfca1477d58be162bULL, 0xce31d07ad1b8f88fULL, 0x280416958f3acb45ULL, 0x7e404bbbcafbd7afULL}

__constant__ uint64_t c_secret[24] = KVDP_SECRET_WORDS;
const uint64_t h_secret[24] = KVDP_SECRET_WORDS;

// Seed-derived values handed to the kernels through the parameter constant bank.
struct HashSecrets {
  uint64_t w[24];    // custom-secret words: k[j] + seed (j even), k[j] - seed (j odd)
  uint64_t last[8];  // rd64(S + 121 + 8l): secret of the final stripe
  uint64_t fin[8];   // rd64(S + 11 + 8k): secret of the merge step
  uint64_t flip16_lo;  // (k[3]^k[4]) + seed : 9..16-byte class
  uint64_t flip16_hi;  // (k[5]^k[6]) - seed
  uint64_t seed;
};

HashSecrets make_secrets(uint64_t seed) {
  HashSecrets s;
  for (int j = 0; j < 24; ++j) s.w[j] = (j & 1) ? h_secret[j] - seed : h_secret[j] + seed;
  // unaligned reads of the byte string whose words are s.w: offset 121 = 15*8 + 1, offset 11 = 1*8 + 3
  for (int l = 0; l < 8; ++l) s.last[l] = (s.w[15 + l] >> 8) | (s.w[16 + l] << 56);
  for (int k = 0; k < 8; ++k) s.fin[k] = (s.w[1 + k] >> 24) | (s.w[2 + k] << 40);
  s.flip16_lo = (h_secret[3] ^ h_secret[4]) + seed;
  s.flip16_hi = (h_secret[5] ^ h_secret[6]) - seed;
  s.seed = seed;
  return s;
}

// ------------------------------------------------------------------------------------------------ device math
__device__ __forceinline__ uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
__device__ __forceinline__ uint32_t bswap32(uint32_t x) { return __byte_perm(x, 0, 0x0123); }
__device__ __forceinline__ uint64_t bswap64(uint64_t x) {
  return ((uint64_t)bswap32((uint32_t)x) << 32) | bswap32((uint32_t)(x >> 32));
}
__device__ __forceinline__ uint64_t fold128(uint64_t a, uint64_t b) { return (a * b) ^ __umul64hi(a, b); }
__device__ __forceinline__ uint64_t avalanche(uint64_t h) {
  h ^= h >> 37;
  h *= MX1;
  h ^= h >> 32;
  return h;
}
__device__ __forceinline__ uint64_t xxh64_avalanche(uint64_t h) {
  h ^= h >> 33;
  h *= P64_2;
  h ^= h >> 29;
  h *= P64_3;
  h ^= h >> 32;
  return h;
}
__device__ __forceinline__ uint64_t rrmxmx(uint64_t h, uint64_t len) {
  h ^= rotl64(h, 49) ^ rotl64(h, 24);
  h *= MX2;
  h ^= (h >> 35) + len;
  h *= MX2;
  h ^= h >> 28;
  return h;
}
__device__ __forceinline__ uint64_t mul32x32(uint64_t k) { return (uint64_t)(uint32_t)k * (uint64_t)(uint32_t)(k >> 32); }

// rd64 of the default secret at an arbitrary byte offset
__device__ __forceinline__ uint64_t ksec64(uint32_t off) {
  const uint32_t j = off >> 3, sh = (off & 7) * 8;
  const uint64_t lo = c_secret[j];
  if (sh == 0) return lo;
  return (lo >> sh) | (c_secret[j + 1] << (64 - sh));
}
// rd64 of the seed-derived secret at an arbitrary byte offset
__device__ __forceinline__ uint64_t dsec64(const HashSecrets& S, uint32_t off) {
  const uint32_t j = off >> 3, sh = (off & 7) * 8;
  const uint64_t lo = (j & 1) ? c_secret[j] - S.seed : c_secret[j] + S.seed;
  if (sh == 0) return lo;
  const uint64_t hi = ((j + 1) & 1) ? c_secret[j + 1] - S.seed : c_secret[j + 1] + S.seed;
  return (lo >> sh) | (hi << (64 - sh));
}

// 16-byte class with both input words in registers: the chain step hash([parent, block_hash]).
__device__ __forceinline__ uint64_t xxh3_16(uint64_t in0, uint64_t in1, const HashSecrets& S) {
  const uint64_t lo = in0 ^ S.flip16_lo;
  const uint64_t hi = in1 ^ S.flip16_hi;
  return avalanche(16ULL + bswap64(lo) + hi + fold128(lo, hi));
}

// Byte readers for the generic path. WordReader needs 4-byte alignment of base and offsets (token data).
struct ByteReader {
  const uint8_t* p;
  __device__ __forceinline__ uint32_t u8(uint32_t o) const { return p[o]; }
  __device__ __forceinline__ uint32_t u32(uint32_t o) const {
    return (uint32_t)p[o] | ((uint32_t)p[o + 1] << 8) | ((uint32_t)p[o + 2] << 16) | ((uint32_t)p[o + 3] << 24);
  }
  __device__ __forceinline__ uint64_t u64(uint32_t o) const { return (uint64_t)u32(o) | ((uint64_t)u32(o + 4) << 32); }
};
struct WordReader {
  const uint32_t* p;
  __device__ __forceinline__ uint32_t u8(uint32_t o) const { return (p[o >> 2] >> ((o & 3) * 8)) & 0xff; }
  __device__ __forceinline__ uint32_t u32(uint32_t o) const { return p[o >> 2]; }
  __device__ __forceinline__ uint64_t u64(uint32_t o) const {
    return (uint64_t)p[o >> 2] | ((uint64_t)p[(o >> 2) + 1] << 32);
  }
};

template <typename R>
__device__ __forceinline__ uint64_t mix16(const R& in, uint32_t o, uint32_t so, uint64_t seed) {
  return fold128(in.u64(o) ^ (ksec64(so) + seed), in.u64(o + 8) ^ (ksec64(so + 8) - seed));
}

template <typename R>
__device__ __forceinline__ void accumulate_stripe(uint64_t acc[8], const R& in, uint32_t o, const HashSecrets& S,
                                                  uint32_t so) {
#pragma unroll
  for (int l = 0; l < 8; l += 2) {
    const uint64_t a0 = in.u64(o + 8 * l), a1 = in.u64(o + 8 * l + 8);
    const uint64_t k0 = a0 ^ dsec64(S, so + 8 * l), k1 = a1 ^ dsec64(S, so + 8 * l + 8);
    acc[l] += a1 + mul32x32(k0);
    acc[l + 1] += a0 + mul32x32(k1);
  }
}

// Every XXH3-64 length class, any length; the slow but complete path.
template <typename R>
__device__ uint64_t xxh3_generic(const R& in, uint32_t len, const HashSecrets& S) {
  const uint64_t seed = S.seed;
  if (len == 0) return xxh64_avalanche(seed ^ ksec64(56) ^ ksec64(64));
  if (len <= 3) {
    const uint32_t c1 = in.u8(0), c2 = in.u8(len >> 1), c3 = in.u8(len - 1);
    const uint32_t comb = (c1 << 16) | (c2 << 24) | c3 | (len << 8);
    const uint64_t flip = (uint64_t)((uint32_t)c_secret[0] ^ (uint32_t)(c_secret[0] >> 32)) + seed;
    return xxh64_avalanche((uint64_t)comb ^ flip);
  }
  if (len <= 8) {
    const uint64_t s2 = seed ^ ((uint64_t)bswap32((uint32_t)seed) << 32);
    const uint64_t in64 = (uint64_t)in.u32(len - 4) + ((uint64_t)in.u32(0) << 32);
    const uint64_t flip = (ksec64(8) ^ ksec64(16)) - s2;
    return rrmxmx(in64 ^ flip, len);
  }
  if (len <= 16) {
    const uint64_t lo = in.u64(0) ^ ((ksec64(24) ^ ksec64(32)) + seed);
    const uint64_t hi = in.u64(len - 8) ^ ((ksec64(40) ^ ksec64(48)) - seed);
    return avalanche((uint64_t)len + bswap64(lo) + hi + fold128(lo, hi));
  }
  if (len <= 128) {
    uint64_t acc = (uint64_t)len * P64_1;
    for (int i = (int)((len - 1) / 32); i >= 0; --i) {
      acc += mix16(in, 16 * i, 32 * i, seed);
      acc += mix16(in, len - 16 * (i + 1), 32 * i + 16, seed);
    }
    return avalanche(acc);
  }
  if (len <= 240) {
    uint64_t acc = (uint64_t)len * P64_1;
    for (int i = 0; i < 8; ++i) acc += mix16(in, 16 * i, 16 * i, seed);
    acc = avalanche(acc);
    uint64_t end = mix16(in, len - 16, 136 - 17, seed);
    const int rounds = (int)(len / 16);
    for (int i = 8; i < rounds; ++i) end += mix16(in, 16 * i, 16 * (i - 8) + 3, seed);
    return avalanche(acc + end);
  }
  uint64_t acc[8] = {P32_3, P64_1, P64_2, P64_3, P64_4, P32_2, P64_5, P32_1};
  const uint32_t nb_blocks = (len - 1) / 1024;
  for (uint32_t b = 0; b < nb_blocks; ++b) {
    for (uint32_t n = 0; n < 16; ++n) accumulate_stripe(acc, in, b * 1024 + 64 * n, S, 8 * n);
#pragma unroll
    for (int l = 0; l < 8; ++l) {
      uint64_t a = acc[l];
      a ^= a >> 47;
      a ^= dsec64(S, 128 + 8 * l);
      a *= P32_1;
      acc[l] = a;
    }
  }
  const uint32_t nb_stripes = ((len - 1) - 1024 * nb_blocks) / 64;
  for (uint32_t n = 0; n < nb_stripes; ++n) accumulate_stripe(acc, in, nb_blocks * 1024 + 64 * n, S, 8 * n);
  accumulate_stripe(acc, in, len - 64, S, 121);
  uint64_t r = (uint64_t)len * P64_1;
#pragma unroll
  for (int i = 0; i < 4; ++i) r += fold128(acc[2 * i] ^ dsec64(S, 11 + 16 * i), acc[2 * i + 1] ^ dsec64(S, 11 + 16 * i + 8));
  return avalanche(r);
}

// ------------------------------------------------------------------------------------------------ block -> request
// Largest r with block_offsets[r] <= gb, skipping zero-block requests so that gb < block_offsets[r+1].
__device__ __forceinline__ int64_t find_request(const int64_t* __restrict__ block_offsets, int64_t n_req, int64_t gb0,
                                                int64_t gb) {
  // warp-uniform binary search for the first block of the warp's tile, then a short per-lane walk
  int64_t lo = 0, hi = n_req;  // invariant: block_offsets[lo] <= gb0 < block_offsets[hi] (= total at n_req)
  while (hi - lo > 1) {
    const int64_t mid = (lo + hi) >> 1;
    if (__ldg(block_offsets + mid) <= gb0)
      lo = mid;
    else
      hi = mid;
  }
  int64_t r = lo;
  while (r + 1 < n_req && __ldg(block_offsets + r + 1) <= gb) ++r;
  return r;
}

// Largest r with block_offsets[r] <= gb and block_offsets[r+1] > gb (gb < total).
__device__ __forceinline__ int64_t boff_upper_request(const int64_t* __restrict__ block_offsets, int64_t n_req,
                                                      int64_t gb) {
  return find_request(block_offsets, n_req, gb, gb);
}

struct HashArgs {
  const uint32_t* tokens;
  const int64_t* offsets;        // [n_req+1] token offsets
  const int64_t* block_offsets;  // [n_req+1] block offsets (scan result)
  uint64_t* block_hash;
  uint64_t* seq_hash;
  int64_t n_req;
  int32_t block_size;
  uint32_t flags;
};

// ------------------------------------------------------------------------------------------------ fast block hash
// LEN = 64*NSTRIPES bytes per block (256 for 64-token blocks, 512 for 128-token blocks).
template <int NSTRIPES, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) kvdp_hash_blocks_tma(const __grid_constant__ HashArgs A,
                                                                   const __grid_constant__ HashSecrets S) {
  constexpr int LEN = 64 * NSTRIPES;
  constexpr int ROW = LEN + 16;  // +16 B pad: 16-byte loads of 8 consecutive lanes hit 8 distinct bank groups
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* wbase = smem + (size_t)warp * (32 * ROW);
  const uint32_t row = smem_u32(wbase) + lane * ROW;
  const uint32_t bar = smem_u32(smem + (size_t)WARPS * 32 * ROW) + 8 * warp;
  if (lane == 0) {
    mbar_init(bar, 32);
    mbar_fence_init();
  }
  __syncwarp();

  // A launch covers the blocks of requests [0, n_req) of the (possibly shifted) offset arrays: global block indices
  // [b_begin, total). For a whole batch b_begin is 0; a staged call passes arrays shifted to its first request.
  const int64_t b_begin = __ldg(A.block_offsets), total = __ldg(A.block_offsets + A.n_req);
  const int64_t tiles = (total - b_begin + 31) >> 5;
  // Each warp owns a CONTIGUOUS range of tiles: one binary search for its first block, then the request index only
  // walks forward (a per-tile binary search over 10^4..10^6 requests cost ~40% of the kernel's time).
  const int64_t warps_total = (int64_t)gridDim.x * WARPS;
  const int64_t tiles_per_warp = (tiles + warps_total - 1) / warps_total;
  const int64_t t_begin = min(tiles, ((int64_t)blockIdx.x * WARPS + warp) * tiles_per_warp);
  const int64_t t_end = min(tiles, t_begin + tiles_per_warp);
  int64_t r_tile = -1;  // request holding the first block of the tile being issued (warp-uniform)
  // Start the copies of tile t: every lane arrives once on the warp's barrier, through its copy's byte count or directly.
  auto issue = [&](int64_t t) {
    const int64_t g0 = b_begin + t * 32;  // first block of the tile
    const int64_t gb = g0 + lane;
    if (r_tile < 0) {
      int64_t lo = 0, hi = A.n_req;  // invariant: block_offsets[lo] <= g0 < block_offsets[hi]
      while (hi - lo > 1) {
        const int64_t mid = (lo + hi) >> 1;
        if (__ldg(A.block_offsets + mid) <= g0)
          lo = mid;
        else
          hi = mid;
      }
      r_tile = lo;
    }
    while (r_tile + 1 < A.n_req && __ldg(A.block_offsets + r_tile + 1) <= g0) ++r_tile;
    if (gb < total) {
      int64_t r = r_tile;
      while (r + 1 < A.n_req && __ldg(A.block_offsets + r + 1) <= gb) ++r;
      const uint32_t* src =
          A.tokens + __ldg(A.offsets + r) + (gb - __ldg(A.block_offsets + r)) * (int64_t)A.block_size;
      if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
        mbar_expect_tx(bar, LEN);
        bulk_g2s(row, src, LEN, bar);
      } else {
        // request does not start on a 16-byte boundary: this lane fills its row with 4-byte loads
        uint32_t* dstw = reinterpret_cast<uint32_t*>(wbase + lane * ROW);
        for (int i = 0; i < LEN / 4; ++i) dstw[i] = __ldg(src + i);
        mbar_arrive(bar);
      }
    } else {
      mbar_arrive(bar);
    }
  };
  uint32_t parity = 0;
  if (t_begin < t_end) issue(t_begin);
  for (int64_t t = t_begin; t < t_end; ++t) {
    const int64_t gb = b_begin + t * 32 + lane;
    const bool valid = gb < total;
    mbar_wait(bar, parity);
    parity ^= 1;
    uint64_t hash = 0;
    if (valid) {
      uint64_t acc[8] = {P32_3, P64_1, P64_2, P64_3, P64_4, P32_2, P64_5, P32_1};
#pragma unroll
      for (int n = 0; n < NSTRIPES; ++n) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int4 v = lds16(row + 64 * n + 16 * q);
          const uint64_t a0 = (uint64_t)(uint32_t)v.x | ((uint64_t)(uint32_t)v.y << 32);
          const uint64_t a1 = (uint64_t)(uint32_t)v.z | ((uint64_t)(uint32_t)v.w << 32);
          const uint64_t s0 = (n == NSTRIPES - 1) ? S.last[2 * q] : S.w[n + 2 * q];
          const uint64_t s1 = (n == NSTRIPES - 1) ? S.last[2 * q + 1] : S.w[n + 2 * q + 1];
          acc[2 * q] += a1 + mul32x32(a0 ^ s0);
          acc[2 * q + 1] += a0 + mul32x32(a1 ^ s1);
        }
      }
      uint64_t h = (uint64_t)LEN * P64_1;
#pragma unroll
      for (int i = 0; i < 4; ++i) h += fold128(acc[2 * i] ^ S.fin[2 * i], acc[2 * i + 1] ^ S.fin[2 * i + 1]);
      hash = avalanche(h);
    }
    __syncwarp();                     // every lane has read its row and has seen this phase complete
    // this lane's generic-proxy reads (and, on the unaligned path, writes) of its row are ordered before the
    // async-proxy bulk copy that overwrites the row next
    fence_proxy_async_smem();
    if (t + 1 < t_end) issue(t + 1);  // the next tile's copies fly while this tile's results are stored
    if (valid) A.block_hash[gb] = hash;
  }
}

// ------------------------------------------------------------------------------------------------ short blocks
// Blocks of 16, 32, 64 or 128 bytes (4, 8, 16 or 32 tokens): XXH3's 9-16 and 17-128-byte classes. Both reduce to a
// SUM of independent 16-byte mixes, so P = LEN/16 consecutive lanes share a block: each loads one 16-byte part (a warp
// instruction reads 512 contiguous bytes), mixes it with the secret words of its position, and the P partial sums
// meet in log2(P) shuffles. No shared memory, kPartsUnroll loads in flight per lane. A warp owns a contiguous range of
// blocks, so the block -> request map is one binary search per lane and forward walks after that.
constexpr int kPartsUnroll = 8;
template <int LEN>
__global__ void __launch_bounds__(256) kvdp_hash_blocks_parts(const __grid_constant__ HashArgs A,
                                                              const __grid_constant__ HashSecrets S) {
  constexpr int P = LEN / 16;    // lanes per block
  constexpr int RPI = 32 / P;    // blocks per warp instruction
  constexpr int STEP = RPI * kPartsUnroll;
  const int lane = threadIdx.x & 31;
  const int sub = lane / P, part = lane % P;
  const int64_t b_begin = __ldg(A.block_offsets), total = __ldg(A.block_offsets + A.n_req);  // see kvdp_hash_blocks_tma
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int64_t gw = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  int64_t per_warp = (total - b_begin + warps_total - 1) / warps_total;
  per_warp = (per_warp + STEP - 1) / STEP * STEP;
  const int64_t w_begin = min(total, b_begin + gw * per_warp), w_end = min(total, w_begin + per_warp);
  if (w_begin >= w_end) return;
  // secret words of this lane's part (word index into S.w): first half of the block counts up, second half down
  const int sw = (part < (P + 1) / 2) ? 4 * part : 4 * (P - 1 - part) + 2;
  const uint64_t s0 = S.w[sw], s1 = S.w[sw + 1];
  int64_t r = boff_upper_request(A.block_offsets, A.n_req, w_begin + sub);
  for (int64_t base = w_begin; base < w_end; base += STEP) {
    uint64_t lo[kPartsUnroll], hi[kPartsUnroll];
    const int64_t g0 = base + sub, g_last = g0 + (kPartsUnroll - 1) * RPI;
    // request of this lane's first block in the trip (usually unchanged from the last trip)
    if (g0 < w_end)
      while (r + 1 < A.n_req && __ldg(A.block_offsets + r + 1) <= g0) ++r;
    const int64_t r_b0 = __ldg(A.block_offsets + r), r_b1 = __ldg(A.block_offsets + r + 1);
    const uint32_t* p0 = A.tokens + __ldg(A.offsets + r) + (g0 - r_b0) * (int64_t)A.block_size + 4 * part;
    if (g_last < w_end && g_last < r_b1 && (reinterpret_cast<uintptr_t>(p0) & 15) == 0) {
      // the whole trip lies inside one request: constant-stride loads, nothing to look up per block
      const size_t stride = (size_t)RPI * (size_t)A.block_size;
#pragma unroll
      for (int u = 0; u < kPartsUnroll; ++u) {
        const int4 v = ld_stream16(p0 + u * stride);
        lo[u] = (uint64_t)(uint32_t)v.x | ((uint64_t)(uint32_t)v.y << 32);
        hi[u] = (uint64_t)(uint32_t)v.z | ((uint64_t)(uint32_t)v.w << 32);
      }
    } else {
      int64_t rr = r;
#pragma unroll
      for (int u = 0; u < kPartsUnroll; ++u) {
        const int64_t gb = base + u * RPI + sub;
        lo[u] = hi[u] = 0;
        if (gb < w_end) {
          while (rr + 1 < A.n_req && __ldg(A.block_offsets + rr + 1) <= gb) ++rr;
          const uint32_t* src = A.tokens + __ldg(A.offsets + rr) +
                                (gb - __ldg(A.block_offsets + rr)) * (int64_t)A.block_size + 4 * part;
          if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
            const int4 v = ld_stream16(src);
            lo[u] = (uint64_t)(uint32_t)v.x | ((uint64_t)(uint32_t)v.y << 32);
            hi[u] = (uint64_t)(uint32_t)v.z | ((uint64_t)(uint32_t)v.w << 32);
          } else {  // the request does not start on a 16-byte boundary
            lo[u] = (uint64_t)__ldg(src) | ((uint64_t)__ldg(src + 1) << 32);
            hi[u] = (uint64_t)__ldg(src + 2) | ((uint64_t)__ldg(src + 3) << 32);
          }
        }
      }
      r = rr;
    }
#pragma unroll
    for (int u = 0; u < kPartsUnroll; ++u) {
      const int64_t gb = base + u * RPI + sub;
      uint64_t h;
      if (P == 1) {
        h = xxh3_16(lo[u], hi[u], S);
      } else {
        uint64_t acc = fold128(lo[u] ^ s0, hi[u] ^ s1);
#pragma unroll
        for (int m = P / 2; m >= 1; m >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, m);
        h = avalanche((uint64_t)LEN * P64_1 + acc);
      }
      if (part == 0 && gb < w_end) A.block_hash[gb] = h;
    }
  }
}

// ------------------------------------------------------------------------------------------------ generic block hash
__global__ void __launch_bounds__(256) kvdp_hash_blocks_generic(const __grid_constant__ HashArgs A,
                                                                const __grid_constant__ HashSecrets S) {
  const int64_t b_begin = __ldg(A.block_offsets), total = __ldg(A.block_offsets + A.n_req);  // see kvdp_hash_blocks_tma
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  for (int64_t gb = b_begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; gb < total; gb += nthreads) {
    const int64_t r = find_request(A.block_offsets, A.n_req, max(b_begin, gb & ~(int64_t)31), gb);
    WordReader in{A.tokens + __ldg(A.offsets + r) + (gb - __ldg(A.block_offsets + r)) * (int64_t)A.block_size};
    A.block_hash[gb] = xxh3_generic(in, (uint32_t)A.block_size * 4u, S);
  }
}

// ------------------------------------------------------------------------------------------------ chain
// One thread per request. The chain is serial by construction (each step hashes the previous step's result), so the
// kernel is latency-bound on the longest request; what can be hidden is the load of the block hashes: a register
// ring keeps the next kChainAhead values in flight while the current ones are chained. (A warp-per-request variant
// with shuffled inputs was measured 2x slower: 32 lanes repeating the same 64-bit multiplies made it issue-bound.)
constexpr int kChainThreads = 64;
constexpr int kChainAhead = 8;
__global__ void __launch_bounds__(kChainThreads) kvdp_hash_chain(const __grid_constant__ HashArgs A,
                                                                 const __grid_constant__ HashSecrets S) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= A.n_req) return;
  const int64_t b0 = __ldg(A.block_offsets + r), b1 = __ldg(A.block_offsets + r + 1);
  if (b0 >= b1) return;
  const uint64_t* __restrict__ bh = A.block_hash + b0;
  uint64_t* __restrict__ sq = A.seq_hash + b0;
  const uint32_t n = (uint32_t)(b1 - b0);  // blocks of one request always fit 32 bits
  // first block: its sequence hash is its block hash (twin variant: chained onto parent 0)
  uint64_t parent = bh[0];
  if (A.flags & KVDP_HASH_TWIN_FIRST) parent = xxh3_16(0, parent, S);
  sq[0] = parent;
  uint32_t i = 1;
  if (n >= 1 + 2 * kChainAhead) {
    uint64_t cur[kChainAhead], nxt[kChainAhead];
#pragma unroll
    for (int k = 0; k < kChainAhead; ++k) cur[k] = bh[i + k];
    // main loop: groups of kChainAhead with the next group's loads already in flight, no bounds predicates
    for (; i + 2 * kChainAhead <= n; i += kChainAhead) {
#pragma unroll
      for (int k = 0; k < kChainAhead; ++k) nxt[k] = bh[i + kChainAhead + k];
#pragma unroll
      for (int k = 0; k < kChainAhead; ++k) {
        parent = xxh3_16(parent, cur[k], S);
        sq[i + k] = parent;
      }
#pragma unroll
      for (int k = 0; k < kChainAhead; ++k) cur[k] = nxt[k];
    }
#pragma unroll
    for (int k = 0; k < kChainAhead; ++k) {  // the last prefetched group
      parent = xxh3_16(parent, cur[k], S);
      sq[i + k] = parent;
    }
    i += kChainAhead;
  }
  for (; i < n; ++i) {
    parent = xxh3_16(parent, bh[i], S);
    sq[i] = parent;
  }
}

// ------------------------------------------------------------------------------------------------ block-count scan
constexpr int kScanThreads = 256;
constexpr int kScanItems = 4;  // requests per thread: 1024-request tiles keep every CTA's loads short
constexpr int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ int64_t req_blocks(const int64_t* __restrict__ offsets, int64_t r, int64_t n_req,
                                              int32_t bs) {
  if (r >= n_req) return 0;
  const int64_t len = __ldg(offsets + r + 1) - __ldg(offsets + r);
  if (len <= 0) return 0;
  // request lengths almost always fit 32 bits: a 32-bit divide is ~5x cheaper than the 64-bit one
  if (len <= 0xffffffffLL) return (int64_t)((uint32_t)len / (uint32_t)bs);
  return len / bs;
}

__device__ __forceinline__ int64_t block_exclusive_scan(int64_t v, int64_t* total, int64_t* warp_sums) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int64_t incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int64_t o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int64_t w = lane < (kScanThreads / 32) ? warp_sums[lane] : 0;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int64_t o = __shfl_up_sync(0xffffffffu, w, d);
      if (lane >= d) w += o;
    }
    if (lane < (kScanThreads / 32)) warp_sums[lane] = w;  // inclusive over warps
  }
  __syncthreads();
  const int64_t warp_prefix = warp ? warp_sums[warp - 1] : 0;
  *total = warp_sums[kScanThreads / 32 - 1];
  return warp_prefix + incl - v;
}

__global__ void __launch_bounds__(kScanThreads) kvdp_block_count_scan_partials(const int64_t* __restrict__ offsets,
                                                                               int64_t n_req, int32_t bs,
                                                                               int64_t* __restrict__ partials) {
  __shared__ int64_t warp_sums[kScanThreads / 32];
  const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  int64_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) s += req_blocks(offsets, base + i, n_req, bs);
  int64_t total;
  block_exclusive_scan(s, &total, warp_sums);
  if (threadIdx.x == 0) partials[blockIdx.x] = total;
}

__global__ void __launch_bounds__(kScanThreads) kvdp_block_count_scan_spine(int64_t* __restrict__ partials,
                                                                            int64_t n_tiles) {
  __shared__ int64_t warp_sums[kScanThreads / 32];
  __shared__ int64_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int64_t base = 0; base < n_tiles; base += kScanThreads) {
    const int64_t i = base + threadIdx.x;
    const int64_t v = i < n_tiles ? partials[i] : 0;
    int64_t total;
    const int64_t ex = block_exclusive_scan(v, &total, warp_sums);
    const int64_t carry = carry_s;
    if (i < n_tiles) partials[i] = carry + ex;
    __syncthreads();
    if (threadIdx.x == 0) carry_s = carry + total;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kScanThreads) kvdp_block_count_scan_apply(const int64_t* __restrict__ offsets,
                                                                            int64_t n_req, int32_t bs,
                                                                            const int64_t* __restrict__ partials,
                                                                            int64_t* __restrict__ block_offsets) {
  __shared__ int64_t warp_sums[kScanThreads / 32];
  const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  int64_t v[kScanItems];
  int64_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    v[i] = req_blocks(offsets, base + i, n_req, bs);
    s += v[i];
  }
  int64_t total;
  int64_t run = partials[blockIdx.x] + block_exclusive_scan(s, &total, warp_sums);
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    if (base + i <= n_req) block_offsets[base + i] = run;  // entry n_req receives the grand total
    run += v[i];
  }
}

// Small batches (<= kScanSmallMax requests): one CTA scans everything in one pass, one launch instead of three.
// Each thread owns a run of consecutive requests: sum them, one block-wide exclusive scan of the per-thread sums,
// then a second walk over the same (now cached) offsets writes the per-request prefix.
constexpr int kScanSmallThreads = 1024;
constexpr int64_t kScanSmallMax = 2048;  // beyond this a single CTA is latency-bound (measured ~3 us per 1000 requests)
__global__ void __launch_bounds__(kScanSmallThreads) kvdp_block_count_scan_small(const int64_t* __restrict__ offsets,
                                                                                 int64_t n_req, int32_t bs,
                                                                                 int64_t* __restrict__ block_offsets) {
  __shared__ int64_t warp_sums[kScanSmallThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t per = (n_req + 1 + kScanSmallThreads - 1) / kScanSmallThreads;
  const int64_t r0 = (int64_t)threadIdx.x * per;
  int64_t s = 0;
  for (int64_t i = 0; i < per; ++i) s += req_blocks(offsets, r0 + i, n_req, bs);
  int64_t incl = s;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int64_t o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int64_t w = warp_sums[lane];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int64_t o = __shfl_up_sync(0xffffffffu, w, d);
      if (lane >= d) w += o;
    }
    warp_sums[lane] = w;
  }
  __syncthreads();
  int64_t run = (warp ? warp_sums[warp - 1] : 0) + incl - s;
  for (int64_t i = 0; i < per; ++i) {
    const int64_t r = r0 + i;
    if (r <= n_req) block_offsets[r] = run;
    run += req_blocks(offsets, r, n_req, bs);
  }
}

__global__ void kvdp_xxh3_batch_kernel(const uint8_t* __restrict__ data, int64_t n, int64_t len, uint64_t* out,
                                       const __grid_constant__ HashSecrets S) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  ByteReader in{data + i * len};
  out[i] = xxh3_generic(in, (uint32_t)len, S);
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// ------------------------------------------------------------------------------------------------ host-side launches
// Block-count scan over the whole batch (needs only the token offsets).
int launch_scan(const int64_t* d_offsets, int64_t n_req, int32_t block_size, int64_t* d_block_offsets,
                void* d_workspace, cudaStream_t stream) {
  int64_t* partials = (int64_t*)d_workspace;
  const int64_t tiles = (n_req + 1 + kScanTile - 1) / kScanTile;
  if (n_req + 1 <= kScanSmallMax) {
    kvdp_block_count_scan_small<<<1, kScanSmallThreads, 0, stream>>>(d_offsets, n_req, block_size, d_block_offsets);
  } else {
    kvdp_block_count_scan_partials<<<(unsigned)tiles, kScanThreads, 0, stream>>>(d_offsets, n_req, block_size,
                                                                                 partials);
    kvdp_block_count_scan_spine<<<1, kScanThreads, 0, stream>>>(partials, tiles);
    kvdp_block_count_scan_apply<<<(unsigned)tiles, kScanThreads, 0, stream>>>(d_offsets, n_req, block_size, partials,
                                                                              d_block_offsets);
  }
  KVDP_CUDA(cudaGetLastError());
  return 0;
}

template <int NSTRIPES, int W>
int launch_tma(const HashArgs& A, const HashSecrets& S, int64_t warp_tiles, const DevInfo& di, int dev,
               cudaStream_t stream) {
  const size_t smem = (size_t)W * 32 * (64 * NSTRIPES + 16) + 8 * W;
  static std::mutex mu;
  static bool attr_set[64] = {};  // the attribute is per device
  {
    std::lock_guard<std::mutex> lk(mu);
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
      KVDP_CUDA(cudaFuncSetAttribute(kvdp_hash_blocks_tma<NSTRIPES, W>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)smem));
      if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
  }
  const int64_t grid = std::min<int64_t>((warp_tiles + W - 1) / W, (int64_t)di.sm_count * 3);
  kvdp_hash_blocks_tma<NSTRIPES, W><<<(unsigned)std::max<int64_t>(grid, 1), W * 32, smem, stream>>>(A, S);
  return 0;
}

// Block hashes of the requests A describes (A may be a sub-range of the batch: shifted offset arrays, smaller n_req).
// max_blocks = host-side upper bound of the number of blocks in that range; it only sizes the grid.
int launch_block_hash(const HashArgs& A, const HashSecrets& S, int64_t max_blocks, const DevInfo& di, int dev,
                      cudaStream_t stream) {
  const int len = A.block_size * 4;
  const int64_t warp_tiles = (max_blocks + 31) / 32;
  if (len == 256) {
    if (int rc = launch_tma<4, 8>(A, S, warp_tiles, di, dev, stream)) return rc;
  } else if (len == 512) {
    if (int rc = launch_tma<8, 4>(A, S, warp_tiles, di, dev, stream)) return rc;
  } else if (len == 1024) {
    if (int rc = launch_tma<16, 2>(A, S, warp_tiles, di, dev, stream)) return rc;
  } else if (len == 16 || len == 32 || len == 64 || len == 128) {
    const int64_t per_cta = 8 * (512 / len) * kPartsUnroll;  // blocks one CTA covers per loop trip
    const int64_t grid =
        std::max<int64_t>(1, std::min<int64_t>((max_blocks + per_cta - 1) / per_cta, (int64_t)di.sm_count * 8));
    switch (len) {
      case 16: kvdp_hash_blocks_parts<16><<<(unsigned)grid, 256, 0, stream>>>(A, S); break;
      case 32: kvdp_hash_blocks_parts<32><<<(unsigned)grid, 256, 0, stream>>>(A, S); break;
      case 64: kvdp_hash_blocks_parts<64><<<(unsigned)grid, 256, 0, stream>>>(A, S); break;
      default: kvdp_hash_blocks_parts<128><<<(unsigned)grid, 256, 0, stream>>>(A, S); break;
    }
  } else {
    const int64_t grid = std::min<int64_t>((max_blocks + 255) / 256, (int64_t)di.sm_count * 8);
    kvdp_hash_blocks_generic<<<(unsigned)std::max<int64_t>(grid, 1), 256, 0, stream>>>(A, S);
  }
  KVDP_CUDA(cudaGetLastError());
  return 0;
}

int launch_chain(const HashArgs& A, const HashSecrets& S, cudaStream_t stream) {
  if (A.n_req <= 0) return 0;
  kvdp_hash_chain<<<(unsigned)((A.n_req + kChainThreads - 1) / kChainThreads), kChainThreads, 0, stream>>>(A, S);
  KVDP_CUDA(cudaGetLastError());
  return 0;
}

HashArgs sub_range(const HashArgs& A, int64_t r0, int64_t r1) {
  HashArgs B = A;
  B.offsets = A.offsets + r0;
  B.block_offsets = A.block_offsets + r0;
  B.n_req = r1 - r0;
  return B;
}

// Side streams for the staged form of a chained call: the chain kernel of one request range is latency-bound (one
// thread per request, serial per request) and runs under the HBM-bound block-hash kernel of the next range.
constexpr int kMaxHashStages = 16;
struct HashFork {
  std::mutex mu;
  cudaStream_t side[2] = {nullptr, nullptr};
  cudaEvent_t hashed[kMaxHashStages] = {};
  cudaEvent_t chained[2] = {nullptr, nullptr};
  bool ok = false;
};
HashFork g_fork[64];

int fork_init(HashFork& f) {
  if (f.ok) return 0;
  int lo = 0, hi = 0;
  KVDP_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));  // hi = numerically lowest = greatest priority
  for (int i = 0; i < 2; ++i) {
    KVDP_CUDA(cudaStreamCreateWithPriority(&f.side[i], cudaStreamNonBlocking, hi));
    KVDP_CUDA(cudaEventCreateWithFlags(&f.chained[i], cudaEventDisableTiming));
  }
  for (int i = 0; i < kMaxHashStages; ++i) KVDP_CUDA(cudaEventCreateWithFlags(&f.hashed[i], cudaEventDisableTiming));
  f.ok = true;
  return 0;
}

// The body of kvdp_hash_token_blocks. `may_fork` = the call may use the library's side streams (created on the first
// such call per device); the allocation-free entry point passes false and stays on the caller's stream alone.
int hash_token_blocks_impl(const uint32_t* d_tokens, const int64_t* d_offsets, int64_t n_req, int64_t total_tokens,
                           int32_t block_size, uint64_t salt, int64_t* d_block_offsets, uint64_t* d_block_hash,
                           uint64_t* d_seq_hash, void* d_workspace, size_t workspace_bytes, uint32_t flags,
                           cudaStream_t stream, bool may_fork) {
  if (n_req < 0 || total_tokens < 0) return fail(KVDP_EINVAL, "hash: negative n_req/total_tokens");
  if (block_size < 1) return fail(KVDP_EINVAL, "hash: block_size must be >= 1");
  if (!d_offsets || !d_block_offsets) return fail(KVDP_EINVAL, "hash: NULL offsets / block_offsets");
  if (total_tokens > 0 && !d_tokens) return fail(KVDP_EINVAL, "hash: NULL tokens");
  if (flags & ~KVDP_HASH_TWIN_FIRST) return fail(KVDP_EINVAL, "hash: unknown flags 0x%x", flags);
  const size_t need = kvdp_hash_workspace_bytes(n_req);
  if (!d_workspace || workspace_bytes < need)
    return fail(KVDP_EINVAL, "hash: workspace of %zu bytes needed, %zu given", need, workspace_bytes);
  if ((uint64_t)block_size * 4 >= ((uint64_t)1 << 31)) return fail(KVDP_ERANGE, "hash: block_size too large");

  int dev = 0;
  KVDP_CUDA(cudaGetDevice(&dev));
  DevInfo di;
  if (int rc = device_info(dev, &di)) return rc;

  // 1. block offsets
  if (int rc = launch_scan(d_offsets, n_req, block_size, d_block_offsets, d_workspace, stream)) return rc;

  const int64_t max_blocks = total_tokens / block_size;
  if (max_blocks == 0 || n_req == 0) return 0;
  if (!d_block_hash) return fail(KVDP_EINVAL, "hash: NULL block_hash");

  HashArgs A;
  A.tokens = d_tokens;
  A.offsets = d_offsets;
  A.block_offsets = d_block_offsets;
  A.block_hash = d_block_hash;
  A.seq_hash = d_seq_hash;
  A.n_req = n_req;
  A.block_size = block_size;
  A.flags = flags;
  const HashSecrets S = make_secrets(salt);

  // 2 + 3. block hashes, then the chain. A big chained batch is cut into request ranges: the chain of range k runs
  // on a side stream under the block hashing of range k+1 and is joined back onto the caller's stream at the end, so
  // the caller sees one ordered call. Only the last range's chain is not hidden.
  int stages = (int)options().hash_stages;
  if (stages <= 0) stages = 1;  // measured (profiles/sweep_r02_hash_stages.txt): every extra range costs ~16 us of
                                // launch ramp and tail and its chain still pays the serial floor; off by default
  stages = std::min(stages, kMaxHashStages);
  const bool staged = may_fork && d_seq_hash && stages > 1 && dev >= 0 && dev < 64 && n_req >= 4096 &&
                      (int64_t)max_blocks * block_size * 4 >= ((int64_t)64 << 20);
  if (!staged) {
    if (int rc = launch_block_hash(A, S, max_blocks, di, dev, stream)) return rc;
    if (d_seq_hash)
      if (int rc = launch_chain(A, S, stream)) return rc;
    return 0;
  }
  HashFork& f = g_fork[dev];
  std::lock_guard<std::mutex> lk(f.mu);  // a fork/join sequence is enqueued as a whole
  if (int rc = fork_init(f)) return rc;
  for (int k = 0; k < stages; ++k) {
    const int64_t r0 = n_req * k / stages, r1 = n_req * (k + 1) / stages;
    if (r1 <= r0) continue;
    const HashArgs B = sub_range(A, r0, r1);
    // the host does not see the offsets: size each range's grid for its share of the batch, with headroom
    const int64_t range_blocks = std::min<int64_t>(max_blocks, max_blocks / stages * 2 + 1024);
    if (int rc = launch_block_hash(B, S, range_blocks, di, dev, stream)) return rc;
    KVDP_CUDA(cudaEventRecord(f.hashed[k], stream));
    cudaStream_t side = f.side[k & 1];
    KVDP_CUDA(cudaStreamWaitEvent(side, f.hashed[k], 0));
    if (int rc = launch_chain(B, S, side)) return rc;
  }
  for (int i = 0; i < 2; ++i) {
    KVDP_CUDA(cudaEventRecord(f.chained[i], f.side[i]));
    KVDP_CUDA(cudaStreamWaitEvent(stream, f.chained[i], 0));
  }
  return 0;
}

}  // namespace

// used by hash_static.cu: the same kernels on the caller's stream only (no side streams, nothing created)
int hash_token_blocks_single_stream(const uint32_t* d_tokens, const int64_t* d_offsets, int64_t n_req,
                                    int64_t total_tokens, int32_t block_size, uint64_t salt, int64_t* d_block_offsets,
                                    uint64_t* d_block_hash, uint64_t* d_seq_hash, void* d_workspace,
                                    size_t workspace_bytes, uint32_t flags, cudaStream_t stream) {
  return hash_token_blocks_impl(d_tokens, d_offsets, n_req, total_tokens, block_size, salt, d_block_offsets,
                                d_block_hash, d_seq_hash, d_workspace, workspace_bytes, flags, stream, false);
}
}  // namespace kvdp

using namespace kvdp;

extern "C" {

size_t kvdp_hash_workspace_bytes(int64_t n_req) {
  if (n_req < 0) n_req = 0;
  const int64_t tiles = (n_req + 1 + kScanTile - 1) / kScanTile;
  return align_up((size_t)(tiles + 1) * sizeof(int64_t), 256);
}

int kvdp_hash_token_blocks(const uint32_t* d_tokens, const int64_t* d_offsets, int64_t n_req, int64_t total_tokens,
                           int32_t block_size, uint64_t salt, int64_t* d_block_offsets, uint64_t* d_block_hash,
                           uint64_t* d_seq_hash, void* d_workspace, size_t workspace_bytes, uint32_t flags,
                           kvdp_stream_t stream_) {
  return hash_token_blocks_impl(d_tokens, d_offsets, n_req, total_tokens, block_size, salt, d_block_offsets,
                                d_block_hash, d_seq_hash, d_workspace, workspace_bytes, flags, (cudaStream_t)stream_,
                                true);
}

int kvdp_xxh3_64_batch(const uint8_t* d_data, int64_t n, int64_t len, uint64_t seed, uint64_t* d_out,
                       kvdp_stream_t stream) {
  if (n < 0 || len < 0 || len >= ((int64_t)1 << 31)) return fail(KVDP_EINVAL, "xxh3 batch: bad n/len");
  if (n == 0) return 0;
  if (!d_out || (len > 0 && !d_data)) return fail(KVDP_EINVAL, "xxh3 batch: NULL pointer");
  const HashSecrets S = make_secrets(seed);
  kvdp_xxh3_batch_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(d_data, n, len, d_out, S);
  KVDP_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ host-buffer form
namespace {
struct HostHashCtx {
  std::mutex mu;
  cudaStream_t up = nullptr, compute = nullptr, down = nullptr;
  cudaEvent_t uploaded = nullptr, hashed = nullptr, scanned = nullptr;
  void* d_tokens = nullptr;
  size_t tok_cap = 0;
  void* d_req = nullptr;  // offsets + block_offsets
  size_t req_cap = 0;
  void* d_out = nullptr;  // block_hash + seq_hash
  size_t out_cap = 0;
  void* d_ws = nullptr;
  size_t ws_cap = 0;
};
HostHashCtx g_host_ctx[64];

int ensure(void** p, size_t* cap, size_t need) {
  if (*cap >= need) return 0;
  if (*p) KVDP_CUDA(cudaFree(*p));
  *p = nullptr;
  *cap = 0;
  const size_t sz = std::max(need, (size_t)1 << 20);
  KVDP_CUDA(cudaMalloc(p, sz));
  *cap = sz;
  return 0;
}

constexpr int64_t kHostChunkTokens = (int64_t)4 << 20;  // 16 MiB of tokens per pipeline step: with 4 MiB pieces the gaps
                                                        // between consecutive DMAs cost ~10 % of the link
}  // namespace

// Upload, hashing and download overlap: the batch is cut at request boundaries into pieces of about 16 MiB of tokens;
// piece k is hashed (and chained) on the compute stream while piece k+1 is still on its way up and the results of piece
// k-1 are on their way down, each on a stream of its own. With pinned host buffers the call runs at the speed of the
// host->device link; pageable buffers work too (the driver stages them).
int kvdp_hash_token_blocks_host(const uint32_t* h_tokens, const int64_t* h_offsets, int64_t n_req, int32_t block_size,
                                uint64_t salt, int64_t* h_block_offsets, uint64_t* h_block_hash, uint64_t* h_seq_hash,
                                uint32_t flags, int32_t device) {
  if (n_req < 0 || !h_offsets || !h_block_offsets) return fail(KVDP_EINVAL, "hash(host): bad arguments");
  if (block_size < 1) return fail(KVDP_EINVAL, "hash(host): block_size must be >= 1");
  if ((uint64_t)block_size * 4 >= ((uint64_t)1 << 31)) return fail(KVDP_ERANGE, "hash: block_size too large");
  if (flags & ~KVDP_HASH_TWIN_FIRST) return fail(KVDP_EINVAL, "hash: unknown flags 0x%x", flags);
  if (device < 0) KVDP_CUDA(cudaGetDevice(&device));
  if (device >= 64) return fail(KVDP_EINVAL, "device ordinal %d not supported", device);
  // validate offsets on the host (the device path trusts them) and cut the batch into pieces
  struct Piece {
    int64_t r0, r1, b0, b1;
  };
  std::vector<Piece> pieces;
  int64_t total_blocks = 0, piece_r0 = 0, piece_b0 = 0;
  for (int64_t r = 0; r < n_req; ++r) {
    if (h_offsets[r + 1] < h_offsets[r]) return fail(KVDP_EINVAL, "hash(host): offsets must be non-decreasing");
    total_blocks += (h_offsets[r + 1] - h_offsets[r]) / block_size;
    if (h_offsets[r + 1] - h_offsets[piece_r0] >= kHostChunkTokens || r + 1 == n_req) {
      pieces.push_back({piece_r0, r + 1, piece_b0, total_blocks});
      piece_r0 = r + 1;
      piece_b0 = total_blocks;
    }
  }
  const int64_t tok_begin = h_offsets[0], tok_end = h_offsets[n_req];
  const int64_t total_tokens = tok_end;  // offsets index into h_tokens from 0
  if (tok_begin < 0) return fail(KVDP_EINVAL, "hash(host): negative offset");
  if (total_tokens > 0 && !h_tokens) return fail(KVDP_EINVAL, "hash(host): NULL tokens");
  if (total_blocks > 0 && !h_block_hash) return fail(KVDP_EINVAL, "hash(host): NULL block_hash");

  DeviceGuard g(device);
  DevInfo di;
  if (int rc = device_info(device, &di)) return rc;
  HostHashCtx& c = g_host_ctx[device];
  std::lock_guard<std::mutex> lk(c.mu);
  if (!c.up) {
    KVDP_CUDA(cudaStreamCreateWithFlags(&c.up, cudaStreamNonBlocking));
    KVDP_CUDA(cudaStreamCreateWithFlags(&c.compute, cudaStreamNonBlocking));
    KVDP_CUDA(cudaStreamCreateWithFlags(&c.down, cudaStreamNonBlocking));
    KVDP_CUDA(cudaEventCreateWithFlags(&c.uploaded, cudaEventDisableTiming));
    KVDP_CUDA(cudaEventCreateWithFlags(&c.hashed, cudaEventDisableTiming));
    KVDP_CUDA(cudaEventCreateWithFlags(&c.scanned, cudaEventDisableTiming));
  }
  if (int rc = ensure(&c.d_tokens, &c.tok_cap, (size_t)total_tokens * 4 + 16)) return rc;
  if (int rc = ensure(&c.d_req, &c.req_cap, (size_t)(n_req + 1) * 16)) return rc;
  if (int rc = ensure(&c.d_out, &c.out_cap, (size_t)total_blocks * 16 + 16)) return rc;
  const size_t ws = kvdp_hash_workspace_bytes(n_req);
  if (int rc = ensure(&c.d_ws, &c.ws_cap, ws)) return rc;
  int64_t* d_off = (int64_t*)c.d_req;
  int64_t* d_boff = d_off + (n_req + 1);
  uint64_t* d_bh = (uint64_t*)c.d_out;
  uint64_t* d_sh = d_bh + total_blocks;
  const uint32_t* d_tok = (const uint32_t*)c.d_tokens;

  // offsets first (small), then the block-count scan; its result goes down while the token pieces go up
  KVDP_CUDA(cudaMemcpyAsync(d_off, h_offsets, (size_t)(n_req + 1) * 8, cudaMemcpyHostToDevice, c.compute));
  if (int rc = launch_scan(d_off, n_req, block_size, d_boff, c.d_ws, c.compute)) return rc;
  KVDP_CUDA(cudaEventRecord(c.scanned, c.compute));
  KVDP_CUDA(cudaStreamWaitEvent(c.down, c.scanned, 0));
  KVDP_CUDA(cudaMemcpyAsync(h_block_offsets, d_boff, (size_t)(n_req + 1) * 8, cudaMemcpyDeviceToHost, c.down));

  HashArgs A;
  A.tokens = d_tok;
  A.offsets = d_off;
  A.block_offsets = d_boff;
  A.block_hash = d_bh;
  A.seq_hash = h_seq_hash ? d_sh : nullptr;
  A.n_req = n_req;
  A.block_size = block_size;
  A.flags = flags;
  const HashSecrets S = make_secrets(salt);
  for (const Piece& p : pieces) {
    const int64_t t0 = h_offsets[p.r0], t1 = h_offsets[p.r1];
    if (t1 > t0)
      KVDP_CUDA(cudaMemcpyAsync((void*)(d_tok + t0), h_tokens + t0, (size_t)(t1 - t0) * 4, cudaMemcpyHostToDevice, c.up));
    if (p.b1 == p.b0) continue;
    KVDP_CUDA(cudaEventRecord(c.uploaded, c.up));
    KVDP_CUDA(cudaStreamWaitEvent(c.compute, c.uploaded, 0));
    const HashArgs B = sub_range(A, p.r0, p.r1);
    if (int rc = launch_block_hash(B, S, p.b1 - p.b0, di, device, c.compute)) return rc;
    if (h_seq_hash)
      if (int rc = launch_chain(B, S, c.compute)) return rc;
    KVDP_CUDA(cudaEventRecord(c.hashed, c.compute));
    KVDP_CUDA(cudaStreamWaitEvent(c.down, c.hashed, 0));
    KVDP_CUDA(cudaMemcpyAsync(h_block_hash + p.b0, d_bh + p.b0, (size_t)(p.b1 - p.b0) * 8, cudaMemcpyDeviceToHost,
                              c.down));
    if (h_seq_hash)
      KVDP_CUDA(cudaMemcpyAsync(h_seq_hash + p.b0, d_sh + p.b0, (size_t)(p.b1 - p.b0) * 8, cudaMemcpyDeviceToHost,
                                c.down));
  }
  // the download stream is last in every dependency chain (uploads -> kernels -> downloads); a batch without any
  // full block still has to see its uploads and the scan through
  if (total_blocks == 0) KVDP_CUDA(cudaStreamSynchronize(c.compute));
  KVDP_CUDA(cudaStreamSynchronize(c.up));  // trailing requests shorter than a block are uploaded but wait for nobody
  KVDP_CUDA(cudaStreamSynchronize(c.down));
  return 0;
}

}  // extern "C"
