/*
 * oracle/xxh3_oracle.c — TEST INFRASTRUCTURE ONLY (CPU restatement, never shipped, never on the product path).
 *
 * What it restates
 *   - XXH3-64-with-seed, the arithmetic behind the reference's `compute_hash`
 *       /root/reference/lib/tokens/src/lib.rs:44-46          (xxh3::xxh3_64_with_seed(data, seed))
 *       /root/reference/lib/llm/src/kv_router/indexer.rs:87-89 (same, seed = XXH3_SEED = 1337, :64)
 *     The arithmetic itself lives in the un-vendored crate xxhash-rust 0.8.15
 *     (/root/reference/Cargo.lock:7618-7621); it is the frozen public XXH3 algorithm (XXH >= 0.8.0),
 *     restated here from the published specification.
 *   - lib/tokens block/sequence hashing semantics
 *       block_hash   : lib/tokens/src/lib.rs:197-215   (hash of the LE bytes of one full block, seed = salt)
 *       sequence_hash: lib/tokens/src/lib.rs:238-261   (first block: = block_hash; else hash of
 *                                                        LE64(parent_seq) || LE64(block_hash), seed = salt)
 *       split_tokens : lib/tokens/src/lib.rs:369-404   (chunks_exact; remainder is not hashed)
 *   - the router indexer's un-chained per-block hash
 *       compute_block_hash_for_seq: lib/llm/src/kv_router/indexer.rs:123-135
 *   - the in-crate twin's first-block push_token quirk (lib/llm/src/tokens.rs:126-144)
 *
 * Parity pin: the three reference known-answer values (lib/tokens/src/lib.rs:539-545) plus cross-checks
 * against two independent XXH3 implementations available in the image (python `xxhash`, pyarrow's vendored
 * xxhash.h) — see tests/test_oracle_hash.py.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include "par.h"

typedef uint64_t u64;
typedef uint32_t u32;
typedef uint8_t u8;

static const u64 P32_1 = 0x9E3779B1ULL, P32_2 = 0x85EBCA77ULL, P32_3 = 0xC2B2AE3DULL;
static const u64 P64_1 = 0x9E3779B185EBCA87ULL, P64_2 = 0xC2B2AE3D27D4EB4FULL, P64_3 = 0x165667B19E3779F9ULL,
                 P64_4 = 0x85EBCA77C2B2AE63ULL, P64_5 = 0x27D4EB2F165667C5ULL;
static const u64 MX1 = 0x165667919E3779F9ULL, MX2 = 0x9FB21C651E98DF25ULL;

/* the 192-byte default secret of the public algorithm */
static const u8 kSecret[192] = {
    0xb8, 0xfe, 0x6c, 0x39, 0x23, 0xa4, 0x4b, 0xbe, 0x7c, 0x01, 0x81, 0x2c, 0xf7, 0x21, 0xad, 0x1c,
    0xde, 0xd4, 0x6d, 0xe9, 0x83, 0x90, 0x97, 0xdb, 0x72, 0x40, 0xa4, 0xa4, 0xb7, 0xb3, 0x67, 0x1f,
    0xcb, 0x79, 0xe6, 0x4e, 0xcc, 0xc0, 0xe5, 0x78, 0x82, 0x5a, 0xd0, 0x7d, 0xcc, 0xff, 0x72, 0x21,
    0xb8, 0x08, 0x46, 0x74, 0xf7, 0x43, 0x24, 0x8e, 0xe0, 0x35, 0x90, 0xe6, 0x81, 0x3a, 0x26, 0x4c,
    0x3c, 0x28, 0x52, 0xbb, 0x91, 0xc3, 0x00, 0xcb, 0x88, 0xd0, 0x65, 0x8b, 0x1b, 0x53, 0x2e, 0xa3,
    0x71, 0x64, 0x48, 0x97, 0xa2, 0x0d, 0xf9, 0x4e, 0x38, 0x19, 0xef, 0x46, 0xa9, 0xde, 0xac, 0xd8,
    0xa8, 0xfa, 0x76, 0x3f, 0xe3, 0x9c, 0x34, 0x3f, 0xf9, 0xdc, 0xbb, 0xc7, 0xc7, 0x0b, 0x4f, 0x1d,
    0x8a, 0x51, 0xe0, 0x4b, 0xcd, 0xb4, 0x59, 0x31, 0xc8, 0x9f, 0x7e, 0xc9, 0xd9, 0x78, 0x73, 0x64,
    0xea, 0xc5, 0xac, 0x83, 0x34, 0xd3, 0xeb, 0xc3, 0xc5, 0x81, 0xa0, 0xff, 0xfa, 0x13, 0x63, 0xeb,
    0x17, 0x0d, 0xdd, 0x51, 0xb7, 0xf0, 0xda, 0x49, 0xd3, 0x16, 0x55, 0x26, 0x29, 0xd4, 0x68, 0x9e,
    0x2b, 0x16, 0xbe, 0x58, 0x7d, 0x47, 0xa1, 0xfc, 0x8f, 0xf8, 0xb8, 0xd1, 0x7a, 0xd0, 0x31, 0xce,
    0x45, 0xcb, 0x3a, 0x8f, 0x95, 0x16, 0x04, 0x28, 0xaf, 0xd7, 0xfb, 0xca, 0xbb, 0x4b, 0x40, 0x7e,
};

static inline u64 rd64(const u8* p) {
  u64 v = 0;
  for (int i = 7; i >= 0; --i) v = (v << 8) | p[i];
  return v;
}
static inline u32 rd32(const u8* p) {
  return (u32)p[0] | ((u32)p[1] << 8) | ((u32)p[2] << 16) | ((u32)p[3] << 24);
}
static inline void wr64(u8* p, u64 v) {
  for (int i = 0; i < 8; ++i) p[i] = (u8)(v >> (8 * i));
}
static inline u64 rotl64(u64 x, int r) { return (x << r) | (x >> (64 - r)); }
static inline u32 swap32(u32 x) {
  return ((x << 24) & 0xff000000u) | ((x << 8) & 0x00ff0000u) | ((x >> 8) & 0x0000ff00u) | ((x >> 24) & 0xffu);
}
static inline u64 swap64(u64 x) { return ((u64)swap32((u32)x) << 32) | swap32((u32)(x >> 32)); }

static inline u64 mul128_fold64(u64 a, u64 b) {
  unsigned __int128 p = (unsigned __int128)a * b;
  return (u64)p ^ (u64)(p >> 64);
}
static inline u64 avalanche(u64 h) {
  h ^= h >> 37;
  h *= MX1;
  h ^= h >> 32;
  return h;
}
static inline u64 rrmxmx(u64 h, u64 len) {
  h ^= rotl64(h, 49) ^ rotl64(h, 24);
  h *= MX2;
  h ^= (h >> 35) + len;
  h *= MX2;
  h ^= h >> 28;
  return h;
}
static inline u64 xxh64_avalanche(u64 h) {
  h ^= h >> 33;
  h *= P64_2;
  h ^= h >> 29;
  h *= P64_3;
  h ^= h >> 32;
  return h;
}
static inline u64 mix16(const u8* in, const u8* sec, u64 seed) {
  return mul128_fold64(rd64(in) ^ (rd64(sec) + seed), rd64(in + 8) ^ (rd64(sec + 8) - seed));
}

static void accumulate_stripe(u64 acc[8], const u8* p, const u8* sec) {
  for (int l = 0; l < 8; ++l) {
    u64 v = rd64(p + 8 * l);
    u64 k = v ^ rd64(sec + 8 * l);
    acc[l ^ 1] += v;
    acc[l] += (k & 0xFFFFFFFFULL) * (k >> 32);
  }
}
static void scramble(u64 acc[8], const u8* sec) {
  for (int l = 0; l < 8; ++l) {
    u64 a = acc[l];
    a ^= a >> 47;
    a ^= rd64(sec + 8 * l);
    a *= P32_1;
    acc[l] = a;
  }
}

static u64 hash_long(const u8* in, size_t len, u64 seed) {
  u8 S[192];
  if (seed == 0) {
    memcpy(S, kSecret, 192);
  } else {
    for (int i = 0; i < 12; ++i) {
      wr64(S + 16 * i, rd64(kSecret + 16 * i) + seed);
      wr64(S + 16 * i + 8, rd64(kSecret + 16 * i + 8) - seed);
    }
  }
  u64 acc[8] = {P32_3, P64_1, P64_2, P64_3, P64_4, P32_2, P64_5, P32_1};
  const size_t stripes_per_block = (192 - 64) / 8; /* 16 */
  const size_t block_len = 64 * stripes_per_block; /* 1024 */
  const size_t nb_blocks = (len - 1) / block_len;
  for (size_t b = 0; b < nb_blocks; ++b) {
    for (size_t n = 0; n < stripes_per_block; ++n) accumulate_stripe(acc, in + b * block_len + 64 * n, S + 8 * n);
    scramble(acc, S + 192 - 64);
  }
  const size_t nb_stripes = ((len - 1) - block_len * nb_blocks) / 64;
  for (size_t n = 0; n < nb_stripes; ++n) accumulate_stripe(acc, in + nb_blocks * block_len + 64 * n, S + 8 * n);
  accumulate_stripe(acc, in + len - 64, S + 192 - 64 - 7);
  u64 r = (u64)len * P64_1;
  for (int i = 0; i < 4; ++i)
    r += mul128_fold64(acc[2 * i] ^ rd64(S + 11 + 16 * i), acc[2 * i + 1] ^ rd64(S + 11 + 16 * i + 8));
  return avalanche(r);
}

/* XXH3-64 with seed over an arbitrary byte string. */
u64 kvo_xxh3_64(const void* data, size_t len, u64 seed) {
  const u8* in = (const u8*)data;
  const u8* s = kSecret;
  if (len == 0) return xxh64_avalanche(seed ^ rd64(s + 56) ^ rd64(s + 64));
  if (len <= 3) {
    u32 c1 = in[0], c2 = in[len >> 1], c3 = in[len - 1];
    u32 comb = (c1 << 16) | (c2 << 24) | c3 | ((u32)len << 8);
    u64 flip = (u64)(rd32(s) ^ rd32(s + 4)) + seed;
    return xxh64_avalanche((u64)comb ^ flip);
  }
  if (len <= 8) {
    seed ^= (u64)swap32((u32)seed) << 32;
    u64 in64 = (u64)rd32(in + len - 4) + ((u64)rd32(in) << 32);
    u64 flip = (rd64(s + 8) ^ rd64(s + 16)) - seed;
    return rrmxmx(in64 ^ flip, (u64)len);
  }
  if (len <= 16) {
    u64 lo = rd64(in) ^ ((rd64(s + 24) ^ rd64(s + 32)) + seed);
    u64 hi = rd64(in + len - 8) ^ ((rd64(s + 40) ^ rd64(s + 48)) - seed);
    return avalanche((u64)len + swap64(lo) + hi + mul128_fold64(lo, hi));
  }
  if (len <= 128) {
    u64 acc = (u64)len * P64_1;
    for (int i = (int)((len - 1) / 32); i >= 0; --i) {
      acc += mix16(in + 16 * i, s + 32 * i, seed);
      acc += mix16(in + len - 16 * (i + 1), s + 32 * i + 16, seed);
    }
    return avalanche(acc);
  }
  if (len <= 240) {
    u64 acc = (u64)len * P64_1;
    for (int i = 0; i < 8; ++i) acc += mix16(in + 16 * i, s + 16 * i, seed);
    acc = avalanche(acc);
    u64 end = mix16(in + len - 16, s + 136 - 17, seed);
    int rounds = (int)(len / 16);
    for (int i = 8; i < rounds; ++i) end += mix16(in + 16 * i, s + 16 * (i - 8) + 3, seed);
    return avalanche(acc + end);
  }
  return hash_long(in, len, seed);
}

/* Hash the LE bytes of `n` u32 tokens (lib/tokens/src/lib.rs:197-215; indexer.rs:123-135 builds the same bytes). */
static u64 hash_tokens(const u32* tok, size_t n, u64 seed) {
  /* explicit little-endian serialisation so the oracle does not depend on host endianness */
  u8 buf[4096];
  if (n * 4 <= sizeof buf) {
    for (size_t i = 0; i < n; ++i) {
      buf[4 * i] = (u8)tok[i];
      buf[4 * i + 1] = (u8)(tok[i] >> 8);
      buf[4 * i + 2] = (u8)(tok[i] >> 16);
      buf[4 * i + 3] = (u8)(tok[i] >> 24);
    }
    return kvo_xxh3_64(buf, n * 4, seed);
  }
  return kvo_xxh3_64(tok, n * 4, seed); /* LE hosts only for very large blocks */
}

static u64 chain_step(u64 parent, u64 block_hash, u64 seed) {
  u8 b[16];
  wr64(b, parent);
  wr64(b + 8, block_hash);
  return kvo_xxh3_64(b, 16, seed);
}

/*
 * TokenBlockSequence::split_tokens for one request (lib/tokens/src/lib.rs:369-404).
 * Writes floor(n/bs) (block_hash, sequence_hash) pairs; returns that count.
 * twin_first_block != 0 reproduces lib/llm/src/tokens.rs:126-144 (push_token path hashes [0, block_hash] for block 0).
 */
size_t kvo_split_tokens(const u32* tok, size_t n, size_t bs, u64 salt, u64* block_hash, u64* seq_hash,
                        int twin_first_block) {
  if (bs == 0) return 0;
  size_t nb = n / bs;
  u64 parent = 0;
  for (size_t b = 0; b < nb; ++b) {
    u64 bh = hash_tokens(tok + b * bs, bs, salt);
    u64 sh;
    if (b == 0)
      sh = twin_first_block ? chain_step(0, bh, salt) : bh;
    else
      sh = chain_step(parent, bh, salt);
    block_hash[b] = bh;
    seq_hash[b] = sh;
    parent = sh;
  }
  return nb;
}

/* Batched (CSR) form over many requests; offsets has n_req+1 entries (token units).
 * out_off[r] = first output slot of request r (prefix sum of floor(len_r/bs)); out_off has n_req+1 entries.
 * threads <= 1: serial loop (mirrors the reference's serial host loop); else pthreads over requests
 * (the reference uses rayon over blocks: lib/tokens/src/lib.rs:374-378). */
struct hash_batch_ctx {
  const u32* tok;
  const int64_t* offsets;
  const int64_t* out_off;
  size_t bs;
  u64 salt;
  u64* block_hash;
  u64* seq_hash;
};
static void hash_batch_body(long b, long e, void* c_) {
  struct hash_batch_ctx* c = (struct hash_batch_ctx*)c_;
  for (long r = b; r < e; ++r)
    kvo_split_tokens(c->tok + c->offsets[r], (size_t)(c->offsets[r + 1] - c->offsets[r]), c->bs, c->salt,
                     c->block_hash + c->out_off[r], c->seq_hash + c->out_off[r], 0);
}
void kvo_hash_batch(const u32* tok, const int64_t* offsets, size_t n_req, size_t bs, u64 salt, const int64_t* out_off,
                    u64* block_hash, u64* seq_hash, int threads) {
  struct hash_batch_ctx c = {tok, offsets, out_off, bs, salt, block_hash, seq_hash};
  /* over-partition so ragged request lengths balance across workers */
  kvo_parallel_for((long)n_req, threads, hash_batch_body, &c);
}

/* compute_block_hash_for_seq (kv_router/indexer.rs:123-135): un-chained, seed 1337, remainder dropped. */
size_t kvo_indexer_block_hashes(const u32* tok, size_t n, size_t bs, u64* out) {
  if (bs == 0) return 0;
  size_t nb = n / bs;
  for (size_t b = 0; b < nb; ++b) out[b] = hash_tokens(tok + b * bs, bs, 1337);
  return nb;
}

int kvo_hw_thread_count(void) { return kvo_hw_threads(); }
