/*
 * oracle/layout_oracle.c — TEST INFRASTRUCTURE ONLY (CPU restatement of the reference's KV block layout maps;
 * never shipped, never on the product path; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may call it).
 *
 * Every function is a plain memcpy loop that follows the reference's host loops with cudaMemcpyAsync
 * replaced by memcpy:
 *   kvo_copy_blocks_3d      <- /root/reference/lib/llm/src/kernels/block_copy.cu:246-305 (copy_blocks_memcpy_3d)
 *                              shapes from /root/reference/lib/llm/src/kv/layer.rs:174-190,322-339
 *   kvo_scatter_6d          <- block_copy.cu:577-728 (permute_scatter_memcpy; perm (3,0,1,2,4,5); block ids ignored)
 *   kvo_rearrange           <- container/deps/vllm/vllm_v0.8.4-dynamo-kv-disagg-patch.patch:934-1031
 *                              (Triton rearrange_kernel_write / _read index map)
 *   kvo_relayout_contig     <- not in the reference; the copy_blocks_3d stride formulas with the
 *                              destination prefix stride replaced
 *   kvo_reshard_push        <- patch :1427-1471 (add_remote_agent descriptor lists) composed with the Triton map
 *                              (Appendix B.4); the TP-down direction is the inverse map (B.5, new in this build)
 *
 * Parity pin: the real reference block_copy.cu is compiled to oracle/_ref/libblock_copy_ref.so (see Makefile) and the
 * GPU tests compare kvo_copy_blocks_3d / kvo_scatter_6d with it on the same inputs; the reference's own test
 * patterns (kv/layer.rs:1536-1581, :1934-2015) are reproduced in tests/.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include "par.h"

typedef uint8_t u8;

/* B.1: view [P][Nblk][suffix]; dst[p][dst_id[i]][:] = src[p][src_id[i]][:] */
int kvo_copy_blocks_3d(const void* src, void* dst, const int* src_ids, const int* dst_ids, int num_pairs,
                       int prefix_dim, int src_blocks_dim, int dst_blocks_dim, int suffix_dim, int elem_size) {
  if (!src || !dst || !src_ids || !dst_ids) return 1;
  if (num_pairs <= 0 || prefix_dim <= 0 || suffix_dim <= 0 || elem_size <= 0) return 1;
  const size_t suffix_bytes = (size_t)suffix_dim * (size_t)elem_size;
  const size_t src_pstride = (size_t)src_blocks_dim * suffix_bytes;
  const size_t dst_pstride = (size_t)dst_blocks_dim * suffix_bytes;
  for (int p = 0; p < prefix_dim; ++p)
    for (int i = 0; i < num_pairs; ++i)
      memcpy((u8*)dst + (size_t)p * dst_pstride + (size_t)dst_ids[i] * suffix_bytes,
             (const u8*)src + (size_t)p * src_pstride + (size_t)src_ids[i] * suffix_bytes, suffix_bytes);
  return 0;
}

/* Multi-layer form of B.1 (what CopyStream::trigger_all_layers does, kv/layer.rs:960-968), optionally threaded. */
struct layers_ctx {
  const uint64_t* src_layers;
  const uint64_t* dst_layers;
  const int* src_ids;
  const int* dst_ids;
  int num_pairs, prefix_dim;
  size_t suffix_bytes, src_pstride, dst_pstride;
};
static void layers_body(long b, long e, void* c_) {
  const struct layers_ctx* c = (const struct layers_ctx*)c_;
  for (long w = b; w < e; ++w) {
    int i = (int)(w % c->num_pairs);
    int p = (int)((w / c->num_pairs) % c->prefix_dim);
    int l = (int)(w / ((long)c->num_pairs * c->prefix_dim));
    memcpy((u8*)(uintptr_t)c->dst_layers[l] + (size_t)p * c->dst_pstride + (size_t)c->dst_ids[i] * c->suffix_bytes,
           (const u8*)(uintptr_t)c->src_layers[l] + (size_t)p * c->src_pstride +
               (size_t)c->src_ids[i] * c->suffix_bytes,
           c->suffix_bytes);
  }
}
int kvo_copy_blocks_layers(const uint64_t* src_layers, const uint64_t* dst_layers, int num_layers, const int* src_ids,
                           const int* dst_ids, int num_pairs, int prefix_dim, int src_blocks_dim, int dst_blocks_dim,
                           int suffix_dim, int elem_size, int threads) {
  struct layers_ctx c;
  c.src_layers = src_layers;
  c.dst_layers = dst_layers;
  c.src_ids = src_ids;
  c.dst_ids = dst_ids;
  c.num_pairs = num_pairs;
  c.prefix_dim = prefix_dim;
  c.suffix_bytes = (size_t)suffix_dim * (size_t)elem_size;
  c.src_pstride = (size_t)src_blocks_dim * c.suffix_bytes;
  c.dst_pstride = (size_t)dst_blocks_dim * c.suffix_bytes;
  kvo_parallel_for((long)num_layers * prefix_dim * num_pairs, threads, layers_body, &c);
  return 0;
}

/* B.2: TP scatter, Rust/C path. dims6 = [kv|blk][blk|kv][bs][sf][dnh][hs]; blocks 0..num_blocks identity-mapped. */
int kvo_scatter_6d(const void* src, void* dst, const uint32_t* dims, uint32_t num_dims, uint32_t elem_size,
                   uint32_t block_dim_index, uint32_t src_block_dim, uint32_t dst_block_dim, uint32_t num_blocks) {
  if (num_dims != 6) return -1;
  if (block_dim_index > 1) return -2;
  const uint32_t kv_idx = block_dim_index == 0 ? 1 : 0;
  if (dims[block_dim_index] != src_block_dim) return -3;
  if (dims[kv_idx] != 2) return -4;
  const size_t bs = dims[2], sf = dims[3];
  const size_t run = (size_t)dims[4] * dims[5] * elem_size;
  /* extents of the two leading axes on each side */
  size_t s_ext[2], d_ext[2];
  s_ext[block_dim_index] = src_block_dim;
  s_ext[kv_idx] = 2;
  d_ext[block_dim_index] = dst_block_dim;
  d_ext[kv_idx] = 2;
  for (size_t b = 0; b < num_blocks; ++b)
    for (size_t kv = 0; kv < 2; ++kv) {
      size_t i0 = block_dim_index == 0 ? b : kv;
      size_t i1 = block_dim_index == 0 ? kv : b;
      for (size_t t = 0; t < bs; ++t)
        for (size_t g = 0; g < sf; ++g) {
          size_t so = ((((i0 * s_ext[1] + i1) * bs + t) * sf) + g) * run;
          size_t dO = ((((g * d_ext[0] + i0) * d_ext[1] + i1) * bs) + t) * run;
          memcpy((u8*)dst + dO, (const u8*)src + so, run);
        }
    }
  return 0;
}

/* B.3: Triton rearrange over one K-or-V tensor [N][B][H][C]; write: t2 viewed [d][N][B][H/d][C] <- t1; read: inverse. */
int kvo_rearrange(void* t1, void* t2, size_t N, size_t B, size_t H, size_t C, size_t elem, size_t d, int is_write) {
  if (d == 0 || H % d != 0) return 1;
  const size_t hd = H / d;
  const size_t run = hd * C * elem;
  for (size_t n = 0; n < N; ++n)
    for (size_t b = 0; b < B; ++b)
      for (size_t g = 0; g < d; ++g) {
        size_t off1 = ((n * B + b) * H + g * hd) * C * elem;
        size_t off2 = (((g * N + n) * B + b) * hd) * C * elem;
        if (is_write)
          memcpy((u8*)t2 + off2, (const u8*)t1 + off1, run);
        else
          memcpy((u8*)t1 + off1, (const u8*)t2 + off2, run);
      }
  return 0;
}

/*
 * B.6: layer-separate <-> fully-contiguous-per-block [block][layer][K|V][token][head][dim].
 * layout: 0 = KvFirst ([2][N][chunk] per layer), 1 = BlockFirst ([N][2][chunk] per layer).
 * to_contig != 0 : contig[dst_id] <- layers[:, :, src_id]; else layers[:, :, dst_id] <- contig[src_id].
 */
struct contig_ctx {
  const uint64_t* layer_ptrs;
  int layout, layer_blocks, num_pairs, to_contig;
  size_t chunk_bytes, blk_stride;
  u8* contig;
  const int* src_ids;
  const int* dst_ids;
};
static void contig_body(long b, long e, void* c_) {
  const struct contig_ctx* c = (const struct contig_ctx*)c_;
  for (long w = b; w < e; ++w) {
    int i = (int)(w % c->num_pairs);
    int kv = (int)((w / c->num_pairs) % 2);
    int l = (int)(w / ((long)c->num_pairs * 2));
    int lid = c->to_contig ? c->src_ids[i] : c->dst_ids[i];
    int cid = c->to_contig ? c->dst_ids[i] : c->src_ids[i];
    size_t loff = c->layout == 0 ? ((size_t)kv * c->layer_blocks + lid) * c->chunk_bytes
                                 : ((size_t)lid * 2 + kv) * c->chunk_bytes;
    u8* lp = (u8*)(uintptr_t)c->layer_ptrs[l] + loff;
    u8* cp = c->contig + (size_t)cid * c->blk_stride + ((size_t)l * 2 + kv) * c->chunk_bytes;
    if (c->to_contig)
      memcpy(cp, lp, c->chunk_bytes);
    else
      memcpy(lp, cp, c->chunk_bytes);
  }
}
int kvo_relayout_contig(const uint64_t* layer_ptrs, int num_layers, int layout, int layer_blocks, size_t chunk_bytes,
                        void* contig, const int* src_ids, const int* dst_ids, int num_pairs, int to_contig,
                        int threads) {
  struct contig_ctx c;
  c.layer_ptrs = layer_ptrs;
  c.layout = layout;
  c.layer_blocks = layer_blocks;
  c.num_pairs = num_pairs;
  c.to_contig = to_contig;
  c.chunk_bytes = chunk_bytes;
  c.blk_stride = (size_t)num_layers * 2 * chunk_bytes;
  c.contig = (u8*)contig;
  c.src_ids = src_ids;
  c.dst_ids = dst_ids;
  kvo_parallel_for((long)num_layers * 2 * num_pairs, threads, contig_body, &c);
  return 0;
}

/*
 * B.4 / B.5: fused reshard "push" of one prefill rank's blocks into decode-rank caches (KvFirst both sides).
 *   src rank p of src_tp, h = H/src_tp heads per token row; row_bytes = h*D*elem.
 *   dst_tp >= src_tp (scatter, m = dst_tp/src_tp): group g of a row -> decode rank p*m+g, full dst row.
 *   dst_tp <  src_tp (gather,  m = src_tp/dst_tp): whole row -> decode rank p/m, sub-row (p mod m).
 * dst_layers is indexed [g * num_layers + l] for the g-th target of this rank (1 target when gathering).
 */
int kvo_reshard_push(const uint64_t* src_layers, const uint64_t* dst_layers, int num_layers, int src_blocks,
                     int dst_blocks, int bs, size_t row_bytes, int src_tp, int dst_tp, int src_rank,
                     const int* src_ids, const int* dst_ids, int num_pairs) {
  if (src_tp <= 0 || dst_tp <= 0) return 1;
  if (dst_tp >= src_tp) {
    if (dst_tp % src_tp) return 1;
    const int m = dst_tp / src_tp;
    if (row_bytes % (size_t)m) return 1;
    const size_t run = row_bytes / m;
    for (int l = 0; l < num_layers; ++l)
      for (int kv = 0; kv < 2; ++kv)
        for (int i = 0; i < num_pairs; ++i)
          for (int t = 0; t < bs; ++t)
            for (int g = 0; g < m; ++g) {
              const u8* s = (const u8*)(uintptr_t)src_layers[l] +
                            (((size_t)kv * src_blocks + src_ids[i]) * bs + t) * row_bytes + (size_t)g * run;
              u8* d = (u8*)(uintptr_t)dst_layers[(size_t)g * num_layers + l] +
                      (((size_t)kv * dst_blocks + dst_ids[i]) * bs + t) * run;
              memcpy(d, s, run);
            }
  } else {
    if (src_tp % dst_tp) return 1;
    const int m = src_tp / dst_tp;
    const int j = src_rank % m;
    const size_t drow = row_bytes * m;
    for (int l = 0; l < num_layers; ++l)
      for (int kv = 0; kv < 2; ++kv)
        for (int i = 0; i < num_pairs; ++i)
          for (int t = 0; t < bs; ++t) {
            const u8* s =
                (const u8*)(uintptr_t)src_layers[l] + (((size_t)kv * src_blocks + src_ids[i]) * bs + t) * row_bytes;
            u8* d = (u8*)(uintptr_t)dst_layers[l] + (((size_t)kv * dst_blocks + dst_ids[i]) * bs + t) * drow +
                    (size_t)j * row_bytes;
            memcpy(d, s, row_bytes);
          }
  }
  return 0;
}
