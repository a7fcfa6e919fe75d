/*
 * kvdp.h — C ABI of libkvdp.so, a B200 (sm_100a) KV-cache data plane.
 *
 * Two groups of entry points:
 *
 *  (A) LEGACY: the 14 `extern "C"` symbols that the reference's Rust modules import from its
 *      `block_copy.cu` (reference: lib/llm/src/kv/layer.rs:33-97 and lib/llm/src/kv/storage.rs:91-101).
 *      Same names, argument order, integer types and return conventions, so `kv/layer.rs` + `kv/storage.rs`
 *      link against libkvdp.so unchanged (swap `libblock_copy.a`).
 *
 *  (B) V2: batched, multi-layer entry points that map one whole request (all layers, all blocks) onto one
 *      kernel launch: block copy, layer-separate <-> contiguous-per-block relayout, TP reshard push into
 *      peer-mapped caches, host-tier offload/onboard, and batched chained xxh3 token-block hashing.
 *
 * Conventions (both groups): plain pointers and sizes only; return 0 on success; a positive return is a
 * `cudaError_t`; a negative return is one of the KVDP_E* codes (legacy scatter keeps the reference's -1..-5).
 * `kvdp_last_error()` returns a thread-local human-readable message for the last failure on this thread.
 * All sizes/strides in V2 are 64-bit (the legacy ABI's `int` dims overflow at 2^31 elements).
 */
#ifndef KVDP_H
#define KVDP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* cudaStream_t / cudaError_t cross the ABI as an opaque pointer / int, exactly as Rust sees them. */
typedef void* kvdp_stream_t;

#define KVDP_OK 0
#define KVDP_EINVAL (-22)     /* bad argument */
#define KVDP_ERANGE (-34)     /* size beyond a stated maximum */
#define KVDP_ESTATE (-71)     /* call out of order for a stateful handle */
#define KVDP_EALIGN (-72)     /* pointer/stride alignment not supported by the requested path */
#define KVDP_EDUP (-73)       /* duplicate destination block id (data race) */
#define KVDP_ENODEV (-19)     /* no CUDA device / extension not usable */

const char* kvdp_last_error(void);
/* major*10000 + minor*100 + patch */
int kvdp_version(void);
/* tunables: "copy_path" (0 auto, 1 vector ld/st, 2 TMA-load + vector store, 3 TMA load + TMA store, 4 warp-per-chunk
 * vectors for chunks <= 4 KiB),
 * "tile_bytes", "stages", "ctas_per_sm", "host_path" (pinned memory: 0 auto = a CopyStream whose
 * host-side ids form runs of >= 1 MiB on average moves them with one DMA per run through an HBM staging buffer plus one
 * kernel, anything else uses the vector kernel; 1 per-chunk DMA loop as the reference does; 2 TMA kernel; 3 vector
 * kernel always), "host_ctas", "host_dma_piece_kib" (the copy-engine path of a CopyStream and the host tier issue each
 * host-side run in pieces of this many KiB, default 8192, 0 = whole runs: an onboard and an offload running at once
 * share the link evenly only when their DMAs are equally large), "l2_hint",
 * "vec32" (copy path 1: 256-bit loads/stores when every pointer and stride is 32-byte aligned; +5 % on that path in HBM,
 * no change over NVLink), "hash_stages" (chained hashing of a big batch: number of request ranges whose chain kernels
 * run under the next range's block hashing; 0 or 1 = off, the default: everything stays on the caller's stream —
 * measured on the router batch every extra range cost more in launch ramp and tail than the overlap returned).
 * Returns KVDP_EINVAL for
 * unknown names. Intended for benchmarking; defaults are what the tests run. */
int kvdp_set_option(const char* name, int64_t value);
int64_t kvdp_get_option(const char* name);

/* ===================================================================================================
 * (A) LEGACY ABI — drop-in for block_copy.cu
 * =================================================================================================== */

/* Opaque; replaces the reference's `struct CopyStream` (block_copy.cu:371-390). */
typedef struct CopyStream CopyStream;

/* replaces copy_blocks_3d, block_copy.cu:309-353 (Rust decl kv/layer.rs:33-45; caller KvLayer::copy_blocks_to
 * :371-450). Blocking on the legacy default stream. Like the reference's default build it reports validation
 * failures on stderr and still returns 0; only a failing final synchronize is returned. */
int copy_blocks_3d(const void* src_data, void* dst_data, const int* h_src_block_ids, const int* h_dst_block_ids,
                   int num_block_pairs, int prefix_dim, int src_blocks_dim, int dst_blocks_dim, int suffix_dim,
                   int elem_size);

/* replaces copy_blocks_launcher_3d, block_copy.cu:167-243 (device id arrays; one kernel launch on `stream`). */
int copy_blocks_launcher_3d(const void* src_data, void* dst_data, const int* d_src_block_ids,
                            const int* d_dst_block_ids, int num_block_pairs, int prefix_dim, int suffix_dim,
                            int elem_size, int src_block_dim, int dst_block_dim, kvdp_stream_t stream);

/* replaces copy_blocks_memcpy_3d, block_copy.cu:246-305 (host id arrays). The reference issues one
 * cudaMemcpyAsync per (prefix, pair); here device-reachable memory (device, pinned, registered) is moved by ONE
 * kernel launch and only pageable host memory falls back to per-chunk DMA. */
int copy_blocks_memcpy_3d(const void* src_data, void* dst_data, const int* h_src_block_ids,
                          const int* h_dst_block_ids, int num_block_pairs, int prefix_dim, int suffix_dim,
                          int elem_size, int src_block_dim, int dst_block_dim, kvdp_stream_t stream);

/* replaces copy_stream_create, block_copy.cu:469-473 (Rust decl kv/layer.rs:47-51). Always returns 0 and stores a
 * handle, as the reference does (Q1); a handle whose resources failed to allocate makes every later call fail. */
int copy_stream_create(CopyStream** stream, int num_layers, int num_blocks);
/* replaces copy_stream_destroy, block_copy.cu:476-480 (exported by the reference but never called from Rust, Q2). */
int copy_stream_destroy(CopyStream* stream);
/* replaces copy_stream_prepare_block_ids, block_copy.cu:484-499 (Rust decl kv/layer.rs:53-58). The ids are copied
 * into pinned staging owned by the handle before the async upload, so the caller's buffers may be freed on return
 * (fixes Q3 without changing the signature). No bound check against the ctor's num_blocks in the reference; here
 * a larger n grows the device arrays. */
int copy_stream_prepare_block_ids(CopyStream* cs, int* src_block_ids, int* dst_block_ids, int num_blocks);
/* replaces copy_stream_launch, block_copy.cu:502-509 (Rust decl kv/layer.rs:60-70, unused there). Kernel path. */
int copy_stream_launch(CopyStream* cs, const void* src_data, void* dst_data, int prefix_dim, int suffix_dim,
                       int elem_size, int src_block_dim, int dst_block_dim);
/* replaces copy_stream_memcpy, block_copy.cu:512-519 (Rust decl kv/layer.rs:72-81; caller CopyStream::trigger_layer
 * :906-958). */
int copy_stream_memcpy(CopyStream* cs, const void* src_data, void* dst_data, int prefix_dim, int suffix_dim,
                       int elem_size, int src_block_dim, int dst_block_dim);
/* replaces copy_stream_sync, block_copy.cu:522-527 (Rust decl kv/layer.rs:83). */
int copy_stream_sync(CopyStream* cs);
/* replaces copy_stream_scatter, block_copy.cu:731-758 + permute_scatter_memcpy :577-728 (Rust decl kv/layer.rs:85-95;
 * caller CopyStream::scatter_copy_layer :1027-1130). dims = 6-D source dims [kv|blk][blk|kv][bs][sf][dnh][hs];
 * permutation (3,0,1,2,4,5); copies blocks 0..n identity-mapped where n is the staged id count (Q4).
 * Returns -1 (num_dims != 6), -2 (block_dim_index not 0/1), -3 (dims[block_dim_index] != src_block_dim),
 * -4 (kv extent != 2), -5 (CUDA failure), 0 otherwise. */
int copy_stream_scatter(CopyStream* cs, const void* src_data, void* dst_data, const uint32_t* dims,
                        uint32_t num_dims, uint32_t elem_size, uint32_t block_dim_index, uint32_t src_block_dim,
                        uint32_t dst_block_dim);

/* replace cuda_malloc_host / cuda_free_host / cuda_memcpy_async / cuda_memcpy_sync, block_copy.cu:529-555
 * (Rust decls kv/storage.rs:91-101). */
int cuda_malloc_host(void** ptr, size_t size);
int cuda_free_host(void* ptr);
int cuda_memcpy_async(void* dst, const void* src, size_t count, kvdp_stream_t stream);
int cuda_memcpy_sync(void* dst, const void* src, size_t count);

/* ===================================================================================================
 * (B) V2 API
 * =================================================================================================== */

/* ---- device slabs (what kv/storage.rs gets from cudarc: DeviceStorageOwned::new, storage.rs:189-230) ---- */
int kvdp_device_count(int* count);
int kvdp_device_alloc(void** ptr, size_t bytes, int device); /* zero-filled, like DeviceStorageOwned (storage.rs:196-205) */
int kvdp_device_free(void* ptr);
int kvdp_memset(void* ptr, int value, size_t bytes);
/* 0 = unregistered/pageable host, 1 = pinned or registered host, 2 = device, 3 = managed. `device` receives the
 * owning device ordinal for types 2/3, else -1. */
int kvdp_pointer_kind(const void* ptr, int* kind, int* device);

/* ---- copy plans: immutable, device-resident description of a (source set, destination set) pair ----------
 * The counterpart of CopyStreamBlockMap (kv/layer.rs:614-690): computed once per block-set pair and direction.
 *
 * One "chunk" is the bytes of one (layer, prefix, block pair). The source chunk is contiguous:
 *   rows * row_groups * run_bytes bytes at  src_layer[l] + p*src_prefix_stride + src_id*src_block_stride.
 * Row r, group g, i.e. source bytes [ (r*row_groups+g)*run_bytes, +run_bytes ), land at
 *   dst_layer[g][l] + p*dst_prefix_stride + dst_id*dst_block_stride + r*dst_row_stride + dst_col_offset.
 * A plain block copy is rows=1,row_groups=1; a TP-up scatter has row_groups = dst_tp/src_tp targets;
 * a TP-down gather has row_groups=1, dst_row_stride = m*run_bytes, dst_col_offset = (rank % m)*run_bytes. */
typedef struct kvdp_plan kvdp_plan;
typedef struct kvdp_plan_desc {
  int32_t num_layers;
  int32_t row_groups;
  int32_t prefix_dim;
  int32_t rows;
  int64_t run_bytes;
  int64_t src_prefix_stride, src_block_stride;
  int64_t dst_prefix_stride, dst_block_stride, dst_row_stride, dst_col_offset;
  int64_t src_row_stride, src_col_offset; /* 0,0 = contiguous source chunk; else row r of the chunk is read from
                                             chunk_base + r*src_row_stride + src_col_offset (a column range of wider
                                             rows, e.g. pulling this rank's heads out of a smaller-TP peer's cache) */
  int64_t src_blocks, dst_blocks;    /* extents of the block axes, used to bounds-check ids (0 = unchecked) */
  const uint64_t* src_layer_ptrs;    /* host array [num_layers] */
  const uint64_t* dst_layer_ptrs;    /* host array [row_groups][num_layers] */
  int32_t device;                    /* device that runs the kernel (owner of the source) */
} kvdp_plan_desc;
int kvdp_plan_create(kvdp_plan** out, const kvdp_plan_desc* desc);
int kvdp_plan_destroy(kvdp_plan* plan);
int64_t kvdp_plan_bytes_per_pair(const kvdp_plan* plan); /* payload bytes moved per block pair over all layers */

/* same-layout multi-layer block copy; arguments as CopyStreamBlockMap::new derives them (kv/layer.rs:648-688) */
int kvdp_plan_block_copy(kvdp_plan** out, const uint64_t* src_layer_ptrs, const uint64_t* dst_layer_ptrs,
                         int32_t num_layers, int32_t prefix_dim, int64_t src_block_dim, int64_t dst_block_dim,
                         int64_t suffix_dim, int32_t elem_size, int32_t device);
/* layer-separate (layout 0 = KvFirst [2][N][chunk], 1 = BlockFirst [N][2][chunk]; kv/layer.rs:99-106,322-339)
 * <-> fully-contiguous-per-block slab [block][layer][K|V][chunk] (not in the reference: SURVEY F7 / App. B.6).
 * to_contig=1: ids are (layer-separate src id, contiguous dst id); 0: (contiguous src id, layer-separate dst id). */
int kvdp_plan_contig(kvdp_plan** out, const uint64_t* layer_ptrs, int32_t num_layers, int32_t layout,
                     int64_t layer_blocks, int64_t chunk_bytes, void* contig_base, int64_t contig_blocks,
                     int32_t to_contig, int32_t device);
/* TP reshard push of prefill rank `src_rank` (KvFirst caches [2][N][bs][h][D]) into decode-rank caches, which may be
 * peer-mapped pointers. Replaces rearrange_tensors + the NIXL descriptor lists of the patched vLLM connector
 * (container/deps/vllm/vllm_v0.8.4-dynamo-kv-disagg-patch.patch:1003-1031, :1427-1471). dst_tp >= src_tp: targets are
 * decode ranks src_rank*m+g, g<m, dst_layer_ptrs is [m][num_layers]; dst_tp < src_tp (the direction the reference
 * asserts against, patch:1437): one target, decode rank src_rank/m, dst_layer_ptrs is [1][num_layers]. */
int kvdp_plan_reshard(kvdp_plan** out, const uint64_t* src_layer_ptrs, const uint64_t* dst_layer_ptrs,
                      int32_t num_layers, int64_t src_blocks, int64_t dst_blocks, int32_t block_size,
                      int64_t src_row_bytes, int32_t src_tp, int32_t dst_tp, int32_t src_rank, int32_t device);

/* ---- transfer engine: one stream + id staging; the v2 counterpart of CopyStream (kv/layer.rs:785-979) ---- */
typedef struct kvdp_xfer kvdp_xfer;
int kvdp_xfer_create(kvdp_xfer** out, int32_t max_pairs, int32_t device);
int kvdp_xfer_destroy(kvdp_xfer* x);
kvdp_stream_t kvdp_xfer_stream(kvdp_xfer* x);
#define KVDP_XFER_CHECK_DST_UNIQUE 1u /* reject duplicate destination ids (kv/layer.rs:848-861 does so in debug) */
/* Asynchronous: stages the ids (host arrays, copied before return), then ONE launch that moves layers
 * [layer_begin, layer_end) of every pair. If done_flag != NULL the kernel itself publishes *done_flag = done_value
 * (release, system scope) once the last CTA has drained its stores; the flag may live in device, peer-mapped or
 * pinned memory. num_pairs == 0 moves nothing but still publishes the flag in stream order (the reference's
 * "no blocks to write, notify only" case, patch:1373-1377). Calls on one engine are ordered; the completion counter
 * is per engine, so do not share one engine between threads without external locking. */
int kvdp_xfer_run(kvdp_xfer* x, const kvdp_plan* plan, const int32_t* h_src_ids, const int32_t* h_dst_ids,
                  int32_t num_pairs, int32_t layer_begin, int32_t layer_end, uint32_t flags, uint32_t* done_flag,
                  uint32_t done_value);
/* After everything previously enqueued on this engine is complete AND visible system-wide, store `value` to
 * *flag (device, peer-mapped or pinned memory). The receiver polls or stream-waits on it: this replaces the NIXL
 * notification of write_blocks (patch:1406-1415) / get_new_notifs (:1421-1425). */
int kvdp_xfer_signal(kvdp_xfer* x, uint32_t* flag, uint32_t value);
int kvdp_xfer_sync(kvdp_xfer* x);
/* Allocation-free launch on a caller stream with ids already resident on the device. */
int kvdp_plan_launch(const kvdp_plan* plan, const int32_t* d_src_ids, const int32_t* d_dst_ids, int32_t num_pairs,
                     int32_t layer_begin, int32_t layer_end, kvdp_stream_t stream);
/* Block the calling host thread until everything enqueued on `stream` has completed. */
int kvdp_stream_synchronize(kvdp_stream_t stream);
/* Make `stream` wait until *flag >= value (receiver side of kvdp_xfer_signal). */
int kvdp_stream_wait_flag(kvdp_stream_t stream, uint32_t* flag, uint32_t value);

/* System-scope compare-and-swap on one 8-byte-aligned 64-bit word of device memory (local, peer-mapped or IPC-mapped),
 * executed by a one-thread kernel on `device`; *old receives the value that was found (the swap happened iff
 * *old == expected). Synchronous. The connector's senders claim their notification lane in a receiver's mailbox with it:
 * the counterpart of NIXL keeping one notification queue per remote agent (patch:1406-1425, nixl_wrapper.get_new_notifs
 * returns {agent_name: [...]}). */
int kvdp_flag_cas64(uint64_t* addr, uint64_t expected, uint64_t desired, uint64_t* old, int32_t device);

/* Page-lock and map a host range the caller owns (page-aligned, e.g. an mmap of a memfd or of POSIX shared memory) so
 * that every entry point treats it like cuda_malloc_host memory (copy engines and SMs reach it; a CopyStream takes its
 * copy-engine path on it). A host tier kept in memory that several worker processes of one node map - each registers it
 * in its own process - lets ANY of their GPUs onboard a block over its own PCIe link: the decode GPU fetches part of a
 * prefill worker's host-resident blocks itself while the prefill GPU relays the rest over NVLink
 * (dynamo_exp_b200/shared_host.py, bench.py --gpus N e2e). The reference pins host KV storage per process
 * (kv/storage.rs:91-101 cuda_malloc_host); this is the counterpart for memory it did not allocate.
 * kvdp_host_unregister before unmapping. */
int kvdp_host_register(void* ptr, size_t bytes);
int kvdp_host_unregister(void* ptr);

/* ---- peers (replaces nixl_agent.register_memory / add_remote_agent metadata exchange, patch:1158-1186,1427-1436) */
#define KVDP_IPC_HANDLE_BYTES 64
int kvdp_peer_enable(int32_t device, int32_t peer_device);            /* same-process peers */
int kvdp_ipc_export(void* device_ptr, uint8_t handle[KVDP_IPC_HANDLE_BYTES]); /* ptr must be a kvdp_device_alloc base */
/* Same for a pointer INSIDE a larger allocation (e.g. a tensor carved out of a framework's caching allocator): the
 * handle names the whole allocation and *offset is device_ptr's distance from its base; the importer adds it back. */
int kvdp_ipc_export_range(void* device_ptr, uint8_t handle[KVDP_IPC_HANDLE_BYTES], uint64_t* offset);
int kvdp_ipc_import(const uint8_t handle[KVDP_IPC_HANDLE_BYTES], int32_t device, void** mapped_ptr);
int kvdp_ipc_close(void* mapped_ptr);

/* ---- host tier (offload / onboard through pinned memory) ----------------------------
 * Host blocks are stored contiguous-per-block in a pinned slab; the device side is layer-separate. Offload =
 * gather into an HBM staging ring (one kernel) + one DMA per ring slot on a side stream; onboard is the reverse. */
typedef struct kvdp_host_tier kvdp_host_tier;
int kvdp_host_tier_create(kvdp_host_tier** out, const uint64_t* layer_ptrs, int32_t num_layers, int32_t layout,
                          int64_t layer_blocks, int64_t chunk_bytes, void* pinned_slab, int64_t host_blocks,
                          int32_t staging_blocks, int32_t device);
int kvdp_host_tier_destroy(kvdp_host_tier* t);
int kvdp_host_tier_offload(kvdp_host_tier* t, const int32_t* h_device_ids, const int32_t* h_host_ids, int32_t n);
int kvdp_host_tier_onboard(kvdp_host_tier* t, const int32_t* h_host_ids, const int32_t* h_device_ids, int32_t n);
int kvdp_host_tier_sync(kvdp_host_tier* t);

/* ---- token-block hashing (lib/tokens + kv_router/indexer) ------------------------------------------------
 * For request r with tokens[offsets[r] .. offsets[r+1]) and nb_r = len_r / block_size full blocks:
 *   block_hash[o_r + b] = xxh3_64(LE bytes of block b, seed = salt)              (lib/tokens/src/lib.rs:197-215)
 *   seq_hash[o_r + 0]   = block_hash[o_r]; seq_hash[o_r + b] = xxh3_64(LE64(seq_hash[o_r+b-1]) || LE64(block_hash[o_r+b]), salt)
 *                                                                                 (lib/tokens/src/lib.rs:238-261,369-404)
 *   o_r = sum_{q<r} nb_q is written to block_offsets[r] (n_req+1 entries). Trailing tokens are not hashed.
 * With salt = 1337 block_hash[] is the router's LocalBlockHash (kv_router/indexer.rs:64,123-135).
 * KVDP_HASH_TWIN_FIRST reproduces lib/llm/src/tokens.rs:126-144 (first block chained onto parent 0). */
#define KVDP_HASH_TWIN_FIRST 1u
size_t kvdp_hash_workspace_bytes(int64_t n_req);
/* all pointers are device pointers; seq_hash may be NULL (un-chained only). total_tokens = offsets[n_req] as known to
 * the host (sizes the grid). Asynchronous and ordered on `stream` like one kernel. With the "hash_stages" option set
 * above 1, a big chained batch (>= 4096 requests, >= 64 MiB of tokens) is cut into that many request ranges: the
 * serial, latency-bound chain kernel of one range runs on a library-owned side stream under the HBM-bound block hashing
 * of the next range and is joined back onto `stream` before the call's work counts as complete (streams and events are
 * created on the first such call per device). The result is the same bits either way. */
int kvdp_hash_token_blocks(const uint32_t* d_tokens, const int64_t* d_offsets, int64_t n_req, int64_t total_tokens,
                           int32_t block_size, uint64_t salt, int64_t* d_block_offsets, uint64_t* d_block_hash,
                           uint64_t* d_seq_hash, void* d_workspace, size_t workspace_bytes, uint32_t flags,
                           kvdp_stream_t stream);
/* The same hashing for a caller that must not allocate device memory: a thread inside an engine whose memory manager
 * took the whole GPU for the KV cache at start-up, or code where a cudaMalloc (which serialises the device) is not
 * acceptable. It takes no workspace; the scan scratch lives in a 64 KiB __device__ array inside the library (held by
 * every GPU that loads libkvdp.so, until exit). Limits and rules:
 *   - at most KVDP_HASH_NOALLOC_MAX_REQUESTS requests per call; more returns KVDP_ERANGE and launches nothing;
 *   - it calls no cudaMalloc / cudaFree / allocating routine; inputs and outputs are still the caller's buffers;
 *   - the first call on each device looks up the scratch's device address (cudaGetSymbolAddress) and sets a kernel
 *     attribute; make that first call outside any stream capture;
 *   - all calls on one device share the one scratch: two calls whose kernels can overlap on the GPU (different
 *     streams, no ordering between them) race on it. Calls ordered on one stream are fine; stale contents from the
 *     previous call are never read.
 * This entry point has no counterpart in the reference. */
#define KVDP_HASH_NOALLOC_MAX_REQUESTS (1 << 22)
int kvdp_hash_token_blocks_noalloc(const uint32_t* d_tokens, const int64_t* d_offsets, int64_t n_req,
                                   int64_t total_tokens, int32_t block_size, uint64_t salt, int64_t* d_block_offsets,
                                   uint64_t* d_block_hash, uint64_t* d_seq_hash, uint32_t flags, kvdp_stream_t stream);
/* Host-buffer form (what a Rust caller of lib/tokens holds): uploads, hashes, downloads, synchronises.
 * h_block_offsets has n_req+1 entries; hash arrays must hold sum(len_r / block_size) entries. */
int kvdp_hash_token_blocks_host(const uint32_t* h_tokens, const int64_t* h_offsets, int64_t n_req, int32_t block_size,
                                uint64_t salt, int64_t* h_block_offsets, uint64_t* h_block_hash, uint64_t* h_seq_hash,
                                uint32_t flags, int32_t device);
/* compute_hash(data, seed) on the device for one byte string (lib/tokens/src/lib.rs:44-46); test/diagnostic entry
 * that exercises every XXH3 length class. n strings of equal length `len`, packed; out[n]. */
int kvdp_xxh3_64_batch(const uint8_t* d_data, int64_t n, int64_t len, uint64_t seed, uint64_t* d_out,
                       kvdp_stream_t stream);

/* ---- prefix index: the KV router's RadixTree with batched GPU lookups (lib/llm/src/kv_router/indexer.rs:187-387) ----
 * Events are applied on the host in call order (RadixTree::apply_event :284-378, remove_worker :380-386) and mirrored
 * into device tables; find_matches (:239-277) runs for a whole batch of requests in one launch. Workers are slots
 * 0..max_workers-1 (<= 64); the host layer maps the reference's i64 WorkerId to a slot. */
typedef struct kvdp_index kvdp_index;
int kvdp_index_create(kvdp_index** out, int64_t max_nodes, int32_t max_workers, int32_t device);
int kvdp_index_destroy(kvdp_index* ix);
kvdp_stream_t kvdp_index_stream(kvdp_index* ix);
/* Stored event: blocks[i] = (block_hashes[i] = engine's ExternalSequenceBlockHash, tokens_hashes[i] = LocalBlockHash).
 * Returns 0 on success, 1 when has_parent is set and this worker has no block `parent_hash` (the reference logs a
 * warning and skips the event, :305-315), negative KVDP_E* on misuse or exhausted capacity. */
int kvdp_index_apply_stored(kvdp_index* ix, int32_t worker, int32_t has_parent, uint64_t parent_hash,
                            const uint64_t* block_hashes, const uint64_t* tokens_hashes, int32_t n);
/* Removed event; unknown block hashes are skipped like in the reference (:354-364). */
int kvdp_index_apply_removed(kvdp_index* ix, int32_t worker, const uint64_t* block_hashes, int32_t n);
int kvdp_index_remove_worker(kvdp_index* ix, int32_t worker);
/* Upload pending changes (find_matches does it implicitly). */
int kvdp_index_flush(kvdp_index* ix);
/* Requests are CSR sequences of LocalBlockHash in device memory (d_seq_offsets has n_req+1 entries). d_scores is
 * [n_req][num_workers] u32 in worker-slot order: how many leading blocks of the request each worker holds;
 * d_matched_depth[n_req] (optional) = number of blocks found in the tree. early_exit as in the reference: stop at the
 * first block held by exactly one worker. Asynchronous on `stream` (NULL = the index's own stream). */
int kvdp_index_find_matches(kvdp_index* ix, const uint64_t* d_hashes, const int64_t* d_seq_offsets, int64_t n_req,
                            int32_t early_exit, int32_t num_workers, uint32_t* d_scores, int32_t* d_matched_depth,
                            kvdp_stream_t stream);
/* The same lookup for host arrays (what a router thread holds): one upload, one launch, one download, synchronous.
 * h_seq_offsets[0] must be 0; h_scores is [n_req][num_workers], h_matched_depth[n_req] may be NULL. */
int kvdp_index_find_matches_host(kvdp_index* ix, const uint64_t* h_hashes, const int64_t* h_seq_offsets, int64_t n_req,
                                 int32_t early_exit, int32_t num_workers, uint32_t* h_scores, int32_t* h_matched_depth);
/* KvIndexer::find_matches_for_request (indexer.rs:639-651) for a batch: token sequences in (CSR, h_offsets[0] == 0), the
 * full blocks are hashed with seed 1337 and matched in the same round trip. Outputs as above. */
int kvdp_index_find_matches_for_tokens_host(kvdp_index* ix, const uint32_t* h_tokens, const int64_t* h_offsets,
                                            int64_t n_req, int32_t block_size, int32_t early_exit, int32_t num_workers,
                                            uint32_t* h_scores, int32_t* h_matched_depth);
/* Access-frequency tracking of a tree built with an expiration duration (indexer.rs:217-223, 251-267): call once per
 * looked-up sequence, in request order, next to the lookup. Every block on the matched path drops visit times older than
 * expiration_ns, reports how many remain (zero counts are skipped, like OverlapScores::add_frequency) and records
 * now_ns. Host-side state. Returns the number of frequencies written to h_frequencies (cap >= n), negative on misuse. */
int64_t kvdp_index_touch(kvdp_index* ix, const uint64_t* h_hashes, int64_t n, int32_t early_exit, int64_t now_ns,
                         int64_t expiration_ns, uint64_t* h_frequencies, int64_t cap);
/* Introspection for tests: what = 0 nodes; 1 children of node `arg`; 2 lookup entries of worker `arg`; 3 workers on node
 * `arg`; 4 root's child for tokens hash `arg` (node id or -1); 5 workers with a lookup table; 6 nodes whose worker set
 * still holds slot `arg` (0 right after kvdp_index_remove_worker means the slot can serve another worker id). */
int64_t kvdp_index_stat(kvdp_index* ix, int32_t what, uint64_t arg);

/* ================================================================== §I. Engine-facing KV-event ABI
 * The five symbols of the reference's libdynamo_llm_capi.so (lib/bindings/c/src/lib.rs:50-311), which the patched vLLM
 * loads from $VLLM_KV_CAPI_PATH with ctypes: same names, argument meaning and 0/1 results, so pointing that variable at
 * libkvdp.so swaps the implementation. Token blocks are hashed with seed 1337 exactly as compute_block_hash_for_seq
 * does (kv_router/indexer.rs:123-135), but in batches on the GPU: the calls only queue; one publisher thread hashes
 * everything that queued up in one launch and emits the events in arrival order. The reference's transport (NATS,
 * subject "namespace.<ns>.component.<component>.kv_events", payload = RouterEvent JSON) is not rebuilt: the sink is
 * chosen with kvdp_event_sink_* below. Differences from the reference, on purpose: publishing before dynamo_llm_init
 * returns ERR (the reference aborts the process); after dynamo_llm_shutdown a new dynamo_llm_init is accepted. */
typedef uint32_t dynamo_llm_result_t;
#define DYNAMO_LLM_OK 0u
#define DYNAMO_LLM_ERR 1u
dynamo_llm_result_t dynamo_llm_init(const char* namespace_c_str, const char* component_c_str, int64_t worker_id,
                                    uint32_t kv_block_size);                                  /* lib.rs:50-113 */
dynamo_llm_result_t dynamo_llm_shutdown(void);                                                /* lib.rs:115-128 */
dynamo_llm_result_t dynamo_llm_load_publisher_create(void);                                   /* lib.rs:130-133 */
/* parent_hash == NULL means "no parent". Blocks whose token count differs from kv_block_size end the list (:191-210). */
dynamo_llm_result_t dynamo_kv_event_publish_stored(uint64_t event_id, const uint32_t* token_ids,
                                                   const size_t* num_block_tokens, const uint64_t* block_ids,
                                                   size_t num_blocks, const uint64_t* parent_hash,
                                                   uint64_t lora_id);                         /* lib.rs:260-295 */
dynamo_llm_result_t dynamo_kv_event_publish_removed(uint64_t event_id, const uint64_t* block_ids,
                                                    size_t num_blocks);                       /* lib.rs:297-311 */
/* Sinks (no reference counterpart). A callback receives every event as (subject, RouterEvent JSON, length) on the
 * publisher thread; an index sink applies the events to a prefix index under worker slot `worker_slot` (either, both or
 * neither may be set; pass NULL to clear; kvdp_index_destroy detaches an index that is still attached). With neither, events wait in an outbox of 65 536 entries for
 * kvdp_event_poll: returns 1 and copies one event (not NUL-terminated, *len bytes), 0 when empty, KVDP_ERANGE with
 * *len = bytes needed when `cap` is too small. kvdp_event_flush blocks until every event published so far has been
 * hashed and delivered. kvdp_event_stat: 0 submitted, 1 delivered, 2 hash batches, 3 blocks hashed, 4 short blocks
 * that ended a list, 5 stored events skipped by the index (unknown parent), 6 events lost to a failed hash batch,
 * 7 outbox overflow drops, 8 outbox length. */
typedef void (*kvdp_event_cb)(void* user, const char* subject, const char* json, size_t len);
int kvdp_event_sink_callback(kvdp_event_cb cb, void* user);
int kvdp_event_sink_index(kvdp_index* ix, int32_t worker_slot);
int kvdp_event_flush(void);
int kvdp_event_poll(char* buf, size_t cap, size_t* len);
int64_t kvdp_event_stat(int32_t what);

#ifdef __cplusplus
}
#endif
#endif /* KVDP_H */
