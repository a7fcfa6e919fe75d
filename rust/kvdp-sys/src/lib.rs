//! Raw bindings to `libkvdp.so` — generated from include/kvdp.h (see the header for the contract of every call).
//!
//! Source only: never compiled in the image this repository is built in (no Rust toolchain there). The legacy block
//! (`copy_blocks_3d` .. `cuda_memcpy_sync`) has exactly the signatures that dynamo-llm's `kv/layer.rs:33-97` and
//! `kv/storage.rs:91-101` declare, so those modules can link against libkvdp.so without this crate.
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_int, c_void};

#[repr(C)] pub struct CopyStream { _p: [u8; 0] }
#[repr(C)] pub struct KvdpPlan { _p: [u8; 0] }
#[repr(C)] pub struct KvdpXfer { _p: [u8; 0] }
#[repr(C)] pub struct KvdpHostTier { _p: [u8; 0] }
#[repr(C)] pub struct KvdpIndex { _p: [u8; 0] }

/// `kvdp_plan_desc` (include/kvdp.h)
#[repr(C)]
pub struct KvdpPlanDesc {
    pub num_layers: i32,
    pub row_groups: i32,
    pub prefix_dim: i32,
    pub rows: i32,
    pub run_bytes: i64,
    pub src_prefix_stride: i64,
    pub src_block_stride: i64,
    pub dst_prefix_stride: i64,
    pub dst_block_stride: i64,
    pub dst_row_stride: i64,
    pub dst_col_offset: i64,
    pub src_row_stride: i64,
    pub src_col_offset: i64,
    pub src_blocks: i64,
    pub dst_blocks: i64,
    pub src_layer_ptrs: *const u64,
    pub dst_layer_ptrs: *const u64,
    pub device: i32,
}

pub const KVDP_OK: c_int = 0;
pub const KVDP_EINVAL: c_int = -22;
pub const KVDP_ERANGE: c_int = -34;
pub const KVDP_ESTATE: c_int = -71;
pub const KVDP_EALIGN: c_int = -72;
pub const KVDP_EDUP: c_int = -73;
pub const KVDP_ENODEV: c_int = -19;
pub const KVDP_XFER_CHECK_DST_UNIQUE: u32 = 1;
pub const KVDP_HASH_TWIN_FIRST: u32 = 1;
pub const KVDP_IPC_HANDLE_BYTES: usize = 64;
pub const KVDP_HASH_NOALLOC_MAX_REQUESTS: i64 = 1 << 22;

extern "C" {
    pub fn kvdp_last_error() -> *const c_char;
    pub fn kvdp_version() -> c_int;
    pub fn kvdp_set_option(name: *const c_char, value: i64) -> c_int;
    pub fn kvdp_get_option(name: *const c_char) -> i64;
    pub fn copy_blocks_3d(src_data: *const c_void, dst_data: *mut c_void, h_src_block_ids: *const c_int, h_dst_block_ids: *const c_int, num_block_pairs: c_int, prefix_dim: c_int, src_blocks_dim: c_int, dst_blocks_dim: c_int, suffix_dim: c_int, elem_size: c_int) -> c_int;
    pub fn copy_blocks_launcher_3d(src_data: *const c_void, dst_data: *mut c_void, d_src_block_ids: *const c_int, d_dst_block_ids: *const c_int, num_block_pairs: c_int, prefix_dim: c_int, suffix_dim: c_int, elem_size: c_int, src_block_dim: c_int, dst_block_dim: c_int, stream: *mut c_void) -> c_int;
    pub fn copy_blocks_memcpy_3d(src_data: *const c_void, dst_data: *mut c_void, h_src_block_ids: *const c_int, h_dst_block_ids: *const c_int, num_block_pairs: c_int, prefix_dim: c_int, suffix_dim: c_int, elem_size: c_int, src_block_dim: c_int, dst_block_dim: c_int, stream: *mut c_void) -> c_int;
    pub fn copy_stream_create(stream: *mut *mut CopyStream, num_layers: c_int, num_blocks: c_int) -> c_int;
    pub fn copy_stream_destroy(stream: *mut CopyStream) -> c_int;
    pub fn copy_stream_prepare_block_ids(cs: *mut CopyStream, src_block_ids: *mut c_int, dst_block_ids: *mut c_int, num_blocks: c_int) -> c_int;
    pub fn copy_stream_launch(cs: *mut CopyStream, src_data: *const c_void, dst_data: *mut c_void, prefix_dim: c_int, suffix_dim: c_int, elem_size: c_int, src_block_dim: c_int, dst_block_dim: c_int) -> c_int;
    pub fn copy_stream_memcpy(cs: *mut CopyStream, src_data: *const c_void, dst_data: *mut c_void, prefix_dim: c_int, suffix_dim: c_int, elem_size: c_int, src_block_dim: c_int, dst_block_dim: c_int) -> c_int;
    pub fn copy_stream_sync(cs: *mut CopyStream) -> c_int;
    pub fn copy_stream_scatter(cs: *mut CopyStream, src_data: *const c_void, dst_data: *mut c_void, dims: *const u32, num_dims: u32, elem_size: u32, block_dim_index: u32, src_block_dim: u32, dst_block_dim: u32) -> c_int;
    pub fn cuda_malloc_host(ptr: *mut *mut c_void, size: usize) -> c_int;
    pub fn cuda_free_host(ptr: *mut c_void) -> c_int;
    pub fn cuda_memcpy_async(dst: *mut c_void, src: *const c_void, count: usize, stream: *mut c_void) -> c_int;
    pub fn cuda_memcpy_sync(dst: *mut c_void, src: *const c_void, count: usize) -> c_int;
    pub fn kvdp_device_count(count: *mut c_int) -> c_int;
    pub fn kvdp_device_alloc(ptr: *mut *mut c_void, bytes: usize, device: c_int) -> c_int;
    pub fn kvdp_device_free(ptr: *mut c_void) -> c_int;
    pub fn kvdp_memset(ptr: *mut c_void, value: c_int, bytes: usize) -> c_int;
    pub fn kvdp_pointer_kind(ptr: *const c_void, kind: *mut c_int, device: *mut c_int) -> c_int;
    pub fn kvdp_plan_create(out: *mut *mut KvdpPlan, desc: *const KvdpPlanDesc) -> c_int;
    pub fn kvdp_plan_destroy(plan: *mut KvdpPlan) -> c_int;
    pub fn kvdp_plan_bytes_per_pair(plan: *const KvdpPlan) -> i64;
    pub fn kvdp_plan_block_copy(out: *mut *mut KvdpPlan, src_layer_ptrs: *const u64, dst_layer_ptrs: *const u64, num_layers: i32, prefix_dim: i32, src_block_dim: i64, dst_block_dim: i64, suffix_dim: i64, elem_size: i32, device: i32) -> c_int;
    pub fn kvdp_plan_contig(out: *mut *mut KvdpPlan, layer_ptrs: *const u64, num_layers: i32, layout: i32, layer_blocks: i64, chunk_bytes: i64, contig_base: *mut c_void, contig_blocks: i64, to_contig: i32, device: i32) -> c_int;
    pub fn kvdp_plan_reshard(out: *mut *mut KvdpPlan, src_layer_ptrs: *const u64, dst_layer_ptrs: *const u64, num_layers: i32, src_blocks: i64, dst_blocks: i64, block_size: i32, src_row_bytes: i64, src_tp: i32, dst_tp: i32, src_rank: i32, device: i32) -> c_int;
    pub fn kvdp_xfer_create(out: *mut *mut KvdpXfer, max_pairs: i32, device: i32) -> c_int;
    pub fn kvdp_xfer_destroy(x: *mut KvdpXfer) -> c_int;
    pub fn kvdp_xfer_stream(x: *mut KvdpXfer) -> *mut c_void;
    pub fn kvdp_xfer_run(x: *mut KvdpXfer, plan: *const KvdpPlan, h_src_ids: *const i32, h_dst_ids: *const i32, num_pairs: i32, layer_begin: i32, layer_end: i32, flags: u32, done_flag: *mut u32, done_value: u32) -> c_int;
    pub fn kvdp_xfer_signal(x: *mut KvdpXfer, flag: *mut u32, value: u32) -> c_int;
    pub fn kvdp_xfer_sync(x: *mut KvdpXfer) -> c_int;
    pub fn kvdp_plan_launch(plan: *const KvdpPlan, d_src_ids: *const i32, d_dst_ids: *const i32, num_pairs: i32, layer_begin: i32, layer_end: i32, stream: *mut c_void) -> c_int;
    pub fn kvdp_stream_synchronize(stream: *mut c_void) -> c_int;
    pub fn kvdp_stream_wait_flag(stream: *mut c_void, flag: *mut u32, value: u32) -> c_int;
    pub fn kvdp_flag_cas64(addr: *mut u64, expected: u64, desired: u64, old: *mut u64, device: i32) -> c_int;
    pub fn kvdp_host_register(ptr: *mut c_void, bytes: usize) -> c_int;
    pub fn kvdp_host_unregister(ptr: *mut c_void) -> c_int;
    pub fn kvdp_peer_enable(device: i32, peer_device: i32) -> c_int;
    pub fn kvdp_ipc_export(device_ptr: *mut c_void, handle: *mut u8) -> c_int;
    pub fn kvdp_ipc_export_range(device_ptr: *mut c_void, handle: *mut u8, offset: *mut u64) -> c_int;
    pub fn kvdp_ipc_import(handle: *const u8, device: i32, mapped_ptr: *mut *mut c_void) -> c_int;
    pub fn kvdp_ipc_close(mapped_ptr: *mut c_void) -> c_int;
    pub fn kvdp_host_tier_create(out: *mut *mut KvdpHostTier, layer_ptrs: *const u64, num_layers: i32, layout: i32, layer_blocks: i64, chunk_bytes: i64, pinned_slab: *mut c_void, host_blocks: i64, staging_blocks: i32, device: i32) -> c_int;
    pub fn kvdp_host_tier_destroy(t: *mut KvdpHostTier) -> c_int;
    pub fn kvdp_host_tier_offload(t: *mut KvdpHostTier, h_device_ids: *const i32, h_host_ids: *const i32, n: i32) -> c_int;
    pub fn kvdp_host_tier_onboard(t: *mut KvdpHostTier, h_host_ids: *const i32, h_device_ids: *const i32, n: i32) -> c_int;
    pub fn kvdp_host_tier_sync(t: *mut KvdpHostTier) -> c_int;
    pub fn kvdp_hash_workspace_bytes(n_req: i64) -> usize;
    pub fn kvdp_hash_token_blocks(d_tokens: *const u32, d_offsets: *const i64, n_req: i64, total_tokens: i64, block_size: i32, salt: u64, d_block_offsets: *mut i64, d_block_hash: *mut u64, d_seq_hash: *mut u64, d_workspace: *mut c_void, workspace_bytes: usize, flags: u32, stream: *mut c_void) -> c_int;
    pub fn kvdp_hash_token_blocks_noalloc(d_tokens: *const u32, d_offsets: *const i64, n_req: i64, total_tokens: i64, block_size: i32, salt: u64, d_block_offsets: *mut i64, d_block_hash: *mut u64, d_seq_hash: *mut u64, flags: u32, stream: *mut c_void) -> c_int;
    pub fn kvdp_hash_token_blocks_host(h_tokens: *const u32, h_offsets: *const i64, n_req: i64, block_size: i32, salt: u64, h_block_offsets: *mut i64, h_block_hash: *mut u64, h_seq_hash: *mut u64, flags: u32, device: i32) -> c_int;
    pub fn kvdp_xxh3_64_batch(d_data: *const u8, n: i64, len: i64, seed: u64, d_out: *mut u64, stream: *mut c_void) -> c_int;
    pub fn kvdp_index_create(out: *mut *mut KvdpIndex, max_nodes: i64, max_workers: i32, device: i32) -> c_int;
    pub fn kvdp_index_destroy(ix: *mut KvdpIndex) -> c_int;
    pub fn kvdp_index_stream(ix: *mut KvdpIndex) -> *mut c_void;
    pub fn kvdp_index_apply_stored(ix: *mut KvdpIndex, worker: i32, has_parent: i32, parent_hash: u64, block_hashes: *const u64, tokens_hashes: *const u64, n: i32) -> c_int;
    pub fn kvdp_index_apply_removed(ix: *mut KvdpIndex, worker: i32, block_hashes: *const u64, n: i32) -> c_int;
    pub fn kvdp_index_remove_worker(ix: *mut KvdpIndex, worker: i32) -> c_int;
    pub fn kvdp_index_flush(ix: *mut KvdpIndex) -> c_int;
    pub fn kvdp_index_find_matches(ix: *mut KvdpIndex, d_hashes: *const u64, d_seq_offsets: *const i64, n_req: i64, early_exit: i32, num_workers: i32, d_scores: *mut u32, d_matched_depth: *mut i32, stream: *mut c_void) -> c_int;
    pub fn kvdp_index_find_matches_host(ix: *mut KvdpIndex, h_hashes: *const u64, h_seq_offsets: *const i64, n_req: i64, early_exit: i32, num_workers: i32, h_scores: *mut u32, h_matched_depth: *mut i32) -> c_int;
    pub fn kvdp_index_find_matches_for_tokens_host(ix: *mut KvdpIndex, h_tokens: *const u32, h_offsets: *const i64, n_req: i64, block_size: i32, early_exit: i32, num_workers: i32, h_scores: *mut u32, h_matched_depth: *mut i32) -> c_int;
    pub fn kvdp_index_touch(ix: *mut KvdpIndex, h_hashes: *const u64, n: i64, early_exit: i32, now_ns: i64, expiration_ns: i64, h_frequencies: *mut u64, cap: i64) -> i64;
    pub fn kvdp_index_stat(ix: *mut KvdpIndex, what: i32, arg: u64) -> i64;
    // §I engine-facing KV-event ABI (same symbols as libdynamo_llm_capi.so)
    pub fn dynamo_llm_init(namespace_c_str: *const c_char, component_c_str: *const c_char, worker_id: i64, kv_block_size: u32) -> u32;
    pub fn dynamo_llm_shutdown() -> u32;
    pub fn dynamo_llm_load_publisher_create() -> u32;
    pub fn dynamo_kv_event_publish_stored(event_id: u64, token_ids: *const u32, num_block_tokens: *const usize, block_ids: *const u64, num_blocks: usize, parent_hash: *const u64, lora_id: u64) -> u32;
    pub fn dynamo_kv_event_publish_removed(event_id: u64, block_ids: *const u64, num_blocks: usize) -> u32;
    pub fn kvdp_event_sink_callback(cb: Option<unsafe extern "C" fn(user: *mut c_void, subject: *const c_char, json: *const c_char, len: usize)>, user: *mut c_void) -> c_int;
    pub fn kvdp_event_sink_index(ix: *mut KvdpIndex, worker_slot: i32) -> c_int;
    pub fn kvdp_event_flush() -> c_int;
    pub fn kvdp_event_poll(buf: *mut c_char, cap: usize, len: *mut usize) -> c_int;
    pub fn kvdp_event_stat(what: i32) -> i64;
}
