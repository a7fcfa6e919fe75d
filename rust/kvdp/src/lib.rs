//! Safe wrappers over `kvdp-sys` for the batched v2 calls. Source only: never compiled in the build image.
//!
//! Where they plug into dynamo-llm (paths relative to the reference tree):
//!   * `CopyPlan::block_copy`  — once per block-set pair, where `CopyStreamBlockMap::new` is built (kv/layer.rs:648)
//!   * `TransferEngine::run`   — instead of prepare_block_ids + trigger_all_layers (kv/layer.rs:836-968)
//!   * `hash_token_blocks`     — instead of `TokenBlockSequence::split_tokens` per request (lib/tokens/src/lib.rs:369)
use kvdp_sys as sys;
use std::ffi::CStr;
use std::ptr;

#[derive(Debug)]
pub struct Error {
    pub code: i32,
    pub message: String,
}

fn check(rc: i32) -> Result<(), Error> {
    if rc == 0 {
        return Ok(());
    }
    let message = unsafe { CStr::from_ptr(sys::kvdp_last_error()) }.to_string_lossy().into_owned();
    Err(Error { code: rc, message })
}

pub struct CopyPlan(*mut sys::KvdpPlan);
unsafe impl Send for CopyPlan {}
unsafe impl Sync for CopyPlan {}

impl CopyPlan {
    /// Same arguments as `CopyStreamBlockMap` carries (kv/layer.rs:592-612).
    #[allow(clippy::too_many_arguments)]
    pub fn block_copy(
        src_layer_ptrs: &[u64], dst_layer_ptrs: &[u64], prefix_dim: i32, src_block_dim: i64, dst_block_dim: i64,
        suffix_dim: i64, elem_size: i32, device: i32,
    ) -> Result<Self, Error> {
        assert_eq!(src_layer_ptrs.len(), dst_layer_ptrs.len());
        let mut p = ptr::null_mut();
        check(unsafe {
            sys::kvdp_plan_block_copy(&mut p, src_layer_ptrs.as_ptr(), dst_layer_ptrs.as_ptr(),
                src_layer_ptrs.len() as i32, prefix_dim, src_block_dim, dst_block_dim, suffix_dim, elem_size, device)
        })?;
        Ok(Self(p))
    }

    /// Layer-separate <-> contiguous-per-block relayout (layout 0 = KvFirst, 1 = BlockFirst).
    #[allow(clippy::too_many_arguments)]
    pub fn contig(
        layer_ptrs: &[u64], layout: i32, layer_blocks: i64, chunk_bytes: i64, contig_base: u64, contig_blocks: i64,
        to_contig: bool, device: i32,
    ) -> Result<Self, Error> {
        let mut p = ptr::null_mut();
        check(unsafe {
            sys::kvdp_plan_contig(&mut p, layer_ptrs.as_ptr(), layer_ptrs.len() as i32, layout, layer_blocks,
                chunk_bytes, contig_base as *mut _, contig_blocks, to_contig as i32, device)
        })?;
        Ok(Self(p))
    }

    /// TP reshard push of prefill rank `src_rank`; `dst_layer_ptrs` is `[targets][layers]` flattened.
    #[allow(clippy::too_many_arguments)]
    pub fn reshard(
        src_layer_ptrs: &[u64], dst_layer_ptrs: &[u64], src_blocks: i64, dst_blocks: i64, block_size: i32,
        src_row_bytes: i64, src_tp: i32, dst_tp: i32, src_rank: i32, device: i32,
    ) -> Result<Self, Error> {
        let mut p = ptr::null_mut();
        check(unsafe {
            sys::kvdp_plan_reshard(&mut p, src_layer_ptrs.as_ptr(), dst_layer_ptrs.as_ptr(),
                src_layer_ptrs.len() as i32, src_blocks, dst_blocks, block_size, src_row_bytes, src_tp, dst_tp,
                src_rank, device)
        })?;
        Ok(Self(p))
    }
}

impl Drop for CopyPlan {
    fn drop(&mut self) {
        unsafe { sys::kvdp_plan_destroy(self.0) };
    }
}

pub struct TransferEngine(*mut sys::KvdpXfer);
unsafe impl Send for TransferEngine {}

impl TransferEngine {
    pub fn new(max_pairs: i32, device: i32) -> Result<Self, Error> {
        let mut p = ptr::null_mut();
        check(unsafe { sys::kvdp_xfer_create(&mut p, max_pairs, device) })?;
        Ok(Self(p))
    }

    /// One launch for layers `[layer_begin, layer_end)` of every (src, dst) pair; asynchronous.
    pub fn run(
        &mut self, plan: &CopyPlan, src_ids: &[i32], dst_ids: &[i32], layer_begin: i32, layer_end: i32,
        check_dst_unique: bool, done_flag: Option<(*mut u32, u32)>,
    ) -> Result<(), Error> {
        assert_eq!(src_ids.len(), dst_ids.len());
        let (flag, value) = done_flag.unwrap_or((ptr::null_mut(), 0));
        check(unsafe {
            sys::kvdp_xfer_run(self.0, plan.0, src_ids.as_ptr(), dst_ids.as_ptr(), src_ids.len() as i32, layer_begin,
                layer_end, if check_dst_unique { sys::KVDP_XFER_CHECK_DST_UNIQUE } else { 0 }, flag, value)
        })
    }

    pub fn sync(&mut self) -> Result<(), Error> {
        check(unsafe { sys::kvdp_xfer_sync(self.0) })
    }
}

impl Drop for TransferEngine {
    fn drop(&mut self) {
        unsafe { sys::kvdp_xfer_destroy(self.0) };
    }
}

/// Batched `split_tokens`: returns (block_offsets, block_hash, sequence_hash) for CSR `offsets` over `tokens`.
pub fn hash_token_blocks(
    tokens: &[u32], offsets: &[i64], block_size: i32, salt: u64, device: i32,
) -> Result<(Vec<i64>, Vec<u64>, Vec<u64>), Error> {
    let n_req = offsets.len() as i64 - 1;
    let total: i64 = offsets.windows(2).map(|w| (w[1] - w[0]) / block_size as i64).sum();
    let mut boff = vec![0i64; offsets.len()];
    let mut bh = vec![0u64; total as usize];
    let mut sh = vec![0u64; total as usize];
    check(unsafe {
        sys::kvdp_hash_token_blocks_host(tokens.as_ptr(), offsets.as_ptr(), n_req, block_size, salt, boff.as_mut_ptr(),
            bh.as_mut_ptr(), sh.as_mut_ptr(), 0, device)
    })?;
    Ok((boff, bh, sh))
}
