// hash_static.cu — token-block hashing for callers that are not allowed to allocate device memory.
//
// Who needs this: a process that has already handed every byte of HBM to something else and must not disturb it.
// The typical case is a router or event-publisher thread living inside an inference engine whose memory manager
// claimed the whole device for the KV cache at start-up (a later cudaMalloc fails or fragments the pool), and code
// that runs where an allocation would be a synchronisation point it cannot afford (cudaMalloc serialises the device).
// Such a caller can still pass its input and output buffers, but it has nowhere to get the scan scratch that
// kvdp_hash_token_blocks() asks for. This entry point carries that scratch inside the library image instead.
//
// The scratch is a file-scope __device__ array, so every GPU that loads libkvdp.so holds kStaticScratchBytes (64 KiB)
// of it from load until exit, whether or not this entry point is ever called.
#include <mutex>

#include "common.cuh"

constexpr size_t kStaticScratchBytes = 64 * 1024;

// Scan partials for the block-count prefix sum (one int64 per 1024 requests). In HOST code this name is only a
// stand-in symbol, not an address a kernel can use; the device address is looked up per device below.
static __device__ __align__(256) unsigned char g_hash_scan_scratch[kStaticScratchBytes];

namespace {

constexpr int kMaxDevices = 64;

std::mutex g_mu;
void* g_scratch_addr[kMaxDevices] = {};

// Device address of the scratch on the CURRENT device; resolved once per device, with that device current.
int scratch_on_current_device(void** out) {
  int dev = -1;
  KVDP_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return kvdp::fail(KVDP_EINVAL, "device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_scratch_addr[dev] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_hash_scan_scratch);
    if (e != cudaSuccess || p == nullptr)
      return kvdp::fail(e != cudaSuccess ? (int)e : KVDP_ENODEV, "cudaGetSymbolAddress(static hash scratch) on device %d: %s",
                        dev, cudaGetErrorString(e));
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return 0;
}

}  // namespace

extern "C" int kvdp_hash_token_blocks_noalloc(const uint32_t* d_tokens, const int64_t* d_offsets, int64_t n_req,
                                              int64_t total_tokens, int32_t block_size, uint64_t salt,
                                              int64_t* d_block_offsets, uint64_t* d_block_hash, uint64_t* d_seq_hash,
                                              uint32_t flags, kvdp_stream_t stream) {
  if (n_req > KVDP_HASH_NOALLOC_MAX_REQUESTS)
    return kvdp::fail(KVDP_ERANGE,
                      "kvdp_hash_token_blocks_noalloc: %lld requests in one call, the built-in scratch holds at most %d; "
                      "split the batch or use kvdp_hash_token_blocks with your own workspace",
                      (long long)n_req, KVDP_HASH_NOALLOC_MAX_REQUESTS);
  if (n_req >= 0 && kvdp_hash_workspace_bytes(n_req) > kStaticScratchBytes)
    return kvdp::fail(KVDP_ERANGE, "kvdp_hash_token_blocks_noalloc: scratch need exceeds the built-in %zu bytes",
                      kStaticScratchBytes);
  void* scratch = nullptr;
  if (int rc = scratch_on_current_device(&scratch)) return rc;
  // Same kernels as the allocating-caller entry point; only the origin of the scratch differs. Nothing in the scratch
  // has to start at zero: every partial sum a call reads was written earlier in that same call, so whatever the
  // previous call left behind is never observed.
  // (the single-stream form: the side streams the allocating entry point may create on a big batch are not used here)
  return kvdp::hash_token_blocks_single_stream(d_tokens, d_offsets, n_req, total_tokens, block_size, salt,
                                               d_block_offsets, d_block_hash, d_seq_hash, scratch, kStaticScratchBytes,
                                               flags, (cudaStream_t)stream);
}
