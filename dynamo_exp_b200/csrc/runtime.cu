// runtime.cu — error plumbing, options, device slabs, peer/IPC mapping, completion flags.
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <unordered_map>

#include "common.cuh"

namespace kvdp {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return code;
}

Options& options() {
  static Options o;
  return o;
}

int device_info(int device, DevInfo* out) {
  static std::mutex mu;
  static std::unordered_map<int, DevInfo> cache;
  std::lock_guard<std::mutex> lk(mu);
  auto it = cache.find(device);
  if (it != cache.end()) {
    *out = it->second;
    return 0;
  }
  DevInfo d;
  KVDP_CUDA(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, device));
  KVDP_CUDA(cudaDeviceGetAttribute(&d.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
  cache[device] = d;
  *out = d;
  return 0;
}

namespace {
__global__ void kvdp_signal_kernel(uint32_t* flag, uint32_t value) {
  __threadfence_system();
  st_release_sys_u32(flag, value);
}
}  // namespace

// result slot of kvdp_flag_cas64 (one per device that loads the library; calls are serialised by a mutex)
__device__ unsigned long long g_cas_old;
namespace {
__global__ void kvdp_cas64_kernel(unsigned long long* addr, unsigned long long expected, unsigned long long desired) {
  g_cas_old = atomicCAS_system(addr, expected, desired);
}
}  // namespace

int enqueue_signal(uint32_t* flag, uint32_t value, cudaStream_t stream) {
  kvdp_signal_kernel<<<1, 1, 0, stream>>>(flag, value);
  KVDP_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace kvdp

using namespace kvdp;

extern "C" {

const char* kvdp_last_error(void) { return g_err; }

int kvdp_version(void) { return 200; /* 0.2.0: tracks the reference workspace version */ }

int kvdp_set_option(const char* name, int64_t value) {
  if (!name) return fail(KVDP_EINVAL, "kvdp_set_option: NULL name");
  Options& o = options();
  if (!strcmp(name, "copy_path")) {
    if (value < 0 || value > 4) return fail(KVDP_EINVAL, "copy_path must be 0..4");
    o.copy_path = value;
  } else if (!strcmp(name, "tile_bytes")) {
    if (value < 0 || value > 200 * 1024) return fail(KVDP_EINVAL, "tile_bytes out of range");
    o.tile_bytes = value;
  } else if (!strcmp(name, "stages")) {
    if (value < 0 || value > 8) return fail(KVDP_EINVAL, "stages must be 0..8");
    o.stages = value;
  } else if (!strcmp(name, "ctas_per_sm")) {
    if (value < 0 || value > 32) return fail(KVDP_EINVAL, "ctas_per_sm must be 0..32");
    o.ctas_per_sm = value;
  } else if (!strcmp(name, "host_path")) {
    if (value < 0 || value > 3) return fail(KVDP_EINVAL, "host_path must be 0..3");
    o.host_path = value;
  } else if (!strcmp(name, "l2_hint")) {
    if (value < 0 || value > 3) return fail(KVDP_EINVAL, "l2_hint must be 0..3");
    o.l2_hint = value;
  } else if (!strcmp(name, "vec32")) {
    if (value < 0 || value > 1) return fail(KVDP_EINVAL, "vec32 must be 0 or 1");
    o.vec32 = value;
  } else if (!strcmp(name, "hash_stages")) {
    if (value < 0 || value > 16) return fail(KVDP_EINVAL, "hash_stages must be 0..16");
    o.hash_stages = value;
  } else if (!strcmp(name, "host_ctas")) {
    if (value < 0 || value > 4096) return fail(KVDP_EINVAL, "host_ctas out of range");
    o.host_ctas = value;
  } else if (!strcmp(name, "host_dma_piece_kib")) {
    if (value < 0 || value > (1 << 20)) return fail(KVDP_EINVAL, "host_dma_piece_kib must be 0..1048576");
    o.host_dma_piece_kib = value;
  } else {
    return fail(KVDP_EINVAL, "unknown option '%s'", name);
  }
  return 0;
}

int64_t kvdp_get_option(const char* name) {
  if (!name) return -1;
  const Options& o = options();
  if (!strcmp(name, "copy_path")) return o.copy_path;
  if (!strcmp(name, "tile_bytes")) return o.tile_bytes;
  if (!strcmp(name, "stages")) return o.stages;
  if (!strcmp(name, "ctas_per_sm")) return o.ctas_per_sm;
  if (!strcmp(name, "host_path")) return o.host_path;
  if (!strcmp(name, "host_ctas")) return o.host_ctas;
  if (!strcmp(name, "host_dma_piece_kib")) return o.host_dma_piece_kib;
  if (!strcmp(name, "l2_hint")) return o.l2_hint;
  if (!strcmp(name, "vec32")) return o.vec32;
  if (!strcmp(name, "hash_stages")) return o.hash_stages;
  return -1;
}

int kvdp_device_count(int* count) {
  if (!count) return fail(KVDP_EINVAL, "kvdp_device_count: NULL out");
  cudaError_t e = cudaGetDeviceCount(count);
  if (e != cudaSuccess) {
    *count = 0;
    return fail(KVDP_ENODEV, "no CUDA device: %s", cudaGetErrorString(e));
  }
  return 0;
}

int kvdp_device_alloc(void** ptr, size_t bytes, int device) {
  if (!ptr) return fail(KVDP_EINVAL, "kvdp_device_alloc: NULL out");
  *ptr = nullptr;
  DeviceGuard g(device);
  KVDP_CUDA(cudaMalloc(ptr, bytes ? bytes : 1));
  KVDP_CUDA(cudaMemset(*ptr, 0, bytes));
  return 0;
}

int kvdp_device_free(void* ptr) {
  if (!ptr) return 0;
  KVDP_CUDA(cudaFree(ptr));
  return 0;
}

int kvdp_memset(void* ptr, int value, size_t bytes) {
  KVDP_CUDA(cudaMemset(ptr, value, bytes));
  return 0;
}

int kvdp_pointer_kind(const void* ptr, int* kind, int* device) {
  if (!kind) return fail(KVDP_EINVAL, "kvdp_pointer_kind: NULL out");
  cudaPointerAttributes attr;
  KVDP_CUDA(cudaPointerGetAttributes(&attr, ptr));
  switch (attr.type) {
    case cudaMemoryTypeHost: *kind = 1; break;
    case cudaMemoryTypeDevice: *kind = 2; break;
    case cudaMemoryTypeManaged: *kind = 3; break;
    default: *kind = 0; break;
  }
  if (device) *device = (*kind >= 2) ? attr.device : -1;
  return 0;
}

int kvdp_peer_enable(int32_t device, int32_t peer_device) {
  if (device == peer_device) return 0;
  int can = 0;
  KVDP_CUDA(cudaDeviceCanAccessPeer(&can, device, peer_device));
  if (!can) return fail(KVDP_ENODEV, "device %d cannot access peer %d", device, peer_device);
  DeviceGuard g(device);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  KVDP_CUDA(e);
  return 0;
}

int kvdp_ipc_export(void* device_ptr, uint8_t handle[KVDP_IPC_HANDLE_BYTES]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == KVDP_IPC_HANDLE_BYTES, "IPC handle size");
  if (!device_ptr || !handle) return fail(KVDP_EINVAL, "kvdp_ipc_export: NULL argument");
  cudaIpcMemHandle_t h;
  KVDP_CUDA(cudaIpcGetMemHandle(&h, device_ptr));
  memcpy(handle, &h, sizeof h);
  return 0;
}

int kvdp_ipc_export_range(void* device_ptr, uint8_t handle[KVDP_IPC_HANDLE_BYTES], uint64_t* offset) {
  if (!device_ptr || !handle || !offset) return fail(KVDP_EINVAL, "kvdp_ipc_export_range: NULL argument");
  typedef CUresult (*range_fn_t)(CUdeviceptr*, size_t*, CUdeviceptr);
  static range_fn_t range_fn = nullptr;
  if (!range_fn) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    KVDP_CUDA(cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &q));
    if (!fn || q != cudaDriverEntryPointSuccess) return fail(KVDP_ENODEV, "cuMemGetAddressRange is not available");
    range_fn = (range_fn_t)fn;
  }
  CUdeviceptr base = 0;
  size_t size = 0;
  CUresult r = range_fn(&base, &size, (CUdeviceptr)(uintptr_t)device_ptr);
  if (r != CUDA_SUCCESS) return fail(KVDP_EINVAL, "cuMemGetAddressRange failed with CUresult %d", (int)r);
  cudaIpcMemHandle_t h;
  KVDP_CUDA(cudaIpcGetMemHandle(&h, (void*)(uintptr_t)base));
  memcpy(handle, &h, sizeof h);
  *offset = (uint64_t)((uintptr_t)device_ptr - (uintptr_t)base);
  return 0;
}

int kvdp_ipc_import(const uint8_t handle[KVDP_IPC_HANDLE_BYTES], int32_t device, void** mapped_ptr) {
  if (!handle || !mapped_ptr) return fail(KVDP_EINVAL, "kvdp_ipc_import: NULL argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof h);
  DeviceGuard g(device);
  KVDP_CUDA(cudaIpcOpenMemHandle(mapped_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int kvdp_ipc_close(void* mapped_ptr) {
  if (!mapped_ptr) return 0;
  KVDP_CUDA(cudaIpcCloseMemHandle(mapped_ptr));
  return 0;
}

int kvdp_flag_cas64(uint64_t* addr, uint64_t expected, uint64_t desired, uint64_t* old, int32_t device) {
  if (!addr || !old) return fail(KVDP_EINVAL, "kvdp_flag_cas64: NULL argument");
  if (((uintptr_t)addr & 7) != 0) return fail(KVDP_EALIGN, "kvdp_flag_cas64: address must be 8-byte aligned");
  static std::mutex mu;
  std::lock_guard<std::mutex> lk(mu);
  DeviceGuard g(device);
  kvdp_cas64_kernel<<<1, 1>>>((unsigned long long*)addr, (unsigned long long)expected, (unsigned long long)desired);
  KVDP_CUDA(cudaGetLastError());
  unsigned long long found = 0;
  KVDP_CUDA(cudaMemcpyFromSymbol(&found, g_cas_old, sizeof found));  // legacy default stream: ordered after the kernel
  *old = (uint64_t)found;
  return 0;
}

int kvdp_host_register(void* ptr, size_t bytes) {
  if (!ptr || bytes == 0) return fail(KVDP_EINVAL, "kvdp_host_register: NULL pointer or zero size");
  cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail((int)e, "kvdp_host_register(%p, %zu): %s", ptr, bytes, cudaGetErrorString(e));
  }
  return 0;
}

int kvdp_host_unregister(void* ptr) {
  if (!ptr) return fail(KVDP_EINVAL, "kvdp_host_unregister: NULL pointer");
  cudaError_t e = cudaHostUnregister(ptr);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail((int)e, "kvdp_host_unregister(%p): %s", ptr, cudaGetErrorString(e));
  }
  return 0;
}

int kvdp_stream_synchronize(kvdp_stream_t stream) {
  KVDP_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return 0;
}

int kvdp_stream_wait_flag(kvdp_stream_t stream, uint32_t* flag, uint32_t value) {
  if (!flag) return fail(KVDP_EINVAL, "kvdp_stream_wait_flag: NULL flag");
  // resolved at run time so that libkvdp.so has no link-time dependency on libcuda.so.1 (it must load on a
  // CPU-only build box for the symbol checks)
  typedef CUresult (*wait_fn_t)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
  static wait_fn_t wait_fn = nullptr;
  if (!wait_fn) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    KVDP_CUDA(cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &q));
    if (!fn || q != cudaDriverEntryPointSuccess) return fail(KVDP_ENODEV, "cuStreamWaitValue32 is not available");
    wait_fn = (wait_fn_t)fn;
  }
  CUresult r = wait_fn((CUstream)stream, (CUdeviceptr)(uintptr_t)flag, value, CU_STREAM_WAIT_VALUE_GEQ);
  if (r != CUDA_SUCCESS) return fail(KVDP_EINVAL, "cuStreamWaitValue32 failed with CUresult %d", (int)r);
  return 0;
}

}  // extern "C"
