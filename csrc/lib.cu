// Library-level plumbing: thread-local error text, launch counter, device info, symmetric memory over
// CUDA IPC (the B200-native replacement for hivemind's CPU shared-memory tensors + gRPC streams,
// SURVEY 1.2 / 5.8), and thin async copy wrappers so Python never needs a second CUDA binding.
#include <stdarg.h>
#include <atomic>
#include <string.h>
#include "common.cuh"
#include "../include/dedloc_b200.h"

namespace dedloc {
static thread_local char g_err[1024] = "";
static std::atomic<long long> g_launches{0};

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace dedloc

using namespace dedloc;

extern "C" const char* dedloc_last_error(void) { return g_err; }
extern "C" long long dedloc_launch_count(void) { return g_launches.load(); }
// Kernels replayed from a CUDA graph never pass through the entry points; the owner of the graph reports them.
extern "C" void dedloc_add_launch_count(long long n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

extern "C" int dedloc_device_info(int* sm_count, int* cc_major, int* cc_minor, long long* total_mem) {
  int dev = 0;
  DEDLOC_CUDA_CHECK(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  DEDLOC_CUDA_CHECK(cudaGetDeviceProperties(&prop, dev));
  if (sm_count) *sm_count = prop.multiProcessorCount;
  if (cc_major) *cc_major = prop.major;
  if (cc_minor) *cc_minor = prop.minor;
  if (total_mem) *total_mem = (long long)prop.totalGlobalMem;
  return 0;
}

static_assert(sizeof(cudaIpcMemHandle_t) == DEDLOC_IPC_HANDLE_BYTES, "IPC handle size");

extern "C" int dedloc_symm_alloc(long long bytes, void** ptr, unsigned char* handle) {
  DEDLOC_REQUIRE(bytes > 0 && ptr != nullptr, "symm_alloc: bad arguments");
  void* p = nullptr;
  DEDLOC_CUDA_CHECK(cudaMalloc(&p, (size_t)bytes));
  DEDLOC_CUDA_CHECK(cudaMemset(p, 0, (size_t)bytes));
  if (handle) {
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
      // IPC can be unavailable (e.g. restricted containers); single-process use still works.
      memset(handle, 0, DEDLOC_IPC_HANDLE_BYTES);
      (void)cudaGetLastError();
    } else {
      memcpy(handle, &h, DEDLOC_IPC_HANDLE_BYTES);
    }
  }
  *ptr = p;
  return 0;
}

extern "C" int dedloc_symm_open(const unsigned char* handle, void** ptr) {
  DEDLOC_REQUIRE(handle != nullptr && ptr != nullptr, "symm_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, DEDLOC_IPC_HANDLE_BYTES);
  DEDLOC_CUDA_CHECK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

extern "C" int dedloc_symm_close(void* ptr) {
  DEDLOC_CUDA_CHECK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

extern "C" int dedloc_symm_free(void* ptr) {
  DEDLOC_CUDA_CHECK(cudaFree(ptr));
  return 0;
}

extern "C" int dedloc_enable_peer_access(int peer_device) {
  int dev = 0, can = 0;
  DEDLOC_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev == peer_device) return 0;
  DEDLOC_CUDA_CHECK(cudaDeviceCanAccessPeer(&can, dev, peer_device));
  DEDLOC_REQUIRE(can, "device %d cannot access peer %d", dev, peer_device);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { (void)cudaGetLastError(); return 0; }
  DEDLOC_CUDA_CHECK(e);
  return 0;
}

extern "C" int dedloc_memset_async(void* ptr, int value, long long bytes, void* stream) {
  DEDLOC_CUDA_CHECK(cudaMemsetAsync(ptr, value, (size_t)bytes, (cudaStream_t)stream));
  return 0;
}
extern "C" int dedloc_memcpy_async(void* dst, const void* src, long long bytes, void* stream) {
  DEDLOC_CUDA_CHECK(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDefault, (cudaStream_t)stream));
  return 0;
}
// Strided copy (rows of `width` bytes): pulls the interleaved chunks an owner holds out of its mapped buffer.
extern "C" int dedloc_memcpy2d_async(void* dst, long long dpitch, const void* src, long long spitch, long long width,
                                     long long height, void* stream) {
  DEDLOC_CUDA_CHECK(cudaMemcpy2DAsync(dst, (size_t)dpitch, src, (size_t)spitch, (size_t)width, (size_t)height,
                                      cudaMemcpyDefault, (cudaStream_t)stream));
  return 0;
}
extern "C" int dedloc_memcpy_peer_async(void* dst, int dst_device, const void* src, int src_device,
                                        long long bytes, void* stream) {
  DEDLOC_CUDA_CHECK(cudaMemcpyPeerAsync(dst, dst_device, src, src_device, (size_t)bytes, (cudaStream_t)stream));
  return 0;
}
