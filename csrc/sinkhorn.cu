// Sinkhorn-Knopp assignment for SwAV in ONE persistent cooperative kernel (reference:
// vissl/losses/swav_loss.py:177-244 = ~25 small torch kernels with host-synchronising nonzero() guards).
//
// The reference repeatedly rescales Q = exp((scores - max) / eps)^T in place. Every rescaling is a row or column
// scaling, so Q_t = diag(a) E diag(b) with E = exp(scores / eps - max) fixed. Tracking only the two vectors:
//     a = r / (E b)        r = 1 / K            (row = prototype normalisation)
//     b = c / (E^T a)      c = 1 / (n * world)  (column = sample normalisation)
// starting from b = 1 / sum(E), and the final column normalisation gives  out[j, i] = a_i E[j, i] / (E^T a)_j.
// The kernel therefore reads the score matrix 2 + iters times (it stays in the 126 MB L2: 4040 x 3000 bf16 is
// 24 MB), never writes an intermediate matrix, needs one grid barrier per pass and no host synchronisation.
// Guards of the reference (inf entries, zero row sums) become clamps, which are identical on finite inputs.
//
// Work split: CTA owns rows (samples) j = blockIdx.x, + gridDim.x, ...; a thread owns columns (prototypes)
// tid, tid + 256, ... so per-column partial sums live in registers and hit global memory once per pass.
#include <cooperative_groups.h>
#include <float.h>
#include "common.cuh"
#include "../include/dedloc_b200.h"

namespace cg = cooperative_groups;

namespace dedloc {

constexpr int kSkThreads = 256;
constexpr int kSkMaxPerThread = 16;  // K <= 4096 prototypes

__device__ __forceinline__ float block_sum1(float v, float* red) {
  float2 r = block_sum2(v, 0.f, red);
  return r.x;
}
__device__ __forceinline__ void atomic_max_float(float* addr, float v) {
  // valid for any sign: compare as ordered ints
  if (v >= 0.f) atomicMax(reinterpret_cast<int*>(addr), __float_as_int(v));
  else atomicMin(reinterpret_cast<unsigned int*>(addr), __float_as_uint(v));
}

template <typename T>
__device__ __forceinline__ float load_score(const T* p) {
  if constexpr (sizeof(T) == 2) return __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(p));
  else return *reinterpret_cast<const float*>(p);
}

template <typename T>
__global__ void __launch_bounds__(kSkThreads) sinkhorn_kernel(const T* __restrict__ scores, long long ld, int n, int K,
                                                              int n_out, float inv_eps, int iters, float world,
                                                              float* __restrict__ out, long long ld_out,
                                                              float* __restrict__ scratch) {
  cg::grid_group grid = cg::this_grid();
  __shared__ float red[64];
  const int tid = threadIdx.x;
  const int per = (K + kSkThreads - 1) / kSkThreads;  // columns per thread (<= kSkMaxPerThread)
  // scratch layout: [0] global max, [1..] (iters + 1) column-sum vectors of length K
  float* gmax = scratch;
  float* colsum = scratch + 32;

  // ---- pass 0: global max of scores / eps (log-sum-exp shift of the reference) ----
  float m = -FLT_MAX;
  for (int j = blockIdx.x; j < n; j += gridDim.x)
    for (int k = 0; k < per; ++k) {
      const int i = tid + k * kSkThreads;
      if (i < K) m = fmaxf(m, load_score(scores + (size_t)j * ld + i) * inv_eps);
    }
  m = warp_max(m);
  if ((tid & 31) == 0) red[tid >> 5] = m;
  __syncthreads();
  if (tid < 32) {
    float v = tid < kSkThreads / 32 ? red[tid] : -FLT_MAX;
    v = warp_max(v);
    if (tid == 0) atomic_max_float(gmax, fminf(v, FLT_MAX));
  }
  __threadfence();
  grid.sync();
  const float M = *reinterpret_cast<volatile float*>(gmax);

  const float r = 1.f / (float)K;
  const float c = 1.f / ((float)n * world);
  float a[kSkMaxPerThread];
#pragma unroll
  for (int k = 0; k < kSkMaxPerThread; ++k) a[k] = 0.f;

  // ---- passes 1 .. iters+1 ----
  // pass p (1-based): T_j = sum_i a_i E_ji (p > 1), b_j = c / T_j, then S_i += E_ji b_j  (p <= iters)
  //                   last pass writes out[j, i] = a_i E_ji / T_j for j < n_out
  for (int p = 1; p <= iters + 1; ++p) {
    float part[kSkMaxPerThread];
#pragma unroll
    for (int k = 0; k < kSkMaxPerThread; ++k) part[k] = 0.f;
    const bool first = (p == 1), last = (p == iters + 1);
    for (int j = blockIdx.x; j < n; j += gridDim.x) {
      if (last && j >= n_out) break;  // only the first n_out rows (the current batch, not the queue) are wanted
      float e[kSkMaxPerThread];
      float dot = 0.f;
#pragma unroll
      for (int k = 0; k < kSkMaxPerThread; ++k) {
        const int i = tid + k * kSkThreads;
        e[k] = 0.f;
        if (k < per && i < K) {
          e[k] = __expf(fminf(load_score(scores + (size_t)j * ld + i) * inv_eps - M, 80.f));
          dot = fmaf(a[k], e[k], dot);
        }
      }
      float bj = 1.f;  // pass 1: b_j = 1 / sum(E) is a scalar folded into `a` below
      float Tj = 1.f;
      if (!first) {
        Tj = fmaxf(block_sum1(dot, red), 1e-37f);
        bj = c / Tj;
      }
      if (!last) {
#pragma unroll
        for (int k = 0; k < kSkMaxPerThread; ++k) part[k] = fmaf(e[k], bj, part[k]);
      } else {
        const float inv = 1.f / Tj;
#pragma unroll
        for (int k = 0; k < kSkMaxPerThread; ++k) {
          const int i = tid + k * kSkThreads;
          if (k < per && i < K) out[(size_t)j * ld_out + i] = a[k] * e[k] * inv;
        }
      }
    }
    if (last) break;
    float* S = colsum + (size_t)(p - 1) * K;
#pragma unroll
    for (int k = 0; k < kSkMaxPerThread; ++k) {
      const int i = tid + k * kSkThreads;
      if (k < per && i < K && part[k] != 0.f) atomicAdd(S + i, part[k]);
    }
    __threadfence();
    grid.sync();
    // a_i = r / S_i ; after pass 1, S is the raw column sum and b = 1/sum(E): a_i = r * sum(E) / S_i
    float total = 1.f;
    if (first) {
      float s = 0.f;
      for (int i = tid; i < K; i += kSkThreads) s += *reinterpret_cast<volatile float*>(S + i);
      total = block_sum1(s, red) * world;  // cross-replica sum of Q in the reference; one replica here
      total = fmaxf(total, 1e-37f);
    }
#pragma unroll
    for (int k = 0; k < kSkMaxPerThread; ++k) {
      const int i = tid + k * kSkThreads;
      if (k < per && i < K) {
        const float Si = fmaxf(*reinterpret_cast<volatile float*>(S + i) * world, 1e-37f);
        a[k] = fminf(first ? r * total / Si : r / Si, FLT_MAX);
      }
    }
  }
}

}  // namespace dedloc

using namespace dedloc;

extern "C" long long dedloc_sinkhorn_scratch_floats(int K, int iters) { return 32 + (long long)(iters + 1) * K; }

extern "C" int dedloc_sinkhorn(const void* scores, int is_bf16, long long ld, int n, int K, int n_out, float epsilon,
                               int iters, float world_size, float* out, long long ld_out, float* scratch,
                               void* stream) {
  DEDLOC_REQUIRE(n > 0 && K > 0 && K <= kSkThreads * kSkMaxPerThread, "sinkhorn: K (%d) must be in [1, %d]", K,
                 kSkThreads * kSkMaxPerThread);
  DEDLOC_REQUIRE(n_out >= 0 && n_out <= n && iters >= 1 && epsilon > 0.f, "sinkhorn: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  // the max slot starts at -FLT_MAX, the column-sum vectors at zero
  DEDLOC_CUDA_CHECK(cudaMemsetAsync(scratch, 0, sizeof(float) * (size_t)dedloc_sinkhorn_scratch_floats(K, iters), st));
  const float neg = -FLT_MAX;
  DEDLOC_CUDA_CHECK(cudaMemcpyAsync(scratch, &neg, sizeof(float), cudaMemcpyHostToDevice, st));
  static int max_blocks_bf16 = 0, max_blocks_f32 = 0;
  int& max_blocks = is_bf16 ? max_blocks_bf16 : max_blocks_f32;
  if (max_blocks == 0) {
    int dev = 0, sms = 0, occ = 0;
    DEDLOC_CUDA_CHECK(cudaGetDevice(&dev));
    DEDLOC_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (is_bf16)
      DEDLOC_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sinkhorn_kernel<__nv_bfloat16>, kSkThreads, 0));
    else
      DEDLOC_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sinkhorn_kernel<float>, kSkThreads, 0));
    DEDLOC_REQUIRE(occ >= 1, "sinkhorn: kernel does not fit on an SM");
    max_blocks = sms * (occ > 4 ? 4 : occ);
  }
  int blocks = n < max_blocks ? n : max_blocks;
  float inv_eps = 1.f / epsilon;
  void* args[] = {(void*)&scores, (void*)&ld, (void*)&n, (void*)&K, (void*)&n_out, (void*)&inv_eps, (void*)&iters,
                  (void*)&world_size, (void*)&out, (void*)&ld_out, (void*)&scratch};
  if (is_bf16)
    DEDLOC_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)sinkhorn_kernel<__nv_bfloat16>, dim3(blocks),
                                                  dim3(kSkThreads), args, 0, st));
  else
    DEDLOC_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)sinkhorn_kernel<float>, dim3(blocks), dim3(kSkThreads),
                                                  args, 0, st));
  count_launch();
  return 0;
}

// ------------------------------------------------------------------------------------------------------------
// Allocation-free variant of the Sinkhorn entry point.
//
// Who needs it: callers that are not allowed to (or simply cannot) obtain device memory at the moment they need
// assignments -- a host application that handed its whole memory budget to an allocator it does not own, a plug-in
// running under a fixed-footprint contract, or a latency-critical loop where cudaMalloc's device-wide
// synchronisation is not acceptable. dedloc_sinkhorn() makes them bring `scratch`; this entry point brings its own:
// a file-scope device array that every GPU loading the library holds until the process exits (128.1 KiB per GPU).
//
// Contract:
//   * capacity: 32 + (iters + 1) * K floats must fit in kSkNoAllocFloats (e.g. K = 4096 with iters <= 7, or
//     K = 3000 with iters <= 9); larger requests are rejected with DEDLOC_EINVAL and a message;
//   * the FIRST call on each device performs one-time set-up (it looks up the device address of the array for the
//     device that is current, and queries occupancy). Make that first call before starting any stream capture;
//   * nothing is allocated or freed: no cudaMalloc / cudaFree, no allocating library routine;
//   * the array is shared by every call and keeps the previous call's contents, so each call first resets the part
//     it uses (max slot to -FLT_MAX, column sums to zero) with a small kernel on the same stream. Two calls whose
//     GPU work overlaps (different streams, or different host threads) race on the array: serialise them.
// ------------------------------------------------------------------------------------------------------------
#include <mutex>

namespace dedloc {

constexpr int kSkNoAllocFloats = 32 + 8 * 4096;   // 32 800 floats
static __device__ __align__(256) float g_sk_noalloc_scratch[kSkNoAllocFloats];

__global__ void sinkhorn_noalloc_reset(float* scratch, int nfloats) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nfloats; i += gridDim.x * blockDim.x)
    scratch[i] = (i == 0) ? -FLT_MAX : 0.f;
}

// In host code `g_sk_noalloc_scratch` is only a handle for the runtime, not a device address: the address valid on
// the CURRENT device comes from cudaGetSymbolAddress (looked up once per device, with that device current).
static int sk_noalloc_scratch(float** out) {
  constexpr int kMaxDevices = 64;
  static float* ptrs[kMaxDevices] = {};
  static std::mutex mu;
  int dev = 0;
  DEDLOC_CUDA_CHECK(cudaGetDevice(&dev));
  DEDLOC_REQUIRE(dev >= 0 && dev < kMaxDevices, "sinkhorn_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    float* p = nullptr;
    DEDLOC_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_sk_noalloc_scratch));
    DEDLOC_REQUIRE(p != nullptr, "sinkhorn_noalloc: cudaGetSymbolAddress returned a null address");
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return 0;
}

}  // namespace dedloc

extern "C" int dedloc_sinkhorn_noalloc_capacity_floats(void) { return dedloc::kSkNoAllocFloats; }

extern "C" int dedloc_sinkhorn_noalloc(const void* scores, int is_bf16, long long ld, int n, int K, int n_out,
                                       float epsilon, int iters, float world_size, float* out, long long ld_out,
                                       void* stream) {
  using namespace dedloc;
  DEDLOC_REQUIRE(n > 0 && K > 0 && K <= kSkThreads * kSkMaxPerThread, "sinkhorn_noalloc: K (%d) must be in [1, %d]", K,
                 kSkThreads * kSkMaxPerThread);
  DEDLOC_REQUIRE(n_out >= 0 && n_out <= n && iters >= 1 && epsilon > 0.f, "sinkhorn_noalloc: bad arguments");
  const long long need = 32 + (long long)(iters + 1) * K;
  DEDLOC_REQUIRE(need <= kSkNoAllocFloats,
                 "sinkhorn_noalloc: K = %d with %d iterations needs %lld scratch floats but the built-in array holds "
                 "%d; use dedloc_sinkhorn() with caller-provided scratch for this size",
                 K, iters, need, kSkNoAllocFloats);
  float* scratch = nullptr;
  { int rc = sk_noalloc_scratch(&scratch); if (rc) return rc; }
  cudaStream_t st = (cudaStream_t)stream;
  sinkhorn_noalloc_reset<<<8, 256, 0, st>>>(scratch, (int)need);
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  static int max_blocks_bf16 = 0, max_blocks_f32 = 0;
  int& max_blocks = is_bf16 ? max_blocks_bf16 : max_blocks_f32;
  if (max_blocks == 0) {
    int dev = 0, sms = 0, occ = 0;
    DEDLOC_CUDA_CHECK(cudaGetDevice(&dev));
    DEDLOC_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (is_bf16)
      DEDLOC_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sinkhorn_kernel<__nv_bfloat16>, kSkThreads, 0));
    else
      DEDLOC_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sinkhorn_kernel<float>, kSkThreads, 0));
    DEDLOC_REQUIRE(occ >= 1, "sinkhorn_noalloc: kernel does not fit on an SM");
    max_blocks = sms * (occ > 4 ? 4 : occ);
  }
  const int blocks = n < max_blocks ? n : max_blocks;
  float inv_eps = 1.f / epsilon;
  void* args[] = {(void*)&scores, (void*)&ld, (void*)&n, (void*)&K, (void*)&n_out, (void*)&inv_eps, (void*)&iters,
                  (void*)&world_size, (void*)&out, (void*)&ld_out, (void*)&scratch};
  if (is_bf16)
    DEDLOC_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)sinkhorn_kernel<__nv_bfloat16>, dim3(blocks),
                                                  dim3(kSkThreads), args, 0, st));
  else
    DEDLOC_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)sinkhorn_kernel<float>, dim3(blocks), dim3(kSkThreads),
                                                  args, 0, st));
  count_launch(2);
  return 0;
}
