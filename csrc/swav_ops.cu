// SwAV head / loss kernels around the tcgen05 GEMMs and the Sinkhorn kernel.
//
//   l2norm128_fwd/bwd   y = x / max(||x||, eps) for rows of 128 (the embedding the prototype GEMM consumes;
//                       vissl/models/heads/swav_prototypes_head.py:100-104). One warp per row.
//   swav_ce_fwd/bwd     the swapped-prediction cross-entropy of vissl/losses/swav_loss.py:275-302 in one pass over
//                       the scores: for every row r = (crop p, sample b)
//                           loss_r = sum over assign crops i != p of  sum_k q_i[b,k] * (s_r[k]/T - lse_r)
//                       (log_softmax is never materialised; the row lives in registers between the max / sum-exp
//                       reduction and the dot products), and its backward
//                           ds_r[k] = g_r * sum_{i != p} (q_i[b,k] - softmax_r[k] * sum_j q_i[b,j]) / T.
//                       One CTA per row; K <= 256 * 16.
#include "common.cuh"
#include "../include/dedloc_b200.h"

namespace dedloc {
namespace swav {

constexpr int kThreads = 256;
constexpr int kMaxPerThread = 16;
constexpr int kMaxAssign = 4;

__device__ __forceinline__ float block_max(float v, float* red) {
  v = warp_max(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float x = (lane < kThreads / 32) ? red[lane] : -INFINITY;
  return warp_max(x);
}
__device__ __forceinline__ float block_sum1(float v, float* red) {
  return block_sum2(v, 0.f, red).x;
}

template <typename T>
__device__ __forceinline__ float ld(const T* p) {
  if constexpr (sizeof(T) == 2) return __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(p));
  else return *p;
}
template <typename T>
__device__ __forceinline__ void st(T* p, float v) {
  if constexpr (sizeof(T) == 2) *reinterpret_cast<__nv_bfloat16*>(p) = __float2bfloat16(v);
  else *p = v;
}

// scores: [num_crops * bs, K] (row pitch ld); q: [n_assign, bs, K] fp32; assign_crop[i] = crop id of assignment i.
// FWD: loss_row[r], lse[r]. BWD: dscores[r, :] (same dtype / pitch as scores), g = upstream gradient * common factor.
template <typename T, bool BWD>
__global__ void __launch_bounds__(kThreads) swav_ce_kernel(const T* __restrict__ scores, long long ld_s,
                                                          const float* __restrict__ q, int n_assign,
                                                          const int* __restrict__ assign_crop, int bs, int K, float inv_t,
                                                          float* __restrict__ loss_row, float* __restrict__ lse_out,
                                                          const float* __restrict__ lse_in, const float* __restrict__ g,
                                                          T* __restrict__ dscores) {
  __shared__ float red[64];
  const int r = blockIdx.x, p = r / bs, b = r % bs, tid = threadIdx.x;
  const T* row = scores + (size_t)r * ld_s;
  float s[kMaxPerThread];
  float mx = -INFINITY;
#pragma unroll
  for (int j = 0; j < kMaxPerThread; ++j) {
    const int k = tid + j * kThreads;
    s[j] = (k < K) ? ld(row + k) * inv_t : -INFINITY;
    mx = fmaxf(mx, s[j]);
  }
  float lse;
  if constexpr (!BWD) {
    mx = block_max(mx, red);
    float se = 0.f;
#pragma unroll
    for (int j = 0; j < kMaxPerThread; ++j) se += (tid + j * kThreads < K) ? __expf(s[j] - mx) : 0.f;
    se = block_sum1(se, red);
    lse = mx + __logf(se);
  } else {
    lse = lse_in[r];
  }
  float total = 0.f;
  float grad[kMaxPerThread];
  if constexpr (BWD) {
#pragma unroll
    for (int j = 0; j < kMaxPerThread; ++j) grad[j] = 0.f;
  }
  for (int i = 0; i < n_assign; ++i) {
    if (assign_crop[i] == p) continue;
    const float* qi = q + ((size_t)i * bs + b) * K;
    float dot = 0.f, qs = 0.f;
    float qv[kMaxPerThread];
#pragma unroll
    for (int j = 0; j < kMaxPerThread; ++j) {
      const int k = tid + j * kThreads;
      qv[j] = (k < K) ? qi[k] : 0.f;
      qs += qv[j];
      if (k < K) dot = fmaf(qv[j], s[j] - lse, dot);
    }
    if constexpr (!BWD) {
      total += dot;
    } else {
      qs = block_sum1(qs, red);
#pragma unroll
      for (int j = 0; j < kMaxPerThread; ++j)
        if (tid + j * kThreads < K) grad[j] += qv[j] - __expf(s[j] - lse) * qs;
    }
  }
  if constexpr (!BWD) {
    total = block_sum1(total, red);
    if (tid == 0) {
      loss_row[r] = total;
      lse_out[r] = lse;
    }
  } else {
    const float c = g[0] * inv_t;
    T* drow = dscores + (size_t)r * ld_s;
#pragma unroll
    for (int j = 0; j < kMaxPerThread; ++j) {
      const int k = tid + j * kThreads;
      if (k < K) st(drow + k, c * grad[j]);
    }
  }
}

template <bool BWD>
__global__ void __launch_bounds__(kThreads) l2norm128_kernel(const __nv_bfloat16* __restrict__ x,
                                                            const __nv_bfloat16* __restrict__ dy,
                                                            __nv_bfloat16* __restrict__ out, int M, float eps) {
  const int lane = threadIdx.x & 31;
  for (int row = blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5); row < M; row += gridDim.x * (kThreads / 32)) {
    const uint2 r = *reinterpret_cast<const uint2*>(x + (size_t)row * 128 + lane * 4);
    const float2 a = unpack_bf16x2(r.x), b = unpack_bf16x2(r.y);
    const float v[4] = {a.x, a.y, b.x, b.y};
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) ss = fmaf(v[i], v[i], ss);
    const float nrm = fmaxf(sqrtf(warp_sum(ss)), eps);
    const float inv = 1.f / nrm;
    float o[4];
    if constexpr (!BWD) {
#pragma unroll
      for (int i = 0; i < 4; ++i) o[i] = v[i] * inv;
    } else {
      // y = x / n  ->  dx = (dy - y (y . dy)) / n
      const uint2 dr = *reinterpret_cast<const uint2*>(dy + (size_t)row * 128 + lane * 4);
      const float2 da = unpack_bf16x2(dr.x), db = unpack_bf16x2(dr.y);
      const float d[4] = {da.x, da.y, db.x, db.y};
      float yd = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) yd = fmaf(v[i] * inv, d[i], yd);
      yd = warp_sum(yd);
#pragma unroll
      for (int i = 0; i < 4; ++i) o[i] = (d[i] - v[i] * inv * yd) * inv;
    }
    *reinterpret_cast<uint2*>(out + (size_t)row * 128 + lane * 4) =
        make_uint2(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]));
  }
}

}  // namespace swav
}  // namespace dedloc

using namespace dedloc;

extern "C" int dedloc_swav_ce_fwd(const void* scores, int is_bf16, long long ld, const float* q, int n_assign,
                                  const int* assign_crop, int num_crops, int bs, int K, float temperature, float* loss_row,
                                  float* lse, void* stream) {
  DEDLOC_REQUIRE(K > 0 && K <= swav::kThreads * swav::kMaxPerThread, "swav_ce: K (%d) exceeds %d", K,
                 swav::kThreads * swav::kMaxPerThread);
  DEDLOC_REQUIRE(n_assign >= 1 && n_assign <= swav::kMaxAssign && num_crops >= 1 && bs >= 1, "swav_ce: bad shape");
  const int rows = num_crops * bs;
  cudaStream_t st = (cudaStream_t)stream;
  if (is_bf16)
    swav::swav_ce_kernel<__nv_bfloat16, false><<<rows, swav::kThreads, 0, st>>>(
        (const __nv_bfloat16*)scores, ld, q, n_assign, assign_crop, bs, K, 1.f / temperature, loss_row, lse, nullptr, nullptr,
        nullptr);
  else
    swav::swav_ce_kernel<float, false><<<rows, swav::kThreads, 0, st>>>((const float*)scores, ld, q, n_assign, assign_crop, bs,
                                                                         K, 1.f / temperature, loss_row, lse, nullptr, nullptr,
                                                                         nullptr);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dedloc_swav_ce_bwd(const void* scores, int is_bf16, long long ld, const float* q, int n_assign,
                                  const int* assign_crop, int num_crops, int bs, int K, float temperature, const float* lse,
                                  const float* grad_scale, void* dscores, void* stream) {
  DEDLOC_REQUIRE(K > 0 && K <= swav::kThreads * swav::kMaxPerThread, "swav_ce: K (%d) exceeds %d", K,
                 swav::kThreads * swav::kMaxPerThread);
  const int rows = num_crops * bs;
  cudaStream_t st = (cudaStream_t)stream;
  if (is_bf16)
    swav::swav_ce_kernel<__nv_bfloat16, true><<<rows, swav::kThreads, 0, st>>>(
        (const __nv_bfloat16*)scores, ld, q, n_assign, assign_crop, bs, K, 1.f / temperature, nullptr, nullptr, lse, grad_scale,
        (__nv_bfloat16*)dscores);
  else
    swav::swav_ce_kernel<float, true><<<rows, swav::kThreads, 0, st>>>((const float*)scores, ld, q, n_assign, assign_crop, bs, K,
                                                                        1.f / temperature, nullptr, nullptr, lse, grad_scale,
                                                                        (float*)dscores);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dedloc_l2norm128_fwd(const void* x, void* y, int M, float eps, void* stream) {
  int blocks = (M + 7) / 8;
  blocks = blocks < 1 ? 1 : (blocks > 1184 ? 1184 : blocks);
  swav::l2norm128_kernel<false><<<blocks, swav::kThreads, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, nullptr,
                                                                                      (__nv_bfloat16*)y, M, eps);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dedloc_l2norm128_bwd(const void* x, const void* dy, void* dx, int M, float eps, void* stream) {
  int blocks = (M + 7) / 8;
  blocks = blocks < 1 ? 1 : (blocks > 1184 ? 1184 : blocks);
  swav::l2norm128_kernel<true><<<blocks, swav::kThreads, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, M, eps);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}
