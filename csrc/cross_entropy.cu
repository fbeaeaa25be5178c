// Masked-LM cross-entropy on bf16 logits: one pass over a [rows, V] logit matrix computes the per-row
// log-sum-exp and loss; the backward pass writes (softmax - onehot) * row_scale as bf16 in one more pass. Replaces
// the fp32 copy of the logits + log_softmax + nll_loss chain (and their backward) -- 3.3 ms -> 0.3 ms per micro-step
// for 4928 x 30000 logits. Behaviour: torch.nn.functional.cross_entropy(logits.float(), labels, reduction="none").
#include "common.cuh"
#include "../include/dedloc_b200.h"

namespace dedloc {

constexpr int kCeThreads = 256;

__device__ __forceinline__ void online_merge(float& m, float& s, float m2, float s2) {
  const float mm = fmaxf(m, m2);
  if (mm == -INFINITY) { m = mm; s = 0.f; return; }
  s = s * __expf(m - mm) + s2 * __expf(m2 - mm);
  m = mm;
}

__global__ void __launch_bounds__(kCeThreads)
ce_fwd_kernel(const __nv_bfloat16* __restrict__ logits, long long ld, const long long* __restrict__ labels, int V,
              float* __restrict__ loss, float* __restrict__ lse_out) {
  const int row = blockIdx.x;
  const __nv_bfloat16* x = logits + (size_t)row * ld;
  float m = -INFINITY, s = 0.f;
  const int nvec = V >> 3;
  for (int i = threadIdx.x; i < nvec; i += kCeThreads) {
    const uint4 u = ld_stream_v4(x + (size_t)i * 8);
    float v[8];
    float2 t;
    t = unpack_bf16x2(u.x); v[0] = t.x; v[1] = t.y;
    t = unpack_bf16x2(u.y); v[2] = t.x; v[3] = t.y;
    t = unpack_bf16x2(u.z); v[4] = t.x; v[5] = t.y;
    t = unpack_bf16x2(u.w); v[6] = t.x; v[7] = t.y;
    float cm = fmaxf(fmaxf(fmaxf(v[0], v[1]), fmaxf(v[2], v[3])), fmaxf(fmaxf(v[4], v[5]), fmaxf(v[6], v[7])));
    const float mm = fmaxf(m, cm);
    if (mm > -INFINITY) {
      float cs = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) cs += __expf(v[k] - mm);
      s = s * __expf(m - mm) + cs;
      m = mm;
    }
  }
  for (int i = (nvec << 3) + threadIdx.x; i < V; i += kCeThreads) online_merge(m, s, __bfloat162float(x[i]), 1.f);
  // block reduction of (m, s)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, o), s2 = __shfl_xor_sync(0xffffffffu, s, o);
    online_merge(m, s, m2, s2);
  }
  __shared__ float sm[kCeThreads / 32], ss[kCeThreads / 32];
  if ((threadIdx.x & 31) == 0) { sm[threadIdx.x >> 5] = m; ss[threadIdx.x >> 5] = s; }
  __syncthreads();
  if (threadIdx.x < 32) {
    m = threadIdx.x < kCeThreads / 32 ? sm[threadIdx.x] : -INFINITY;
    s = threadIdx.x < kCeThreads / 32 ? ss[threadIdx.x] : 0.f;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
      const float m2 = __shfl_xor_sync(0xffffffffu, m, o), s2 = __shfl_xor_sync(0xffffffffu, s, o);
      online_merge(m, s, m2, s2);
    }
    if (threadIdx.x == 0) {
      const float lse = m + __logf(s);
      const long long lab = labels[row];
      lse_out[row] = lse;
      loss[row] = (lab >= 0 && lab < V) ? lse - __bfloat162float(x[lab]) : 0.f;
    }
  }
}

__global__ void __launch_bounds__(kCeThreads)
ce_bwd_kernel(const __nv_bfloat16* __restrict__ logits, long long ld, const long long* __restrict__ labels,
              const float* __restrict__ lse, const float* __restrict__ row_scale, __nv_bfloat16* __restrict__ dlogits,
              long long ldd, int V) {
  const int row = blockIdx.x;
  const __nv_bfloat16* x = logits + (size_t)row * ld;
  __nv_bfloat16* d = dlogits + (size_t)row * ldd;
  const float l = lse[row], g = row_scale[row];
  const long long lab = labels[row];
  const int nvec = V >> 3;
  for (int i = threadIdx.x; i < nvec; i += kCeThreads) {
    const uint4 u = ld_stream_v4(x + (size_t)i * 8);
    float v[8];
    float2 t;
    t = unpack_bf16x2(u.x); v[0] = t.x; v[1] = t.y;
    t = unpack_bf16x2(u.y); v[2] = t.x; v[3] = t.y;
    t = unpack_bf16x2(u.z); v[4] = t.x; v[5] = t.y;
    t = unpack_bf16x2(u.w); v[6] = t.x; v[7] = t.y;
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = __expf(v[k] - l) * g;
    const long long rel = lab - (long long)i * 8;
    if (rel >= 0 && rel < 8) {
#pragma unroll
      for (int k = 0; k < 8; ++k) if (k == rel) v[k] -= g;
    }
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(d + (size_t)i * 8) = o;
  }
  for (int i = (nvec << 3) + threadIdx.x; i < V; i += kCeThreads) {
    float p = __expf(__bfloat162float(x[i]) - l) * g;
    if (i == lab) p -= g;
    d[i] = __float2bfloat16(p);
  }
}

}  // namespace dedloc

using namespace dedloc;

extern "C" int dedloc_ce_fwd(const void* logits, long long ld, const long long* labels, int rows, int V, float* loss,
                             float* lse, void* stream) {
  DEDLOC_REQUIRE(rows >= 0 && V > 0, "cross_entropy: bad shape %d x %d", rows, V);
  DEDLOC_REQUIRE(ld % 8 == 0 && ((uintptr_t)logits & 15) == 0, "cross_entropy: rows must be 16-byte aligned (ld %% 8 == 0)");
  if (rows == 0) return 0;
  ce_fwd_kernel<<<rows, kCeThreads, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)logits, ld, labels, V, loss, lse);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dedloc_ce_bwd(const void* logits, long long ld, const long long* labels, const float* lse,
                             const float* row_scale, void* dlogits, long long ldd, int rows, int V, void* stream) {
  DEDLOC_REQUIRE(rows >= 0 && V > 0, "cross_entropy: bad shape %d x %d", rows, V);
  DEDLOC_REQUIRE(ld % 8 == 0 && ldd % 8 == 0 && ((uintptr_t)logits & 15) == 0 && ((uintptr_t)dlogits & 15) == 0,
                 "cross_entropy: rows must be 16-byte aligned (ld %% 8 == 0)");
  if (rows == 0) return 0;
  ce_bwd_kernel<<<rows, kCeThreads, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)logits, ld, labels, lse, row_scale,
                                                               (__nv_bfloat16*)dlogits, ldd, V);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}
