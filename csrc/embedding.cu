// ALBERT's factorised embedding front end and the 128-wide LayerNorm of its MLM head.
//
//   embed_ln_fwd   y[t] = LayerNorm_128(word[ids[t]] + type[tt[t]] + pos[t % S])      one kernel instead of three
//                  gathers, two adds and a LayerNorm; bf16 out feeds the K = 128 projection GEMM
//   embed_ln_bwd   LayerNorm backward + the three embedding backwards: the row sum is recomputed from the tables,
//                  dx is added straight into the fp32 main-gradient rows with red.add (no sort, no index_put kernels)
//   ln128_fwd/bwd  the same LayerNorm for a plain [M, 128] bf16 input (MLM head: dense -> gelu -> LayerNorm(128))
// Reference: HF AlbertEmbeddings / AlbertMLMHead under AlbertForPreTraining, built at albert/run_trainer.py:64-68.
//
// One warp per row: 128 elements = 32 lanes x 4. Parameter gradients (gamma, beta) are accumulated per warp in
// registers over all its rows, reduced per CTA in shared memory and added to the fp32 buffers with one red.add per
// element and CTA.
#include "common.cuh"
#include "../include/dedloc_b200.h"

namespace dedloc {
namespace emb {

constexpr int kE = 128;
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

__device__ __forceinline__ void load4_bf16(const __nv_bfloat16* p, float (&x)[4]) {
  const uint2 r = *reinterpret_cast<const uint2*>(p);
  const float2 a = unpack_bf16x2(r.x), b = unpack_bf16x2(r.y);
  x[0] = a.x; x[1] = a.y; x[2] = b.x; x[3] = b.y;
}
__device__ __forceinline__ void store4_bf16(__nv_bfloat16* p, const float (&x)[4]) {
  *reinterpret_cast<uint2*>(p) = make_uint2(pack_bf16x2(x[0], x[1]), pack_bf16x2(x[2], x[3]));
}

// row statistics of 128 values spread over a warp (4 per lane); two-pass for accuracy
__device__ __forceinline__ void row_stats(const float (&x)[4], float eps, float& mean, float& rstd) {
  mean = warp_sum(x[0] + x[1] + x[2] + x[3]) * (1.f / kE);
  float v = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) v = fmaf(x[i] - mean, x[i] - mean, v);
  rstd = rsqrtf(warp_sum(v) * (1.f / kE) + eps);
}

template <bool GATHER>
__global__ void __launch_bounds__(kThreads) ln128_fwd_kernel(const __nv_bfloat16* __restrict__ x, const long long* __restrict__ ids,
                                                            const long long* __restrict__ tts,
                                                            const __nv_bfloat16* __restrict__ word,
                                                            const __nv_bfloat16* __restrict__ pos,
                                                            const __nv_bfloat16* __restrict__ type,
                                                            const __nv_bfloat16* __restrict__ gamma,
                                                            const __nv_bfloat16* __restrict__ beta,
                                                            __nv_bfloat16* __restrict__ y, float* __restrict__ mean_out,
                                                            float* __restrict__ rstd_out, int M, int S, float eps) {
  const int lane = threadIdx.x & 31;
  const int col = lane * 4;
  float g[4], b[4];
  load4_bf16(gamma + col, g);
  load4_bf16(beta + col, b);
  for (int row = blockIdx.x * kWarps + (threadIdx.x >> 5); row < M; row += gridDim.x * kWarps) {
    float v[4];
    if constexpr (GATHER) {
      float w[4], p[4], t[4];
      load4_bf16(word + (size_t)ids[row] * kE + col, w);
      load4_bf16(pos + (size_t)(row % S) * kE + col, p);
      load4_bf16(type + (size_t)tts[row] * kE + col, t);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        // the reference adds bf16 tensors: round after every add like torch does
        const float s = __bfloat162float(__float2bfloat16(w[i] + t[i]));
        v[i] = __bfloat162float(__float2bfloat16(s + p[i]));
      }
    } else {
      load4_bf16(x + (size_t)row * kE + col, v);
    }
    float mean, rstd;
    row_stats(v, eps, mean, rstd);
    float o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) o[i] = fmaf((v[i] - mean) * rstd, g[i], b[i]);
    store4_bf16(y + (size_t)row * kE + col, o);
    if (lane == 0) {
      mean_out[row] = mean;
      rstd_out[row] = rstd;
    }
  }
}

template <bool GATHER>
__global__ void __launch_bounds__(kThreads) ln128_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                                                            const long long* __restrict__ ids, const long long* __restrict__ tts,
                                                            const __nv_bfloat16* __restrict__ word,
                                                            const __nv_bfloat16* __restrict__ pos,
                                                            const __nv_bfloat16* __restrict__ type,
                                                            const __nv_bfloat16* __restrict__ gamma,
                                                            const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
                                                            __nv_bfloat16* __restrict__ dx, float* __restrict__ gword,
                                                            float* __restrict__ gpos, float* __restrict__ gtype,
                                                            float* __restrict__ dgamma, float* __restrict__ dbeta, int M, int S,
                                                            long long pad_id, int n_types) {
  __shared__ float red[2][kWarps][kE];
  // the token-type table has 2 rows: every token of the batch would hit the same 256 addresses. Collect per CTA.
  constexpr int kTypeRows = 8;
  __shared__ float tacc[GATHER ? kTypeRows : 1][GATHER ? kE : 1];
  const bool type_in_smem = GATHER && n_types <= kTypeRows;
  if constexpr (GATHER) {
    for (int i = threadIdx.x; i < kTypeRows * kE; i += kThreads) tacc[i / kE][i % kE] = 0.f;
    __syncthreads();
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col = lane * 4;
  float g[4];
  load4_bf16(gamma + col, g);
  float ag[4] = {0.f, 0.f, 0.f, 0.f}, ab[4] = {0.f, 0.f, 0.f, 0.f};
  for (int row = blockIdx.x * kWarps + warp; row < M; row += gridDim.x * kWarps) {
    float v[4], d[4];
    long long id = 0, tt = 0;
    if constexpr (GATHER) {
      id = ids[row];
      tt = tts[row];
      float w[4], p[4], t[4];
      load4_bf16(word + (size_t)id * kE + col, w);
      load4_bf16(pos + (size_t)(row % S) * kE + col, p);
      load4_bf16(type + (size_t)tt * kE + col, t);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float s = __bfloat162float(__float2bfloat16(w[i] + t[i]));
        v[i] = __bfloat162float(__float2bfloat16(s + p[i]));
      }
    } else {
      load4_bf16(x + (size_t)row * kE + col, v);
    }
    load4_bf16(dy + (size_t)row * kE + col, d);
    const float mean = mean_in[row], rstd = rstd_in[row];
    float xh[4], dxh[4], s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      xh[i] = (v[i] - mean) * rstd;
      dxh[i] = d[i] * g[i];
      s1 += dxh[i];
      s2 = fmaf(dxh[i], xh[i], s2);
      ag[i] = fmaf(d[i], xh[i], ag[i]);
      ab[i] += d[i];
    }
    s1 = warp_sum(s1) * (1.f / kE);
    s2 = warp_sum(s2) * (1.f / kE);
    float o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) o[i] = rstd * (dxh[i] - s1 - xh[i] * s2);
    if constexpr (GATHER) {
      // the three embedding backwards: the same dx goes to one row of each table (fp32 main gradients)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (id != pad_id) atomicAdd(gword + (size_t)id * kE + col + i, o[i]);
        atomicAdd(gpos + (size_t)(row % S) * kE + col + i, o[i]);
        if (type_in_smem) atomicAdd(&tacc[tt][col + i], o[i]);
        else atomicAdd(gtype + (size_t)tt * kE + col + i, o[i]);
      }
    } else {
      store4_bf16(dx + (size_t)row * kE + col, o);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    red[0][warp][col + i] = ag[i];
    red[1][warp][col + i] = ab[i];
  }
  __syncthreads();
  if (threadIdx.x < kE) {
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      a += red[0][w][threadIdx.x];
      b += red[1][w][threadIdx.x];
    }
    atomicAdd(dgamma + threadIdx.x, a);
    atomicAdd(dbeta + threadIdx.x, b);
  }
  if constexpr (GATHER) {
    if (type_in_smem)
      for (int i = threadIdx.x; i < n_types * kE; i += kThreads) atomicAdd(gtype + i, tacc[i / kE][i % kE]);
  }
}

static int grid_for(int M) {
  int blocks = (M + kWarps - 1) / kWarps;
  return blocks < 1 ? 1 : (blocks > 148 * 8 ? 148 * 8 : blocks);
}

}  // namespace emb
}  // namespace dedloc

using namespace dedloc;

extern "C" int dedloc_embed_ln_fwd(const long long* ids, const long long* token_types, const void* word, const void* pos,
                                   const void* type, const void* gamma, const void* beta, void* y, float* mean, float* rstd,
                                   int M, int S, int E, float eps, void* stream) {
  DEDLOC_REQUIRE(E == emb::kE, "embed_ln: embedding size %d is not supported (128)", E);
  DEDLOC_REQUIRE(M > 0 && S > 0, "embed_ln: empty input");
  emb::ln128_fwd_kernel<true><<<emb::grid_for(M), emb::kThreads, 0, (cudaStream_t)stream>>>(
      nullptr, ids, token_types, (const __nv_bfloat16*)word, (const __nv_bfloat16*)pos, (const __nv_bfloat16*)type,
      (const __nv_bfloat16*)gamma, (const __nv_bfloat16*)beta, (__nv_bfloat16*)y, mean, rstd, M, S, eps);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dedloc_embed_ln_bwd(const void* dy, const long long* ids, const long long* token_types, const void* word,
                                   const void* pos, const void* type, const void* gamma, const float* mean, const float* rstd,
                                   float* gword, float* gpos, float* gtype, float* dgamma, float* dbeta, int M, int S, int E,
                                   long long pad_id, int n_types, void* stream) {
  DEDLOC_REQUIRE(E == emb::kE, "embed_ln: embedding size %d is not supported (128)", E);
  emb::ln128_bwd_kernel<true><<<emb::grid_for(M), emb::kThreads, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dy, nullptr, ids, token_types, (const __nv_bfloat16*)word, (const __nv_bfloat16*)pos,
      (const __nv_bfloat16*)type, (const __nv_bfloat16*)gamma, mean, rstd, nullptr, gword, gpos, gtype, dgamma, dbeta, M, S,
      pad_id, n_types);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dedloc_ln128_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd, int M,
                                float eps, void* stream) {
  DEDLOC_REQUIRE(M > 0, "ln128: empty input");
  emb::ln128_fwd_kernel<false><<<emb::grid_for(M), emb::kThreads, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, nullptr, nullptr, nullptr, nullptr, nullptr, (const __nv_bfloat16*)gamma,
      (const __nv_bfloat16*)beta, (__nv_bfloat16*)y, mean, rstd, M, 1, eps);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dedloc_ln128_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                                void* dx, float* dgamma, float* dbeta, int M, void* stream) {
  emb::ln128_bwd_kernel<false><<<emb::grid_for(M), emb::kThreads, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, nullptr, nullptr, nullptr, nullptr, nullptr,
      (const __nv_bfloat16*)gamma, mean, rstd, (__nv_bfloat16*)dx, nullptr, nullptr, nullptr, dgamma, dbeta, M, 1, -1, 0);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}
