// Row-wise kernels around the GEMMs of the shared ALBERT layer (SURVEY C4, C6 "fused GELU/LayerNorm"):
//   ln_fwd      y = LayerNorm(x) * gamma + beta            (bf16 in/out, fp32 statistics saved for backward)
//   ln_bwd      dx, and dgamma / dbeta ACCUMULATED (fp32 red.add) into main-grad buffers -- the shared layer is
//               applied 24 times per forward, so its LayerNorm parameters collect 24 contributions per step
//   colsum_acc  out[n] += sum_m dy[m, n]                   (bias gradients of the fused linears)
// One warp owns one row at a time (H / 32 elements per lane, 16-byte loads); HBM-bound by construction:
// forward moves 2 bytes in + 2 bytes out per element, backward 4 in + 2 out.
#include "common.cuh"
#include "../include/dedloc_b200.h"

namespace dedloc {

constexpr int kLnThreads = 256;  // 8 warps = 8 rows in flight per CTA

template <int CHUNKS>  // H = CHUNKS * 256; lane handles CHUNKS groups of 8 consecutive elements
__global__ void __launch_bounds__(kLnThreads) ln_fwd_kernel(const __nv_bfloat16* __restrict__ x,
                                                            const __nv_bfloat16* __restrict__ gamma,
                                                            const __nv_bfloat16* __restrict__ beta,
                                                            __nv_bfloat16* __restrict__ y, float* __restrict__ mean_out,
                                                            float* __restrict__ rstd_out, int M, float eps) {
  constexpr int H = CHUNKS * 256;
  const int lane = threadIdx.x & 31;
  const int warps_per_grid = gridDim.x * (kLnThreads / 32);
  float g[CHUNKS][8], b[CHUNKS][8];
#pragma unroll
  for (int c = 0; c < CHUNKS; ++c) {
    const int col = (c * 32 + lane) * 8;
    const uint4 gu = *reinterpret_cast<const uint4*>(gamma + col);
    const uint4 bu = *reinterpret_cast<const uint4*>(beta + col);
    float2 t;
    t = unpack_bf16x2(gu.x); g[c][0] = t.x; g[c][1] = t.y;
    t = unpack_bf16x2(gu.y); g[c][2] = t.x; g[c][3] = t.y;
    t = unpack_bf16x2(gu.z); g[c][4] = t.x; g[c][5] = t.y;
    t = unpack_bf16x2(gu.w); g[c][6] = t.x; g[c][7] = t.y;
    t = unpack_bf16x2(bu.x); b[c][0] = t.x; b[c][1] = t.y;
    t = unpack_bf16x2(bu.y); b[c][2] = t.x; b[c][3] = t.y;
    t = unpack_bf16x2(bu.z); b[c][4] = t.x; b[c][5] = t.y;
    t = unpack_bf16x2(bu.w); b[c][6] = t.x; b[c][7] = t.y;
  }
  for (int row = blockIdx.x * (kLnThreads / 32) + (threadIdx.x >> 5); row < M; row += warps_per_grid) {
    const __nv_bfloat16* xr = x + (size_t)row * H;
    float v[CHUNKS][8];
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) {
      const uint4 u = ld_stream_v4(xr + (c * 32 + lane) * 8);
      float2 t;
      t = unpack_bf16x2(u.x); v[c][0] = t.x; v[c][1] = t.y;
      t = unpack_bf16x2(u.y); v[c][2] = t.x; v[c][3] = t.y;
      t = unpack_bf16x2(u.z); v[c][4] = t.x; v[c][5] = t.y;
      t = unpack_bf16x2(u.w); v[c][6] = t.x; v[c][7] = t.y;
#pragma unroll
      for (int i = 0; i < 8; ++i) s += v[c][i];
    }
    const float mean = warp_sum(s) * (1.f / H);
    float s2 = 0.f;
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c)
#pragma unroll
      for (int i = 0; i < 8; ++i) { const float d = v[c][i] - mean; s2 = fmaf(d, d, s2); }
    const float rstd = rsqrtf(warp_sum(s2) * (1.f / H) + eps);
    __nv_bfloat16* yr = y + (size_t)row * H;
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) {
      float o[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = fmaf((v[c][i] - mean) * rstd, g[c][i], b[c][i]);
      st_stream_v4(yr + (c * 32 + lane) * 8,
                   make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7])));
    }
    if (lane == 0) { mean_out[row] = mean; rstd_out[row] = rstd; }
  }
}

template <int CHUNKS>
__global__ void __launch_bounds__(kLnThreads) ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy,
                                                            const __nv_bfloat16* __restrict__ x,
                                                            const __nv_bfloat16* __restrict__ gamma,
                                                            const float* __restrict__ mean_in,
                                                            const float* __restrict__ rstd_in,
                                                            __nv_bfloat16* __restrict__ dx, float* __restrict__ dgamma,
                                                            float* __restrict__ dbeta, int M) {
  constexpr int H = CHUNKS * 256;
  __shared__ float sred[kLnThreads / 32][32 * 8 + 8];  // per-warp staging for the final cross-warp reduction
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int warps_per_grid = gridDim.x * (kLnThreads / 32);
  float g[CHUNKS][8], dg[CHUNKS][8], db[CHUNKS][8];
#pragma unroll
  for (int c = 0; c < CHUNKS; ++c) {
    const uint4 gu = *reinterpret_cast<const uint4*>(gamma + (c * 32 + lane) * 8);
    float2 t;
    t = unpack_bf16x2(gu.x); g[c][0] = t.x; g[c][1] = t.y;
    t = unpack_bf16x2(gu.y); g[c][2] = t.x; g[c][3] = t.y;
    t = unpack_bf16x2(gu.z); g[c][4] = t.x; g[c][5] = t.y;
    t = unpack_bf16x2(gu.w); g[c][6] = t.x; g[c][7] = t.y;
#pragma unroll
    for (int i = 0; i < 8; ++i) { dg[c][i] = 0.f; db[c][i] = 0.f; }
  }
  // software pipeline: the 2 x CHUNKS 16-byte loads of the NEXT row are in flight while this row is reduced (the
  // register footprint allows one block per SM, so one row per warp in flight covers only ~60 % of HBM's
  // latency x bandwidth product)
  constexpr bool kPipe = CHUNKS <= 4;            // wider rows would spill
  uint4 xn[CHUNKS], dn[CHUNKS];
  int row = blockIdx.x * (kLnThreads / 32) + warp;
  if (kPipe && row < M) {
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) {
      const size_t off = (size_t)row * H + (c * 32 + lane) * 8;
      xn[c] = ld_stream_v4(x + off);
      dn[c] = ld_stream_v4(dy + off);
    }
  }
  for (; row < M; row += warps_per_grid) {
    const float mean = mean_in[row], rstd = rstd_in[row];
    float xh[CHUNKS][8], gy[CHUNKS][8];
    float s1 = 0.f, s2 = 0.f;
    uint4 xc[CHUNKS], dc[CHUNKS];
    if (!kPipe) {
#pragma unroll
      for (int c = 0; c < CHUNKS; ++c) {
        const size_t off = (size_t)row * H + (c * 32 + lane) * 8;
        xn[c] = ld_stream_v4(x + off);
        dn[c] = ld_stream_v4(dy + off);
      }
    }
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) { xc[c] = xn[c]; dc[c] = dn[c]; }
    const int nrow = row + warps_per_grid;
    if (kPipe && nrow < M) {
#pragma unroll
      for (int c = 0; c < CHUNKS; ++c) {
        const size_t off = (size_t)nrow * H + (c * 32 + lane) * 8;
        xn[c] = ld_stream_v4(x + off);
        dn[c] = ld_stream_v4(dy + off);
      }
    }
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) {
      const uint4 xu = xc[c];
      const uint4 du = dc[c];
      float xv[8], dv[8];
      float2 t;
      t = unpack_bf16x2(xu.x); xv[0] = t.x; xv[1] = t.y;
      t = unpack_bf16x2(xu.y); xv[2] = t.x; xv[3] = t.y;
      t = unpack_bf16x2(xu.z); xv[4] = t.x; xv[5] = t.y;
      t = unpack_bf16x2(xu.w); xv[6] = t.x; xv[7] = t.y;
      t = unpack_bf16x2(du.x); dv[0] = t.x; dv[1] = t.y;
      t = unpack_bf16x2(du.y); dv[2] = t.x; dv[3] = t.y;
      t = unpack_bf16x2(du.z); dv[4] = t.x; dv[5] = t.y;
      t = unpack_bf16x2(du.w); dv[6] = t.x; dv[7] = t.y;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        xh[c][i] = (xv[i] - mean) * rstd;
        gy[c][i] = dv[i] * g[c][i];
        dg[c][i] = fmaf(dv[i], xh[c][i], dg[c][i]);
        db[c][i] += dv[i];
        s1 += gy[c][i];
        s2 = fmaf(gy[c][i], xh[c][i], s2);
      }
    }
    s1 = warp_sum(s1) * (1.f / H);
    s2 = warp_sum(s2) * (1.f / H);
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) {
      float o[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = rstd * (gy[c][i] - s1 - xh[c][i] * s2);
      st_stream_v4(dx + (size_t)row * H + (c * 32 + lane) * 8,
                   make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7])));
    }
  }
  // cross-warp reduction of the per-lane parameter gradients, one chunk at a time, then one red.add per column
  for (int pass = 0; pass < 2; ++pass) {
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) {
      __syncthreads();
#pragma unroll
      for (int i = 0; i < 8; ++i) sred[warp][lane * 8 + i] = pass == 0 ? dg[c][i] : db[c][i];
      __syncthreads();
      {
        const int col = threadIdx.x;  // 256 threads <-> 256 columns of this chunk
        float acc = 0.f;
#pragma unroll
        for (int w = 0; w < kLnThreads / 32; ++w) acc += sred[w][col];
        atomicAdd((pass == 0 ? dgamma : dbeta) + c * 256 + col, acc);
      }
    }
  }
}

// out[n] += sum over rows of dy[m, n]. Grid: (N / 256 column blocks, row splits); 8 bf16 per lane, 32 lanes = 256 cols.
__global__ void __launch_bounds__(256) colsum_acc_kernel(const __nv_bfloat16* __restrict__ dy, float* __restrict__ out,
                                                         int M, int N, long long ld) {
  __shared__ float sred[8][256 + 8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col0 = blockIdx.x * 256 + lane * 8;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (col0 < N) {
    // four rows in flight per warp: one 16-byte load per lane does not cover the HBM latency-bandwidth product
    const int stride = gridDim.y * 8;
    int row = blockIdx.y * 8 + warp;
    for (; row + 3 * stride < M; row += 4 * stride) {
      uint4 u[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) u[k] = ld_stream_v4(dy + (size_t)(row + k * stride) * ld + col0);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float2 t;
        t = unpack_bf16x2(u[k].x); acc[0] += t.x; acc[1] += t.y;
        t = unpack_bf16x2(u[k].y); acc[2] += t.x; acc[3] += t.y;
        t = unpack_bf16x2(u[k].z); acc[4] += t.x; acc[5] += t.y;
        t = unpack_bf16x2(u[k].w); acc[6] += t.x; acc[7] += t.y;
      }
    }
    for (; row < M; row += stride) {
      const uint4 u = ld_stream_v4(dy + (size_t)row * ld + col0);
      float2 t;
      t = unpack_bf16x2(u.x); acc[0] += t.x; acc[1] += t.y;
      t = unpack_bf16x2(u.y); acc[2] += t.x; acc[3] += t.y;
      t = unpack_bf16x2(u.z); acc[4] += t.x; acc[5] += t.y;
      t = unpack_bf16x2(u.w); acc[6] += t.x; acc[7] += t.y;
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) sred[warp][lane * 8 + i] = acc[i];
  __syncthreads();
  const int col = blockIdx.x * 256 + threadIdx.x;
  if (col < N) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += sred[w][threadIdx.x];
    atomicAdd(out + col, s);
  }
}

static int ln_grid(int M) {
  int want = (M + 7) / 8;
  int cap = 148 * 4;
  return want < cap ? (want < 1 ? 1 : want) : cap;
}

}  // namespace dedloc

using namespace dedloc;

extern "C" int dedloc_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean,
                                    float* rstd, int M, int H, float eps, void* stream) {
  DEDLOC_REQUIRE(H % 256 == 0 && H >= 256 && H <= 2048, "layernorm: H (%d) must be a multiple of 256 in [256, 2048]", H);
  cudaStream_t st = (cudaStream_t)stream;
  const __nv_bfloat16 *xp = (const __nv_bfloat16*)x, *gp = (const __nv_bfloat16*)gamma, *bp = (const __nv_bfloat16*)beta;
  __nv_bfloat16* yp = (__nv_bfloat16*)y;
  const int grid = ln_grid(M);
  switch (H / 256) {
    case 1: ln_fwd_kernel<1><<<grid, kLnThreads, 0, st>>>(xp, gp, bp, yp, mean, rstd, M, eps); break;
    case 2: ln_fwd_kernel<2><<<grid, kLnThreads, 0, st>>>(xp, gp, bp, yp, mean, rstd, M, eps); break;
    case 3: ln_fwd_kernel<3><<<grid, kLnThreads, 0, st>>>(xp, gp, bp, yp, mean, rstd, M, eps); break;
    case 4: ln_fwd_kernel<4><<<grid, kLnThreads, 0, st>>>(xp, gp, bp, yp, mean, rstd, M, eps); break;
    case 8: ln_fwd_kernel<8><<<grid, kLnThreads, 0, st>>>(xp, gp, bp, yp, mean, rstd, M, eps); break;
    default: set_last_error("layernorm: unsupported H %d", H); return DEDLOC_EUNSUPPORTED;
  }
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dedloc_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                                    const float* rstd, void* dx, float* dgamma, float* dbeta, int M, int H,
                                    void* stream) {
  DEDLOC_REQUIRE(H % 256 == 0 && H >= 256 && H <= 2048, "layernorm: H (%d) must be a multiple of 256 in [256, 2048]", H);
  cudaStream_t st = (cudaStream_t)stream;
  const __nv_bfloat16 *dyp = (const __nv_bfloat16*)dy, *xp = (const __nv_bfloat16*)x, *gp = (const __nv_bfloat16*)gamma;
  __nv_bfloat16* dxp = (__nv_bfloat16*)dx;
  int grid = ln_grid(M);
  if (grid > 148 * 2) grid = 148 * 2;  // fewer CTAs -> fewer atomics on dgamma / dbeta
  switch (H / 256) {
    case 1: ln_bwd_kernel<1><<<grid, kLnThreads, 0, st>>>(dyp, xp, gp, mean, rstd, dxp, dgamma, dbeta, M); break;
    case 2: ln_bwd_kernel<2><<<grid, kLnThreads, 0, st>>>(dyp, xp, gp, mean, rstd, dxp, dgamma, dbeta, M); break;
    case 3: ln_bwd_kernel<3><<<grid, kLnThreads, 0, st>>>(dyp, xp, gp, mean, rstd, dxp, dgamma, dbeta, M); break;
    case 4: ln_bwd_kernel<4><<<grid, kLnThreads, 0, st>>>(dyp, xp, gp, mean, rstd, dxp, dgamma, dbeta, M); break;
    case 8: ln_bwd_kernel<8><<<grid, kLnThreads, 0, st>>>(dyp, xp, gp, mean, rstd, dxp, dgamma, dbeta, M); break;
    default: set_last_error("layernorm: unsupported H %d", H); return DEDLOC_EUNSUPPORTED;
  }
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dedloc_colsum_accumulate(const void* dy, float* out, int M, int N, long long ld, void* stream) {
  DEDLOC_REQUIRE(N % 8 == 0 && ld % 8 == 0, "colsum: N and ld must be multiples of 8");
  const int col_blocks = (N + 255) / 256;
  int row_splits = (148 * 4 + col_blocks - 1) / col_blocks;
  const int max_splits = (M + 63) / 64;
  if (row_splits > max_splits) row_splits = max_splits;
  if (row_splits < 1) row_splits = 1;
  colsum_acc_kernel<<<dim3(col_blocks, row_splits), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)dy, out, M, N, ld);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}
