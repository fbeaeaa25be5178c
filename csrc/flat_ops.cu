// Flat-vector kernels on the gradient path of the collaborative optimizer.
//
//   K0a  sum of squares of the local (bf16 or fp32) gradient   -> global L2 norm for max_grad_norm clipping
//   K0b  acc += clip_coef * grad ; grad = 0                    -> the "accumulate a local step" of H3
//   ENC  payload/aux = encode(acc * scale)                     -> wire codecs of H8, fused with the 1/n_steps
//
// All vectors are padded to a multiple of kChunk (2048) elements; padding is always zero.
#include "codec.cuh"
#include "../include/dedloc_b200.h"

namespace dedloc {

// ------------------------------------------------------------------------------------------------
// K0a: sum of squares with a two-stage reduction. Stage 1 writes one partial per block into `scratch`,
// stage 2 (one block) folds them in a fixed order, so the result is run-to-run deterministic.
// ------------------------------------------------------------------------------------------------
constexpr int kSumsqThreads = 256;
constexpr int kSumsqMaxBlocks = 1184;  // 148 SMs x 8

template <typename T>
__global__ void __launch_bounds__(kSumsqThreads) sumsq_stage1(const T* __restrict__ x, long long n,
                                                              float* __restrict__ partials) {
  __shared__ float red[64];
  float s = 0.f;
  const long long nvec = n / 8;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    if constexpr (sizeof(T) == 2) {
      uint4 r = ld_stream_v4(reinterpret_cast<const uint4*>(x) + i);
      float2 t;
      t = unpack_bf16x2(r.x); s = fmaf(t.x, t.x, s); s = fmaf(t.y, t.y, s);
      t = unpack_bf16x2(r.y); s = fmaf(t.x, t.x, s); s = fmaf(t.y, t.y, s);
      t = unpack_bf16x2(r.z); s = fmaf(t.x, t.x, s); s = fmaf(t.y, t.y, s);
      t = unpack_bf16x2(r.w); s = fmaf(t.x, t.x, s); s = fmaf(t.y, t.y, s);
    } else {
      const uint4* p = reinterpret_cast<const uint4*>(x) + 2 * i;
      uint4 a = ld_stream_v4(p), b = ld_stream_v4(p + 1);
      float v[8] = {__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(a.z), __uint_as_float(a.w),
                    __uint_as_float(b.x), __uint_as_float(b.y), __uint_as_float(b.z), __uint_as_float(b.w)};
#pragma unroll
      for (int k = 0; k < 8; ++k) s = fmaf(v[k], v[k], s);
    }
  }
  float2 r = block_sum2(s, 0.f, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = r.x;
}

__global__ void __launch_bounds__(kSumsqThreads) sumsq_stage2(const float* __restrict__ partials, int nparts,
                                                              float* __restrict__ out) {
  __shared__ float red[64];
  float s = 0.f;
  for (int i = threadIdx.x; i < nparts; i += blockDim.x) s += partials[i];
  float2 r = block_sum2(s, 0.f, red);
  if (threadIdx.x == 0) out[0] = r.x;
}

template <typename T>
static int launch_sumsq(const T* x, long long n, float* scratch, int scratch_floats, float* out,
                        cudaStream_t st) {
  DEDLOC_REQUIRE(n % 8 == 0, "sumsq: n (%lld) must be a multiple of 8", n);
  long long want = (n / 8 + kSumsqThreads - 1) / kSumsqThreads;
  int blocks = (int)(want < 1 ? 1 : (want > kSumsqMaxBlocks ? kSumsqMaxBlocks : want));
  if (blocks > scratch_floats) blocks = scratch_floats;
  DEDLOC_REQUIRE(blocks >= 1, "sumsq: scratch must hold at least one float");
  sumsq_stage1<T><<<blocks, kSumsqThreads, 0, st>>>(x, n, scratch);
  sumsq_stage2<<<1, kSumsqThreads, 0, st>>>(scratch, blocks, out);
  count_launch(2);
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------
// K0b: acc += coef * grad; grad = 0.  coef = min(1, max_norm / (sqrt(normsq) + 1e-6)) when max_norm > 0
// (the clip_grad_norm_ rule that HF Trainer applies before every optimizer.step()).
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) accumulate_clip_kernel(float* __restrict__ acc, T* __restrict__ grad,
                                                              long long n, const float* __restrict__ normsq,
                                                              float max_norm, int zero_grad) {
  float coef = 1.f;
  if (max_norm > 0.f && normsq != nullptr) {
    const float nrm = sqrtf(*normsq);
    coef = fminf(1.f, max_norm / (nrm + 1e-6f));
    if (!isfinite(nrm)) coef = nrm;  // propagate NaN/Inf so the non-finite guard downstream fires
  }
  // Four elements per thread and step: consecutive threads touch consecutive 16-byte (acc, fp32 grad) or 8-byte
  // (bf16 grad) words, so every warp-wide access covers whole sectors (the earlier 32-bytes-per-thread layout used
  // half of each sector per instruction and ran at 0.29 of the copy bandwidth; torch's add_ + zero_ at 0.87).
  // All loads of kU steps are issued before the first dependent store.
  constexpr int kU = 4;
  const long long nvec = n / 4;
  const long long stride = (long long)gridDim.x * blockDim.x;
  float4* __restrict__ a4 = reinterpret_cast<float4*>(acc);
  for (long long i0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; i0 < nvec; i0 += stride * kU) {
    float4 a[kU], g[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = i0 + u * stride;
      if (i < nvec) {
        a[u] = a4[i];
        if constexpr (sizeof(T) == 2) {
          const uint2 r = reinterpret_cast<const uint2*>(grad)[i];
          const float2 lo = unpack_bf16x2(r.x), hi = unpack_bf16x2(r.y);
          g[u] = make_float4(lo.x, lo.y, hi.x, hi.y);
        } else {
          g[u] = reinterpret_cast<const float4*>(grad)[i];
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = i0 + u * stride;
      if (i < nvec) {
        float4 r = a[u];
        r.x = fmaf(coef, g[u].x, r.x); r.y = fmaf(coef, g[u].y, r.y);
        r.z = fmaf(coef, g[u].z, r.z); r.w = fmaf(coef, g[u].w, r.w);
        a4[i] = r;
        if (zero_grad) {
          if constexpr (sizeof(T) == 2) reinterpret_cast<uint2*>(grad)[i] = make_uint2(0u, 0u);
          else reinterpret_cast<float4*>(grad)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// ENC: per-chunk codecs (NONE / FLOAT16 / BFLOAT16 / MEANSTD_16BIT / BLOCKWISE_FP8). One CTA per chunk.
// ------------------------------------------------------------------------------------------------
template <int CODEC>
__global__ void __launch_bounds__(256) encode_chunk_kernel(const float* __restrict__ src, float scale,
                                                           uint8_t* __restrict__ payload,
                                                           float* __restrict__ aux) {
  __shared__ float red[64];
  const long long c = blockIdx.x;
  const long long e = c * kChunk + threadIdx.x * 8;
  const float4* sp = reinterpret_cast<const float4*>(src + e);
  float4 a = sp[0], b = sp[1];
  float x[8] = {a.x * scale, a.y * scale, a.z * scale, a.w * scale,
                b.x * scale, b.y * scale, b.z * scale, b.w * scale};
  if constexpr (CODEC == CODEC_NONE) {
    uint4* p = reinterpret_cast<uint4*>(payload + e * 4);
    p[0] = make_uint4(__float_as_uint(x[0]), __float_as_uint(x[1]), __float_as_uint(x[2]), __float_as_uint(x[3]));
    p[1] = make_uint4(__float_as_uint(x[4]), __float_as_uint(x[5]), __float_as_uint(x[6]), __float_as_uint(x[7]));
  } else if constexpr (CODEC == CODEC_FLOAT16) {
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = x[i] > 65504.f ? 65504.f : (x[i] < -65504.f ? -65504.f : x[i]);  // NaN stays NaN
    *reinterpret_cast<uint4*>(payload + e * 2) =
        make_uint4(pack_f16x2(x[0], x[1]), pack_f16x2(x[2], x[3]), pack_f16x2(x[4], x[5]), pack_f16x2(x[6], x[7]));
  } else if constexpr (CODEC == CODEC_BFLOAT16) {
    *reinterpret_cast<uint4*>(payload + e * 2) =
        make_uint4(pack_bf16x2(x[0], x[1]), pack_bf16x2(x[2], x[3]), pack_bf16x2(x[4], x[5]), pack_bf16x2(x[6], x[7]));
  } else if constexpr (CODEC == CODEC_MEANSTD_16BIT) {
    float s = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { s += x[i]; s2 = fmaf(x[i], x[i], s2); }
    float2 r = block_sum2(s, s2, red);
    const float mean = r.x / kChunk;
    const float var = fmaxf(r.y / kChunk - mean * mean, 0.f);
    const float sd = fmaxf(sqrtf(var), 1e-6f);
    const float inv = 1.f / sd;
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = (x[i] - mean) * inv;
    *reinterpret_cast<uint4*>(payload + e * 2) =
        make_uint4(pack_f16x2(x[0], x[1]), pack_f16x2(x[2], x[3]), pack_f16x2(x[4], x[5]), pack_f16x2(x[6], x[7]));
    if (threadIdx.x == 0) { aux[2 * c] = mean; aux[2 * c + 1] = sd; }
  } else {  // BLOCKWISE_FP8: 16 consecutive threads (128 elements) share one scale
    float amax = 0.f;
    bool bad = false;
#pragma unroll
    for (int i = 0; i < 8; ++i) { amax = fmaxf(amax, fabsf(x[i])); bad |= !isfinite(x[i]); }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
      amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
      bad |= (bool)__shfl_xor_sync(0xffffffffu, (int)bad, o);
    }
    // (a block of denormals: amax / 448 underflows to 0 and 1 / 0 would poison the block; FLT_MIN keeps it finite)
    float s = amax > 0.f ? fmaxf(amax / 448.f, 1.17549435e-38f) : 1.f;
    if (bad) s = __int_as_float(0x7fc00000);  // NaN scale: decoder reproduces a non-finite block
    const float inv = bad ? 0.f : 1.f / s;
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      lo |= (uint32_t)float_to_fp8_e4m3(x[i] * inv) << (8 * i);
      hi |= (uint32_t)float_to_fp8_e4m3(x[4 + i] * inv) << (8 * i);
    }
    *reinterpret_cast<uint2*>(payload + e) = make_uint2(lo, hi);
    if ((threadIdx.x & 15) == 0) aux[e / kFp8Block] = s;
  }
}

// ------------------------------------------------------------------------------------------------
// ENC: 8-bit codebook codecs. One 1024-thread CTA per 65536-element block.
//   UNIFORM:  bucket = clamp(round((x-mean)/(6 sigma/256)) + 128, 0, 255)
//   QUANTILE: borders = quantiles k/256 of the block, estimated from a 2048-bin histogram over [min,max]
// codebook[b] = mean of the values that fell in bucket b (0 when empty).
// ------------------------------------------------------------------------------------------------
constexpr int kLutThreads = 1024;
constexpr int kHistBins = 2048;

template <bool QUANTILE>
__global__ void __launch_bounds__(kLutThreads) encode_lut_kernel(const float* __restrict__ src, long long n,
                                                                 float scale, uint8_t* __restrict__ payload,
                                                                 float* __restrict__ aux) {
  __shared__ float red[64];
  __shared__ float bsum[256];
  __shared__ unsigned int bcnt[256];
  __shared__ float borders[256];
  __shared__ unsigned int hist[QUANTILE ? kHistBins : 1];
  __shared__ float s_lo, s_hi;

  const long long base = (long long)blockIdx.x * kCodecBlock;
  const long long rem = n - base;
  const int nvalid = (int)(rem < kCodecBlock ? rem : kCodecBlock);  // multiple of kChunk
  const int tid = threadIdx.x;
  const int iters = nvalid / (kLutThreads * 8) + ((nvalid % (kLutThreads * 8)) ? 1 : 0);

  // pass 1: mean / std (+ min / max)
  float s = 0.f, s2 = 0.f, mn = INFINITY, mx = -INFINITY;
  for (int it = 0; it < iters; ++it) {
    const int off = (it * kLutThreads + tid) * 8;
    if (off < nvalid) {
      const float4* sp = reinterpret_cast<const float4*>(src + base + off);
      float4 a = sp[0], b = sp[1];
      float x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float v = x[i] * scale;
        s += v; s2 = fmaf(v, v, s2); mn = fminf(mn, v); mx = fmaxf(mx, v);
      }
    }
  }
  float2 r = block_sum2(s, s2, red);
  const float mean = r.x / nvalid;
  const float sd = sqrtf(fmaxf(r.y / nvalid - mean * mean, 0.f));
  if (tid < 256) { bsum[tid] = 0.f; bcnt[tid] = 0u; }

  float inv_step = 0.f;
  if constexpr (!QUANTILE) {
    const float step = 6.f * sd / 256.f;
    inv_step = step > 0.f ? 1.f / step : 0.f;
    __syncthreads();
  } else {
    mn = -warp_max(-mn);
    mx = warp_max(mx);
    __syncthreads();
    if ((tid & 31) == 0) { red[tid >> 5] = mn; red[32 + (tid >> 5)] = mx; }
    __syncthreads();
    if (tid < 32) {
      float a = red[tid], b = red[32 + tid];
      a = -warp_max(-a);
      b = warp_max(b);
      if (tid == 0) { s_lo = a; s_hi = b; }
    }
    for (int i = tid; i < kHistBins; i += kLutThreads) hist[i] = 0u;
    __syncthreads();
    const float lo = s_lo, hi = s_hi;
    const float width = (hi - lo) / kHistBins;
    const float inv_w = width > 0.f ? 1.f / width : 0.f;
    for (int it = 0; it < iters; ++it) {
      const int off = (it * kLutThreads + tid) * 8;
      if (off < nvalid) {
        const float4* sp = reinterpret_cast<const float4*>(src + base + off);
        float4 a = sp[0], b = sp[1];
        float x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          int bin = (int)((x[i] * scale - lo) * inv_w);
          bin = bin < 0 ? 0 : (bin >= kHistBins ? kHistBins - 1 : bin);
          atomicAdd(&hist[bin], 1u);
        }
      }
    }
    __syncthreads();
    // inclusive prefix sum of the histogram by one warp (2048 bins, 64 per lane)
    if (tid < 32) {
      unsigned int local = 0;
      for (int i = 0; i < kHistBins / 32; ++i) local += hist[tid * (kHistBins / 32) + i];
      unsigned int incl = local;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (tid >= o) incl += t;
      }
      unsigned int run = incl - local;
      for (int i = 0; i < kHistBins / 32; ++i) {
        run += hist[tid * (kHistBins / 32) + i];
        hist[tid * (kHistBins / 32) + i] = run;
      }
    }
    __syncthreads();
    // border k (k = 0..254) separates bucket k from bucket k+1: value at CDF = (k+1)/256
    if (tid < 255) {
      const unsigned int target = (unsigned int)(((long long)(tid + 1) * nvalid) / 256);
      int lo_b = 0, hi_b = kHistBins - 1;
      while (lo_b < hi_b) {
        int mid = (lo_b + hi_b) >> 1;
        if (hist[mid] >= target) hi_b = mid; else lo_b = mid + 1;
      }
      const unsigned int below = lo_b > 0 ? hist[lo_b - 1] : 0u;
      const unsigned int inbin = hist[lo_b] - below;
      const float frac = inbin > 0 ? (float)(target - below) / (float)inbin : 0.f;
      borders[tid] = lo + ((float)lo_b + frac) * width;
    }
    if (tid == 255) borders[255] = INFINITY;
    __syncthreads();
  }

  // pass 2: quantise, write payload, accumulate bucket sums
  for (int it = 0; it < iters; ++it) {
    const int off = (it * kLutThreads + tid) * 8;
    if (off < nvalid) {
      const float4* sp = reinterpret_cast<const float4*>(src + base + off);
      float4 a = sp[0], b = sp[1];
      float x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
      uint32_t w[2] = {0u, 0u};
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float v = x[i] * scale;
        int q;
        if constexpr (!QUANTILE) {
          q = __float2int_rn((v - mean) * inv_step) + 128;
          q = q < 0 ? 0 : (q > 255 ? 255 : q);
        } else {
          // number of borders <= v  (borders ascending; borders[255] = +inf)
          int lo_b = 0, hi_b = 255;
          while (lo_b < hi_b) {
            int mid = (lo_b + hi_b) >> 1;
            if (borders[mid] > v) hi_b = mid; else lo_b = mid + 1;
          }
          q = lo_b;
        }
        if (!isfinite(v)) q = 255;
        atomicAdd(&bsum[q], v);
        atomicAdd(&bcnt[q], 1u);
        w[i >> 2] |= (uint32_t)q << (8 * (i & 3));
      }
      *reinterpret_cast<uint2*>(payload + base + off) = make_uint2(w[0], w[1]);
    }
  }
  __syncthreads();
  if (tid < 256) {
    const unsigned int cnt = bcnt[tid];
    aux[(long long)blockIdx.x * 256 + tid] = cnt ? bsum[tid] / (float)cnt : 0.f;
  }
}

}  // namespace dedloc

using namespace dedloc;

extern "C" int dedloc_sumsq(const void* x, int is_bf16, long long n, float* scratch, int scratch_floats,
                            float* out, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (is_bf16) return launch_sumsq<__nv_bfloat16>((const __nv_bfloat16*)x, n, scratch, scratch_floats, out, st);
  return launch_sumsq<float>((const float*)x, n, scratch, scratch_floats, out, st);
}

extern "C" int dedloc_accumulate_clip(float* acc, void* grad, int is_bf16, long long n, const float* normsq,
                                      float max_norm, int zero_grad, void* stream) {
  DEDLOC_REQUIRE(n % 8 == 0, "accumulate_clip: n (%lld) must be a multiple of 8", n);
  cudaStream_t st = (cudaStream_t)stream;
  long long want = (n / 4 + 256 * 4 - 1) / (256 * 4);     // 4 elements x 4 unrolled steps per thread
  int blocks = (int)(want < 1 ? 1 : (want > 148 * 8 ? 148 * 8 : want));
  if (is_bf16)
    accumulate_clip_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(acc, (__nv_bfloat16*)grad, n, normsq, max_norm, zero_grad);
  else
    accumulate_clip_kernel<float><<<blocks, 256, 0, st>>>(acc, (float*)grad, n, normsq, max_norm, zero_grad);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dedloc_encode(int codec, const float* src, long long n, float scale, void* payload, float* aux,
                             void* stream) {
  DEDLOC_REQUIRE(n > 0 && n % kChunk == 0, "encode: n (%lld) must be a positive multiple of %d", n, kChunk);
  cudaStream_t st = (cudaStream_t)stream;
  const int chunks = (int)(n / kChunk);
  uint8_t* pl = (uint8_t*)payload;
  switch (codec) {
    case CODEC_NONE: encode_chunk_kernel<CODEC_NONE><<<chunks, 256, 0, st>>>(src, scale, pl, aux); break;
    case CODEC_FLOAT16: encode_chunk_kernel<CODEC_FLOAT16><<<chunks, 256, 0, st>>>(src, scale, pl, aux); break;
    case CODEC_BFLOAT16: encode_chunk_kernel<CODEC_BFLOAT16><<<chunks, 256, 0, st>>>(src, scale, pl, aux); break;
    case CODEC_MEANSTD_16BIT: encode_chunk_kernel<CODEC_MEANSTD_16BIT><<<chunks, 256, 0, st>>>(src, scale, pl, aux); break;
    case CODEC_BLOCKWISE_FP8: encode_chunk_kernel<CODEC_BLOCKWISE_FP8><<<chunks, 256, 0, st>>>(src, scale, pl, aux); break;
    case CODEC_UNIFORM_8BIT: {
      int blocks = (int)((n + kCodecBlock - 1) / kCodecBlock);
      encode_lut_kernel<false><<<blocks, kLutThreads, 0, st>>>(src, n, scale, pl, aux);
      break;
    }
    case CODEC_QUANTILE_8BIT: {
      int blocks = (int)((n + kCodecBlock - 1) / kCodecBlock);
      encode_lut_kernel<true><<<blocks, kLutThreads, 0, st>>>(src, n, scale, pl, aux);
      break;
    }
    default:
      set_last_error("encode: unknown codec %d", codec);
      return DEDLOC_EINVAL;
  }
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" long long dedloc_codec_aux_floats(int codec, long long n) { return codec_aux_floats(codec, n); }
extern "C" int dedloc_codec_payload_bytes_per_elem(int codec) { return codec_payload_bytes_per_elem(codec); }
extern "C" int dedloc_chunk_elems(void) { return kChunk; }
