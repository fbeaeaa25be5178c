// Wire codecs for the averaging path (SURVEY H8 + the B200-native additions).
// A "payload" is the encoded flat gradient vector in a peer's symmetric send buffer; "aux" holds the
// per-chunk / per-block side information (statistics, codebooks, scales) as fp32.
//
//   NONE            fp32 payload
//   FLOAT16         fp16 payload, values clamped to +-65504
//   BFLOAT16        bf16 payload
//   MEANSTD_16BIT   fp16 payload of (x-mean)/std, aux = (mean, std) per 2048-element chunk
//   UNIFORM_8BIT    u8 bucket index, aux = 256-entry fp32 codebook (bucket means) per 65536-element block;
//                   buckets are uniform over mean +- 3 sigma  (6 sigma / 256 wide)
//   QUANTILE_8BIT   same payload/decoder, bucket borders are (approximate) quantiles of the block
//   BLOCKWISE_FP8   e4m3 payload, aux = one fp32 scale per 128 elements
//
// Decoding always happens in registers inside the reducing kernel (fused_allreduce.cu); encoding is
// fused with the "finalize gradient accumulators" pass (flat_ops.cu).
#pragma once
#include "common.cuh"

namespace dedloc {

enum Codec : int {
  CODEC_NONE = 0,
  CODEC_MEANSTD_16BIT = 1,
  CODEC_FLOAT16 = 2,
  CODEC_QUANTILE_8BIT = 3,
  CODEC_UNIFORM_8BIT = 4,
  CODEC_BFLOAT16 = 5,
  CODEC_BLOCKWISE_FP8 = 6,
  CODEC_COUNT = 7,
};

constexpr int kChunk = 2048;            // elements per chunk: 256 threads x 8 elements
constexpr int kCodecBlockChunks = 32;   // 8-bit codebooks cover 32 chunks = 65536 elements
constexpr int kCodecBlock = kChunk * kCodecBlockChunks;
constexpr int kFp8Block = 128;          // elements per fp8 scale

__host__ __device__ inline int codec_payload_bytes_per_elem(int codec) {
  switch (codec) {
    case CODEC_NONE: return 4;
    case CODEC_FLOAT16:
    case CODEC_BFLOAT16:
    case CODEC_MEANSTD_16BIT: return 2;
    default: return 1;
  }
}
// number of fp32 aux entries for a vector of n (padded) elements
__host__ __device__ inline long long codec_aux_floats(int codec, long long n) {
  switch (codec) {
    case CODEC_MEANSTD_16BIT: return 2 * (n / kChunk);
    case CODEC_UNIFORM_8BIT:
    case CODEC_QUANTILE_8BIT: return 256 * ((n + kCodecBlock - 1) / kCodecBlock);
    case CODEC_BLOCKWISE_FP8: return n / kFp8Block;
    default: return 0;
  }
}

#ifdef __CUDACC__

__device__ __forceinline__ float fp8_e4m3_to_float(uint8_t b) {
  __half_raw h = __nv_cvt_fp8_to_halfraw((__nv_fp8_storage_t)b, __NV_E4M3);
  return __half2float(*reinterpret_cast<__half*>(&h));
}
__device__ __forceinline__ uint8_t float_to_fp8_e4m3(float x) {
  return (uint8_t)__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3);
}

// Raw (still encoded) 8 elements of one peer: at most 32 bytes.
struct Raw8 {
  uint4 a, b;
};

// Issue the loads for 8 consecutive elements starting at element index e (multiple of 8).
template <int CODEC>
__device__ __forceinline__ void load_raw8(const uint8_t* __restrict__ payload, long long e, Raw8& r) {
  if constexpr (CODEC == CODEC_NONE) {
    const uint4* p = reinterpret_cast<const uint4*>(payload + e * 4);
    r.a = ld_stream_v4(p);
    r.b = ld_stream_v4(p + 1);
  } else if constexpr (CODEC == CODEC_FLOAT16 || CODEC == CODEC_BFLOAT16 || CODEC == CODEC_MEANSTD_16BIT) {
    r.a = ld_stream_v4(payload + e * 2);
  } else {
    uint2 t = ld_stream_v2(payload + e);
    r.a.x = t.x;
    r.a.y = t.y;
  }
}

// Decode 8 elements. `aux` is the peer's aux array; `lut` is that peer's 256-entry codebook staged in
// shared memory (8-bit codecs only); c = chunk index of e.
template <int CODEC>
__device__ __forceinline__ void decode8(const Raw8& r, const float* __restrict__ aux, const float* lut,
                                        long long e, long long c, float (&x)[8]) {
  if constexpr (CODEC == CODEC_NONE) {
    x[0] = __uint_as_float(r.a.x); x[1] = __uint_as_float(r.a.y);
    x[2] = __uint_as_float(r.a.z); x[3] = __uint_as_float(r.a.w);
    x[4] = __uint_as_float(r.b.x); x[5] = __uint_as_float(r.b.y);
    x[6] = __uint_as_float(r.b.z); x[7] = __uint_as_float(r.b.w);
  } else if constexpr (CODEC == CODEC_FLOAT16 || CODEC == CODEC_MEANSTD_16BIT) {
    float2 t;
    t = unpack_f16x2(r.a.x); x[0] = t.x; x[1] = t.y;
    t = unpack_f16x2(r.a.y); x[2] = t.x; x[3] = t.y;
    t = unpack_f16x2(r.a.z); x[4] = t.x; x[5] = t.y;
    t = unpack_f16x2(r.a.w); x[6] = t.x; x[7] = t.y;
    if constexpr (CODEC == CODEC_MEANSTD_16BIT) {
      const float mean = __ldg(aux + 2 * c), sd = __ldg(aux + 2 * c + 1);
#pragma unroll
      for (int i = 0; i < 8; ++i) x[i] = fmaf(x[i], sd, mean);
    }
  } else if constexpr (CODEC == CODEC_BFLOAT16) {
    float2 t;
    t = unpack_bf16x2(r.a.x); x[0] = t.x; x[1] = t.y;
    t = unpack_bf16x2(r.a.y); x[2] = t.x; x[3] = t.y;
    t = unpack_bf16x2(r.a.z); x[4] = t.x; x[5] = t.y;
    t = unpack_bf16x2(r.a.w); x[6] = t.x; x[7] = t.y;
  } else if constexpr (CODEC == CODEC_UNIFORM_8BIT || CODEC == CODEC_QUANTILE_8BIT) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      x[i] = lut[(r.a.x >> (8 * i)) & 0xFF];
      x[4 + i] = lut[(r.a.y >> (8 * i)) & 0xFF];
    }
  } else {  // BLOCKWISE_FP8
    const float s = __ldg(aux + e / kFp8Block);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      x[i] = fp8_e4m3_to_float((r.a.x >> (8 * i)) & 0xFF) * s;
      x[4 + i] = fp8_e4m3_to_float((r.a.y >> (8 * i)) & 0xFF) * s;
    }
  }
}

#endif  // __CUDACC__
}  // namespace dedloc
