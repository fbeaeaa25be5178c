// Shared pieces of the tcgen05 GEMM kernels: kernel parameter block and the fused epilogue applied to 32
// accumulator columns of one output row (bias / gelu_new / gelu' / residual / scaling, bf16 or fp32 stores, or
// fp32 red.add for split-K weight-gradient accumulation).
#pragma once
#include <cuda.h>
#include "common.cuh"
#include "../include/dedloc_b200.h"

namespace dedloc {

struct GemmKernelParams {
  int M, N, K;
  int a_mn, b_mn;       // operand majors
  int splits;           // split-K factor (output must be an atomic-add epilogue when > 1)
  int kb_per_split;
  int epi;              // DEDLOC_EPI_* flags
  int out_kind;         // 0 bf16 store, 1 fp32 store, 2 fp32 atomic add
  int fmt;              // 1 = bf16 inputs, 0 = fp16 inputs
  int a_mn3d, b_mn3d;   // MN-major operand described by a 3-D tensor map (one TMA per stage instead of one per 64 columns)
  float alpha;
  void* C;              // [M, ldc]
  long long ldc;
  const void* bias;     // [N], bf16 or fp32 (epi flag)
  const void* residual; // [M, ld_res] bf16
  long long ld_res;
  const void* aux_in;   // [M, ld_aux] bf16: pre-activation for the gelu' epilogue
  long long ld_aux;
  void* pre_out;        // [M, ld_pre] bf16: pre-activation saved by the forward gelu epilogue
  long long ld_pre;
};


// Grouped tile order: consecutive tile ids walk kGroupM row-tiles before moving to the next column-tile, so the
// CTAs resident at the same time share both A row-panels and B column-panels out of L2 (instead of streaming a
// different A panel per CTA against one B panel).
constexpr int kGroupM = 8;
__device__ __forceinline__ void tile_coords(int rem, int m_tiles, int n_tiles, int& mt, int& nt) {
  const int per_group = kGroupM * n_tiles;
  const int group = rem / per_group;
  const int first_m = group * kGroupM;
  const int gm = min(m_tiles - first_m, kGroupM);
  const int in_group = rem - group * per_group;
  mt = first_m + in_group % gm;
  nt = in_group / gm;
}

__device__ __forceinline__ void unpack8_bf16(const uint4& u, float (&o)[8]) {
  float2 t;
  t = unpack_bf16x2(u.x); o[0] = t.x; o[1] = t.y;
  t = unpack_bf16x2(u.y); o[2] = t.x; o[3] = t.y;
  t = unpack_bf16x2(u.z); o[4] = t.x; o[5] = t.y;
  t = unpack_bf16x2(u.w); o[6] = t.x; o[7] = t.y;
}

// Epilogue math on 32 accumulator columns [col0, col0+32) of output row `row`:
//   v = acc * alpha (+ bias) ; pre = v ; (gelu_new | * gelu'(aux)) ; (+ residual)
// `pre` (the pre-activation, only meaningful with SAVE_PRE) is returned separately so that the caller decides how
// to store it. Loads of row-indexed operands are skipped for rows beyond M.
__device__ __forceinline__ void gemm_epilogue_math(const GemmKernelParams& p, int row, bool row_ok, int col0,
                                                   const uint32_t (&r)[32], float (&v)[32], float (&pre)[32]) {
  const bool full_cols = (col0 + 32 <= p.N);
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]) * p.alpha;
  if (p.epi & DEDLOC_EPI_BIAS) {
    if (p.epi & DEDLOC_EPI_BIAS_F32) {
      const float* b = reinterpret_cast<const float*>(p.bias) + col0;
      if (full_cols) {
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const float4 t = __ldg(reinterpret_cast<const float4*>(b + i));   // same address in every lane: broadcast
          v[i] += t.x; v[i + 1] += t.y; v[i + 2] += t.z; v[i + 3] += t.w;
        }
      } else {
        _Pragma("unroll") for (int i = 0; i < 32; ++i) if (col0 + i < p.N) v[i] += __ldg(b + i);
      }
    } else {
      const __nv_bfloat16* b = reinterpret_cast<const __nv_bfloat16*>(p.bias) + col0;
      if (full_cols) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          float t[8];
          unpack8_bf16(__ldg(reinterpret_cast<const uint4*>(b + i)), t);
#pragma unroll
          for (int k = 0; k < 8; ++k) v[i + k] += t[k];
        }
      } else {
        _Pragma("unroll") for (int i = 0; i < 32; ++i) if (col0 + i < p.N) v[i] += __bfloat162float(b[i]);
      }
    }
  }
  if ((p.epi & DEDLOC_EPI_GELU) && (p.epi & DEDLOC_EPI_SAVE_DGELU)) {
#pragma unroll
    for (int i = 0; i < 32; ++i) gelu_tanh_both(v[i], v[i], pre[i]);
  } else {
    if (p.epi & DEDLOC_EPI_SAVE_PRE) {
#pragma unroll
      for (int i = 0; i < 32; ++i) pre[i] = v[i];
    }
    if (p.epi & DEDLOC_EPI_GELU) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = gelu_tanh(v[i]);
    }
  }
  if ((p.epi & (DEDLOC_EPI_DGELU | DEDLOC_EPI_MUL_AUX)) && row_ok) {
    const bool mul_aux = (p.epi & DEDLOC_EPI_MUL_AUX) != 0;
    const __nv_bfloat16* ai = reinterpret_cast<const __nv_bfloat16*>(p.aux_in) + (size_t)row * p.ld_aux + col0;
    if (full_cols) {
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        float t[8];
        unpack8_bf16(*reinterpret_cast<const uint4*>(ai + i), t);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[i + k] *= mul_aux ? t[k] : gelu_tanh_grad(t[k]);
      }
    } else {
      _Pragma("unroll") for (int i = 0; i < 32; ++i) if (col0 + i < p.N) {
        const float a = __bfloat162float(ai[i]);
        v[i] *= mul_aux ? a : gelu_tanh_grad(a);
      }
    }
  }
  if ((p.epi & DEDLOC_EPI_RESIDUAL) && row_ok) {
    const __nv_bfloat16* rs = reinterpret_cast<const __nv_bfloat16*>(p.residual) + (size_t)row * p.ld_res + col0;
    if (full_cols) {
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        float t[8];
        unpack8_bf16(*reinterpret_cast<const uint4*>(rs + i), t);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[i + k] += t[k];
      }
    } else {
      _Pragma("unroll") for (int i = 0; i < 32; ++i) if (col0 + i < p.N) v[i] += __bfloat162float(rs[i]);
    }
  }
}

// bf16 row-segment store (32 columns) straight to global memory.
__device__ __forceinline__ void store_bf16_32(__nv_bfloat16* dst, const float (&v)[32], bool full_cols, int valid) {
  if (full_cols) {
#pragma unroll
    for (int i = 0; i < 32; i += 8)
      *reinterpret_cast<uint4*>(dst + i) = make_uint4(pack_bf16x2(v[i], v[i + 1]), pack_bf16x2(v[i + 2], v[i + 3]),
                                                      pack_bf16x2(v[i + 4], v[i + 5]), pack_bf16x2(v[i + 6], v[i + 7]));
  } else {
    _Pragma("unroll") for (int i = 0; i < 32; ++i) if (i < valid) dst[i] = __float2bfloat16(v[i]);
  }
}

// Direct (register -> global) store of the epilogue result for every output kind.
__device__ __forceinline__ void gemm_epilogue_store(const GemmKernelParams& p, int row, int col0, const float (&v)[32]) {
  const bool full_cols = (col0 + 32 <= p.N);
  const int valid = p.N - col0;
  if (p.out_kind == 0) {
    store_bf16_32(reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)row * p.ldc + col0, v, full_cols, valid);
  } else if (p.out_kind == 1) {
    float* co = reinterpret_cast<float*>(p.C) + (size_t)row * p.ldc + col0;
    if (full_cols) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) *reinterpret_cast<float4*>(co + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
    } else {
      _Pragma("unroll") for (int i = 0; i < 32; ++i) if (i < valid) co[i] = v[i];
    }
  } else {
    float* co = reinterpret_cast<float*>(p.C) + (size_t)row * p.ldc + col0;
    if (full_cols) {
#pragma unroll
      for (int i = 0; i < 32; i += 4)
        asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(co + i), "f"(v[i]), "f"(v[i + 1]),
                     "f"(v[i + 2]), "f"(v[i + 3])
                     : "memory");
    } else {
      _Pragma("unroll") for (int i = 0; i < 32; ++i) if (i < valid) atomicAdd(co + i, v[i]);
    }
  }
}

// r[32]: raw fp32 accumulator bits of columns [col0, col0 + 32) of output row `row`; everything stored directly.
__device__ __forceinline__ void gemm_epilogue_32cols(const GemmKernelParams& p, int row, bool row_ok, int col0,
                                                     const uint32_t (&r)[32]) {
  float v[32], pre[32];
  gemm_epilogue_math(p, row, row_ok, col0, r, v, pre);
  if (!row_ok) return;
  if (p.epi & DEDLOC_EPI_SAVE_PRE)
    store_bf16_32(reinterpret_cast<__nv_bfloat16*>(p.pre_out) + (size_t)row * p.ld_pre + col0, pre, col0 + 32 <= p.N,
                  p.N - col0);
  gemm_epilogue_store(p, row, col0, v);
}

// Host helper shared by both kernels (defined in gemm_tcgen05.cu).
int make_tmap_2d_16bit(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld,
                       int box_cols, int box_rows, int is_bf16);
int bind_context_for_driver_calls();
// MN-major operand [k_rows, mn_cols] (mn contiguous, mn_cols % 64 == 0) as (64 elements, k rows, mn_cols / 64 groups):
// one box of (64, box_k, groups) lands in shared memory as `groups` consecutive [box_k x 128 B] swizzled tiles.
int make_tmap_mn_3d_16bit(CUtensorMap* map, const void* base, long long k_rows, long long mn_cols, long long ld,
                          int box_k, int groups, int is_bf16);

}  // namespace dedloc
