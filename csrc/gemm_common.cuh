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


// r[32]: raw fp32 accumulator bits of columns [col0, col0 + 32) of output row `row`.
__device__ __forceinline__ void gemm_epilogue_32cols(const GemmKernelParams& p, int row, bool row_ok, int col0,
                                                     const uint32_t (&r)[32]) {
    const bool has_bias = (p.epi & DEDLOC_EPI_BIAS) != 0;
    const bool bias_f32 = (p.epi & DEDLOC_EPI_BIAS_F32) != 0;
    const bool do_gelu = (p.epi & DEDLOC_EPI_GELU) != 0;
    const bool do_dgelu = (p.epi & DEDLOC_EPI_DGELU) != 0;
    const bool has_res = (p.epi & DEDLOC_EPI_RESIDUAL) != 0;
    const bool save_pre = (p.epi & DEDLOC_EPI_SAVE_PRE) != 0;
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]) * p.alpha;
    const bool full_cols = (col0 + 32 <= p.N);
    if (has_bias) {
      if (bias_f32) {
        const float* b = reinterpret_cast<const float*>(p.bias) + col0;
#pragma unroll
        for (int i = 0; i < 32; ++i) if (full_cols || col0 + i < p.N) v[i] += __ldg(b + i);
      } else {
        const __nv_bfloat16* b = reinterpret_cast<const __nv_bfloat16*>(p.bias) + col0;
#pragma unroll
        for (int i = 0; i < 32; ++i) if (full_cols || col0 + i < p.N) v[i] += __bfloat162float(b[i]);
      }
    }
    if (row_ok) {
      if (save_pre) {
        __nv_bfloat16* po = reinterpret_cast<__nv_bfloat16*>(p.pre_out) + (size_t)row * p.ld_pre + col0;
        if (full_cols) {
#pragma unroll
          for (int i = 0; i < 32; i += 8)
            *reinterpret_cast<uint4*>(po + i) = make_uint4(pack_bf16x2(v[i], v[i + 1]), pack_bf16x2(v[i + 2], v[i + 3]),
                                                           pack_bf16x2(v[i + 4], v[i + 5]), pack_bf16x2(v[i + 6], v[i + 7]));
        } else {
          for (int i = 0; i < 32 && col0 + i < p.N; ++i) po[i] = __float2bfloat16(v[i]);
        }
      }
      if (do_gelu) {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = gelu_tanh(v[i]);
      }
      if (do_dgelu) {
        const __nv_bfloat16* ai = reinterpret_cast<const __nv_bfloat16*>(p.aux_in) + (size_t)row * p.ld_aux + col0;
        if (full_cols) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            const uint4 u = *reinterpret_cast<const uint4*>(ai + i);
            float2 t;
            t = unpack_bf16x2(u.x); v[i] *= gelu_tanh_grad(t.x); v[i + 1] *= gelu_tanh_grad(t.y);
            t = unpack_bf16x2(u.y); v[i + 2] *= gelu_tanh_grad(t.x); v[i + 3] *= gelu_tanh_grad(t.y);
            t = unpack_bf16x2(u.z); v[i + 4] *= gelu_tanh_grad(t.x); v[i + 5] *= gelu_tanh_grad(t.y);
            t = unpack_bf16x2(u.w); v[i + 6] *= gelu_tanh_grad(t.x); v[i + 7] *= gelu_tanh_grad(t.y);
          }
        } else {
          for (int i = 0; i < 32 && col0 + i < p.N; ++i) v[i] *= gelu_tanh_grad(__bfloat162float(ai[i]));
        }
      }
      if (has_res) {
        const __nv_bfloat16* rs = reinterpret_cast<const __nv_bfloat16*>(p.residual) + (size_t)row * p.ld_res + col0;
        if (full_cols) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            const uint4 u = *reinterpret_cast<const uint4*>(rs + i);
            float2 t;
            t = unpack_bf16x2(u.x); v[i] += t.x; v[i + 1] += t.y;
            t = unpack_bf16x2(u.y); v[i + 2] += t.x; v[i + 3] += t.y;
            t = unpack_bf16x2(u.z); v[i + 4] += t.x; v[i + 5] += t.y;
            t = unpack_bf16x2(u.w); v[i + 6] += t.x; v[i + 7] += t.y;
          }
        } else {
          for (int i = 0; i < 32 && col0 + i < p.N; ++i) v[i] += __bfloat162float(rs[i]);
        }
      }
      if (p.out_kind == 0) {
        __nv_bfloat16* co = reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)row * p.ldc + col0;
        if (full_cols) {
#pragma unroll
          for (int i = 0; i < 32; i += 8)
            *reinterpret_cast<uint4*>(co + i) = make_uint4(pack_bf16x2(v[i], v[i + 1]), pack_bf16x2(v[i + 2], v[i + 3]),
                                                           pack_bf16x2(v[i + 4], v[i + 5]), pack_bf16x2(v[i + 6], v[i + 7]));
        } else {
          for (int i = 0; i < 32 && col0 + i < p.N; ++i) co[i] = __float2bfloat16(v[i]);
        }
      } else if (p.out_kind == 1) {
        float* co = reinterpret_cast<float*>(p.C) + (size_t)row * p.ldc + col0;
        if (full_cols) {
#pragma unroll
          for (int i = 0; i < 32; i += 4) *reinterpret_cast<float4*>(co + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        } else {
          for (int i = 0; i < 32 && col0 + i < p.N; ++i) co[i] = v[i];
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
          for (int i = 0; i < 32 && col0 + i < p.N; ++i) atomicAdd(co + i, v[i]);
        }
      }
    }
}

// Host helper shared by both kernels (defined in gemm_tcgen05.cu).
int make_tmap_2d_16bit(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld,
                       int box_cols, int box_rows, int is_bf16);
int bind_context_for_driver_calls();

}  // namespace dedloc
