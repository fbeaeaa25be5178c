// Shared-layer attention, forward: softmax(Q K^T / sqrt(d) + mask) V for head_dim 64, S <= 512,
// reading Q, K, V straight from the PACKED output of the fused QKV GEMM ([B, S, 3, H, 64] bf16) and writing the
// context in [B, S, H*64] -- no head transposes, no unpack copies. Non-causal, optional key-padding mask.
//
// Because S <= 512 the whole score row fits in tensor memory (128 lanes x 512 fp32 columns), so softmax is exact
// two-pass (no online rescaling):
//   warp 0   TMA: K and V of one (batch, head) [S x 64] once, Q blocks [128 x 64] double-buffered
//   warp 1   tcgen05.mma: S = Q K^T (N = 128 per instruction, 4 k-steps), later O = P V with P read FROM TENSOR
//            MEMORY as the A operand (32 k-steps over the keys, B = V in MN-major form)
//   warp 2   TMEM allocator (all 512 columns)
//   warps 4-7  one thread per query row: row max, exp2, row sum, P packed to bf16 and stored back IN PLACE over the
//            scores (tcgen05.st), then the epilogue O / sum -> bf16 -> global, log-sum-exp saved for backward
// A CTA walks a contiguous range of (batch, head, q-block) units so K/V are fetched once per (batch, head).
#include "gemm_common.cuh"

namespace dedloc {

constexpr int kAttThreads = 256;
constexpr int kAttD = 64;
constexpr int kAttBQ = 128;
constexpr int kAttMaxS = 512;
constexpr int kAttKBytes = kAttMaxS * 128;          // 64 KB: S rows x 128 B
constexpr int kAttQBytes = kAttBQ * 128;            // 16 KB
constexpr int kAttSmem = 2 * kAttKBytes + 2 * kAttQBytes + kAttMaxS * 4 + 256 + 1024;

struct AttFwdParams {
  int B, S, H;
  float scale_log2;          // softmax scale * log2(e)
  float scale;               // softmax scale
  const unsigned char* mask; // [B, S] 1 = attend, 0 = padding; may be null
  __nv_bfloat16* out;        // [B, S, H*64]
  float* lse;                // [B, H, S]
};

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(kAttThreads, 1)
attention_fwd_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ AttFwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = smem + kAttKBytes;
  uint8_t* sQ = smem + 2 * kAttKBytes;                       // 2 buffers
  float* sBias = reinterpret_cast<float*>(smem + 2 * kAttKBytes + 2 * kAttQBytes);   // [S] 0 or -inf
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * kAttKBytes + 2 * kAttQBytes + kAttMaxS * 4);
  uint64_t* kv_full = bars + 0;
  uint64_t* kv_empty = bars + 1;
  uint64_t* q_full = bars + 2;     // [2]
  uint64_t* q_empty = bars + 4;    // [2]
  uint64_t* s_full = bars + 6;
  uint64_t* p_full = bars + 7;
  uint64_t* o_full = bars + 8;
  uint64_t* o_empty = bars + 9;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = p.S, H = p.H;
  const int qblocks = S / kAttBQ;
  const int total_units = p.B * H * qblocks;
  const int per_cta = (total_units + gridDim.x - 1) / gridDim.x;
  const int u0 = blockIdx.x * per_cta;
  const int u1 = min(total_units, u0 + per_cta);

  if (warp == 0 && lane == 0) tma_prefetch_desc(&tmap_qkv);
  if (warp == 1 && lane == 0) {
    mbar_init(kv_full, 1);
    mbar_init(kv_empty, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 1); }
    mbar_init(s_full, 1);
    mbar_init(p_full, 4);
    mbar_init(o_full, 1);
    mbar_init(o_empty, 4);
    mbar_fence_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_o = tmem_base + 448;   // O accumulator: the last 64 columns (free once P has replaced S)

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int cur_bh = -1;
      uint32_t kv_phase = 0;
      int qs = 0;
      uint32_t qphase = 0;
      for (int u = u0; u < u1; ++u) {
        const int bh = u / qblocks, qb = u % qblocks;
        const int b = bh / H, h = bh % H;
        if (bh != cur_bh) {
          mbar_wait(kv_empty, kv_phase ^ 1);
          mbar_expect_tx(kv_full, 2 * S * 128);
          for (int r = 0; r < S; r += 128) {
            tma_load_2d(sK + r * 128, &tmap_qkv, kv_full, (1 * H + h) * kAttD, b * S + r);
            tma_load_2d(sV + r * 128, &tmap_qkv, kv_full, (2 * H + h) * kAttD, b * S + r);
          }
          kv_phase ^= 1;
          cur_bh = bh;
        }
        mbar_wait(&q_empty[qs], qphase ^ 1);
        mbar_expect_tx(&q_full[qs], kAttQBytes);
        tma_load_2d(sQ + qs * kAttQBytes, &tmap_qkv, &q_full[qs], (0 * H + h) * kAttD, b * S + qb * kAttBQ);
        if (++qs == 2) { qs = 0; qphase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t idesc_s = make_idesc_f16(128, 128, 0, 0, 1);   // S: A = Q (K-major), B = K (K-major)
    const uint32_t idesc_o = make_idesc_f16(128, 64, 0, 1, 1);    // O: A = P (TMEM), B = V (MN-major)
    // descriptor low words (address | LBO) + immediates; shared high word; elect.sync single-thread regions
    constexpr uint32_t kHi = 0x40004040u;
    const uint32_t bQ = ((smem_u32(sQ) & 0x3FFFF) >> 4) | (1u << 16);
    const uint32_t bK = ((smem_u32(sK) & 0x3FFFF) >> 4) | (1u << 16);
    const uint32_t bV = ((smem_u32(sV) & 0x3FFFF) >> 4) | ((uint32_t)(128 * 128 >> 4) << 16);
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t tb_o = tb + 448;
    int cur_bh = -1;
    uint32_t kv_phase = 0, unit_phase = 0;
    int qs = 0;
    uint32_t qphase = 0;
    for (int u = u0; u < u1; ++u) {
      const int bh = u / qblocks;
      if (bh != cur_bh) {
        mbar_wait(kv_full, kv_phase);
        kv_phase ^= 1;
        cur_bh = bh;
      }
      const bool last_of_bh = (u + 1 == u1) || ((u + 1) / qblocks != bh);
      mbar_wait(&q_full[qs], qphase);
      mbar_wait(o_empty, unit_phase ^ 1);      // previous unit's O has been read out of tensor memory
      tc_fence_after();
      if (elect_one()) {
        const uint32_t aq = bQ + qs * (kAttQBytes >> 4);
        for (int kb = 0; kb < S / 128; ++kb) {
#pragma unroll
          for (int j = 0; j < kAttD / 16; ++j)
            umma_f16_ss_lohi(tb + kb * 128, aq + j * 2, bK + kb * (128 * 128 >> 4) + j * 2, kHi, idesc_s, j > 0 ? 1u : 0u);
        }
        umma_commit(&q_empty[qs]);
        umma_commit(s_full);
      }
      __syncwarp();
      mbar_wait(p_full, unit_phase);           // softmax warps have written P (bf16) into tensor memory
      tc_fence_after();
      if (elect_one()) {
        // 16 keys = 2 groups of 8 rows x 128 B of V; P occupies 8 TMEM columns per 16 keys
#pragma unroll 8
        for (int k = 0; k < S / 16; ++k)
          umma_f16_ts_lohi(tb_o, tb + k * 8, bV + k * 128, kHi, idesc_o, k > 0 ? 1u : 0u);
        umma_commit(o_full);
        if (last_of_bh) umma_commit(kv_empty);
      }
      __syncwarp();
      unit_phase ^= 1;
      if (++qs == 2) { qs = 0; qphase ^= 1; }
    }
  } else if (warp >= 4) {
    // ===================== softmax + epilogue: one thread per query row =====================
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16);
    uint32_t unit_phase = 0;
    int cur_b = -1;
    const bool has_mask = p.mask != nullptr;
    for (int u = u0; u < u1; ++u) {
      const int bh = u / qblocks, qb = u % qblocks;
      const int b = bh / H, h = bh % H;
      if (has_mask && b != cur_b) {
        // the 128 softmax threads refresh the additive key bias of this batch row
        asm volatile("bar.sync 1, 128;");
        for (int k = threadIdx.x - 128; k < S; k += 128) sBias[k] = p.mask[(size_t)b * S + k] ? 0.f : -INFINITY;
        asm volatile("bar.sync 1, 128;");
        cur_b = b;
      }
      mbar_wait(s_full, unit_phase);
      tc_fence_after();
      // Both passes are software-pipelined: the tcgen05.ld of chunk c+1 is in flight while chunk c is processed
      // (tcgen05.wait::ld waits for every outstanding load, so the next load is issued right after the wait).
      // pass 1: row maximum
      float m = -INFINITY;
      {
        uint32_t ra[32], rb[32];
        tmem_ld_32x32b_x32(t_row, ra);
        for (int c = 0; c < S; c += 64) {
          tmem_ld_wait();
          tmem_ld_32x32b_x32(t_row + c + 32, rb);
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            float v = __uint_as_float(ra[i]);
            if (has_mask) v += sBias[c + i];
            m = fmaxf(m, v);
          }
          tmem_ld_wait();
          if (c + 64 < S) tmem_ld_32x32b_x32(t_row + c + 64, ra);
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            float v = __uint_as_float(rb[i]);
            if (has_mask) v += sBias[c + 32 + i];
            m = fmaxf(m, v);
          }
        }
      }
      const float m_scaled = (m == -INFINITY) ? 0.f : m * p.scale_log2;
      // pass 2: p = exp2(s * scale_log2 - m_scaled); packed bf16 pairs overwrite the scores in place
      float sum = 0.f;
      {
        uint32_t ra[32], rb[32];
        tmem_ld_32x32b_x32(t_row, ra);
        for (int c = 0; c < S; c += 64) {
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            tmem_ld_wait();
            uint32_t (&cur)[32] = half == 0 ? ra : rb;
            const int cc = c + half * 32;
            if (half == 0) tmem_ld_32x32b_x32(t_row + c + 32, rb);
            else if (c + 64 < S) tmem_ld_32x32b_x32(t_row + c + 64, ra);
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              float v0 = __uint_as_float(cur[i]), v1 = __uint_as_float(cur[i + 1]);
              if (has_mask) { v0 += sBias[cc + i]; v1 += sBias[cc + i + 1]; }
              const float e0 = ex2f(fmaf(v0, p.scale_log2, -m_scaled));
              const float e1 = ex2f(fmaf(v1, p.scale_log2, -m_scaled));
              // the row sum uses the bf16-rounded probabilities that the P V product will actually see
              const uint32_t packed = pack_bf16x2(e0, e1);
              const float2 back = unpack_bf16x2(packed);
              sum += back.x + back.y;
              pk[i >> 1] = packed;
            }
            // in-place: columns [cc/2, cc/2 + 16) were read two or more chunks ago; the load in flight reads
            // columns >= cc + 32
            tmem_st_32x32b_x16(t_row + (cc >> 1), pk);
          }
        }
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
      // epilogue
      mbar_wait(o_full, unit_phase);
      tc_fence_after();
      const float inv = sum > 0.f ? 1.f / sum : 0.f;
      const int srow = qb * kAttBQ + row;
      __nv_bfloat16* dst = p.out + ((size_t)(b * S + srow) * H + h) * kAttD;
#pragma unroll
      for (int c = 0; c < kAttD; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(t_row + 448 + c, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 v;
          v.x = pack_bf16x2(__uint_as_float(r[i]) * inv, __uint_as_float(r[i + 1]) * inv);
          v.y = pack_bf16x2(__uint_as_float(r[i + 2]) * inv, __uint_as_float(r[i + 3]) * inv);
          v.z = pack_bf16x2(__uint_as_float(r[i + 4]) * inv, __uint_as_float(r[i + 5]) * inv);
          v.w = pack_bf16x2(__uint_as_float(r[i + 6]) * inv, __uint_as_float(r[i + 7]) * inv);
          *reinterpret_cast<uint4*>(dst + c + i) = v;
        }
      }
      if (p.lse) p.lse[((size_t)b * H + h) * S + srow] = (sum > 0.f) ? m * p.scale + __logf(sum) : -INFINITY;
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(o_empty);
      unit_phase ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace dedloc

using namespace dedloc;

extern "C" int dedloc_attention_fwd(const void* qkv, const unsigned char* mask, void* out, float* lse, int B, int S,
                                    int H, int D, float scale, void* stream) {
  DEDLOC_REQUIRE(D == kAttD, "attention: head_dim must be 64 (got %d)", D);
  DEDLOC_REQUIRE(S % 128 == 0 && S >= 128 && S <= kAttMaxS, "attention: S (%d) must be a multiple of 128 up to %d", S,
                 kAttMaxS);
  DEDLOC_REQUIRE(((uintptr_t)qkv & 15) == 0 && ((uintptr_t)out & 15) == 0, "attention: pointers must be 16-byte aligned");
  { int rc0 = bind_context_for_driver_calls(); if (rc0) return rc0; }
  CUtensorMap tm;
  // packed qkv as a 2-D matrix: rows = B*S tokens, cols = 3*H*64, box = 64 columns x 128 rows
  int rc = make_tmap_2d_16bit(&tm, qkv, (long long)B * S, 3LL * H * D, 3LL * H * D, 64, 128, 1);
  if (rc) return rc;
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    DEDLOC_CUDA_CHECK(cudaGetDevice(&dev));
    DEDLOC_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    DEDLOC_CUDA_CHECK(cudaFuncSetAttribute(attention_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttSmem));
  }
  AttFwdParams p;
  p.B = B; p.S = S; p.H = H;
  p.scale = scale;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.mask = mask;
  p.out = (__nv_bfloat16*)out;
  p.lse = lse;
  const int units = B * H * (S / kAttBQ);
  const int grid = units < sms ? units : sms;
  attention_fwd_kernel<<<grid, kAttThreads, kAttSmem, (cudaStream_t)stream>>>(tm, p);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}
