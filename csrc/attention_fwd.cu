// Shared-layer attention, forward: softmax(Q K^T / sqrt(d) + mask) V for head_dim 64, S <= 512,
// reading Q, K, V straight from the PACKED output of the fused QKV GEMM ([B, S, 3, H, 64] bf16) and writing the
// context in [B, S, H*64] -- no head transposes, no unpack copies. Non-causal, optional key-padding mask.
//
// The exponentials (MUFU, 16 / clock / SM) are the bound at head_dim 64, so the design keeps the MUFU pipe fed:
// two query blocks (128 rows each) of one (batch, head) are in flight, one per softmax warpgroup, and the tensor
// core serves whichever of them is ready. Softmax is online over 128-key score blocks with a LAZY rescale: the exponent
// offset only follows the running row maximum when it grew by more than 2^8, so the accumulated O row (tensor memory)
// is rarely touched after the first block; P goes straight back to the tensor core from tensor memory:
//   warp 0      TMA: K / V blocks of 128 keys into fixed per-(batch, head) slots (a slot is refilled for the next
//               (batch, head) as soon as its last MMA retired), Q blocks double-buffered per warpgroup
//   warp 1      tcgen05.mma issuer, polling the barriers of both warpgroups:
//                 S_w = Q K_kb^T (128 x 128, fp32 in 128 TMEM columns); O_w += P_w V_kb with P (bf16, written in
//                 place over S_w by the softmax warps) as the TMEM A operand and V as an MN-major B operand
//   warp 2      TMEM allocator        warp 3  key-padding bias of the (batch) row, one (batch, head) ahead
//   warps 4-7 / 8-11  one thread per query row: block max, exp2 / row sum / bf16 pack (two reads of the block), rare O
//               rescale, then O / sum -> bf16 -> swizzled shared memory (the Q buffer of the finished block) ->
//               one TMA store per block; log-sum-exp saved for the backward kernel
#include <type_traits>

#include "gemm_common.cuh"

namespace dedloc {

constexpr int kAttThreads = 384;
constexpr int kAttD = 64;
constexpr int kAttBQ = 128;
constexpr int kAttMaxS = 512;
constexpr int kAttTile = 128 * 128;                 // 16 KB: 128 rows x 128 B
constexpr int kAttBiasBytes = kAttMaxS * 4;
// smem: K x 4 slots, V x 4 slots, Q x 2 warpgroups x 2 stages (12 tiles) + bias x 2 + barriers
constexpr int kAttSmem = 12 * kAttTile + 2 * kAttBiasBytes + 512 + 1024;

struct AttFwdParams {
  int B, S, H;
  float scale_log2;          // softmax scale * log2(e)
  float scale;               // softmax scale
  const unsigned char* mask; // [B, S] 1 = attend, 0 = padding; may be null
  float* lse;                // [B, H, S]
};

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <bool HAS_MASK>
__global__ void __launch_bounds__(kAttThreads, 1)
attention_fwd_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_out,
                     const __grid_constant__ AttFwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;                               // 4 slots (key blocks)
  uint8_t* sV = smem + 4 * kAttTile;                // 4 slots
  uint8_t* sQ = smem + 8 * kAttTile;                // [warpgroup][stage]
  uint8_t* sBias = smem + 12 * kAttTile;            // [2][S] 0 or -inf per key
  uint64_t* bars = reinterpret_cast<uint64_t*>(sBias + 2 * kAttBiasBytes);
  uint64_t* k_full = bars + 0;      // [4]
  uint64_t* k_free = bars + 4;      // [4]
  uint64_t* v_full = bars + 8;      // [4]
  uint64_t* v_free = bars + 12;     // [4]
  uint64_t* q_full = bars + 16;     // [wg*2 + stage]
  uint64_t* q_free = bars + 20;     // [wg*2 + stage]
  uint64_t* s_full = bars + 24;     // [wg]
  uint64_t* pv_done = bars + 26;    // [wg] committed after every P V product
  uint64_t* p_full = bars + 28;     // [wg]
  uint64_t* o_full = bars + 30;     // [wg]
  uint64_t* o_free = bars + 32;     // [wg]
  uint64_t* bias_full = bars + 34;  // [2]
  uint64_t* bias_free = bars + 36;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 38);
  uint32_t* blk_pad = reinterpret_cast<uint32_t*>(bars + 40);   // [2][4]: does key block kb of this batch row contain padding?

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = p.S, H = p.H;
  const int nkb = S / 128;                 // key blocks == query blocks of a (batch, head)
  const int total_bh = p.B * H;

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmap_qkv); tma_prefetch_desc(&tmap_out); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 4; ++i) {
      mbar_init(&k_full[i], 1); mbar_init(&k_free[i], 1);
      mbar_init(&v_full[i], 1); mbar_init(&v_free[i], 1);
      mbar_init(&q_full[i], 1); mbar_init(&q_free[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1); mbar_init(&pv_done[i], 1); mbar_init(&p_full[i], 4);
      mbar_init(&o_full[i], 1); mbar_init(&o_free[i], 4);
      mbar_init(&bias_full[i], 1); mbar_init(&bias_free[i], 8);
    }
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
  // columns: warpgroup w owns [192 w, 192 w + 128) for S / P and [192 w + 128, 192 w + 192) for O

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int bc = 0, uq0 = 0, uq1 = 0;      // (batch, head) counter, per-warpgroup block counters
      for (int bh = blockIdx.x; bh < total_bh; bh += gridDim.x, ++bc) {
        const int b = bh / H, h = bh % H;
        auto load_q = [&](int qb) {
          const int w = qb & 1;
          int& uq = (w == 0) ? uq0 : uq1;
          const int slot = w * 2 + (uq & 1);
          mbar_wait(&q_free[slot], ((uq >> 1) & 1) ^ 1);
          mbar_expect_tx(&q_full[slot], kAttTile);
          tma_load_2d(sQ + slot * kAttTile, &tmap_qkv, &q_full[slot], (0 * H + h) * kAttD, b * S + qb * kAttBQ);
          ++uq;
        };
        // order = the order in which the consumers free / need things: first-round Q, K blocks, V blocks, later Q
        load_q(0);
        if (nkb > 1) load_q(1);
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&k_free[kb], (bc & 1) ^ 1);
          mbar_expect_tx(&k_full[kb], kAttTile);
          tma_load_2d(sK + kb * kAttTile, &tmap_qkv, &k_full[kb], (1 * H + h) * kAttD, b * S + kb * 128);
        }
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&v_free[kb], (bc & 1) ^ 1);
          mbar_expect_tx(&v_full[kb], kAttTile);
          tma_load_2d(sV + kb * kAttTile, &tmap_qkv, &v_full[kb], (2 * H + h) * kAttD, b * S + kb * 128);
        }
        for (int qb = 2; qb < nkb; ++qb) load_q(qb);
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: serves whichever warpgroup is ready =====================
    const uint32_t idesc_s = make_idesc_f16(128, 128, 0, 0, 1);   // S: A = Q (K-major), B = K (K-major)
    const uint32_t idesc_o = make_idesc_f16(128, 64, 0, 1, 1);    // O: A = P (TMEM), B = V (MN-major)
    constexpr uint32_t kHi = 0x40004040u;                         // SBO = 1024, version 1, SWIZZLE_128B
    constexpr uint32_t kT = kAttTile >> 4;                        // one tile in descriptor address units
    const uint32_t bQ = ((smem_u32(sQ) & 0x3FFFF) >> 4) | (1u << 16);
    const uint32_t bK = ((smem_u32(sK) & 0x3FFFF) >> 4) | (1u << 16);
    const uint32_t bV = ((smem_u32(sV) & 0x3FFFF) >> 4) | (kT << 16);
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const int my_bh = (total_bh - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    // per-warpgroup progress: (batch, head) index, block index inside it, event index inside the block, counters
    int bi[2] = {0, 0}, qb[2] = {0, 1}, ev[2] = {0, 0}, uq[2] = {0, 0}, n_pfull[2] = {0, 0};
    bool done[2] = {my_bh == 0, my_bh == 0 || nkb < 2};
    uint32_t cnt_k = 0, cnt_v = 0;       // uses of each K / V slot inside the current (batch, head), 8 bits per slot
    auto issue_s = [&](int w, int kb) {  // S_w = Q K_kb^T
      if (elect_one()) {
        const uint32_t aq = bQ + (w * 2 + (uq[w] & 1)) * kT, ak = bK + kb * kT, td = tb + w * 192;
#pragma unroll
        for (int j = 0; j < kAttD / 16; ++j) umma_f16_ss_lohi(td, aq + j * 2, ak + j * 2, kHi, idesc_s, j > 0 ? 1u : 0u);
        umma_commit(&s_full[w]);
      }
      __syncwarp();
    };
    // events of one query block: 0 = S(0); 1 + kb = { O += P(kb) V(kb), then S(kb + 1) }
    while (!(done[0] && done[1])) {
      bool progressed = false;
#pragma unroll
      for (int w = 0; w < 2; ++w) {
        if (done[w]) continue;
        const int e = ev[w];
        if (e == 0) {
          bool ready = mbar_test_wait(&k_full[0], bi[w] & 1);
          if (ready) ready = mbar_test_wait(&q_full[w * 2 + (uq[w] & 1)], (uq[w] >> 1) & 1);
          if (!ready) continue;
          tc_fence_after();
          issue_s(w, 0);
          cnt_k += 1u;
          if ((cnt_k & 0xFF) == (uint32_t)nkb) { cnt_k &= ~0xFFu; if (elect_one()) umma_commit(&k_free[0]); __syncwarp(); }
          ev[w] = 1;
        } else {
          const int kb = e - 1;
          bool ready = mbar_test_wait(&p_full[w], n_pfull[w] & 1);
          if (ready) ready = mbar_test_wait(&v_full[kb], bi[w] & 1);
          if (ready && kb == 0 && uq[w] > 0) ready = mbar_test_wait(&o_free[w], (uq[w] - 1) & 1);
          if (ready && kb < nkb - 1) ready = mbar_test_wait(&k_full[kb + 1], bi[w] & 1);
          if (!ready) continue;
          ++n_pfull[w];
          tc_fence_after();
          if (elect_one()) {
            const uint32_t av = bV + kb * kT, ts = tb + w * 192, to = ts + 128;
            // 16 keys per k-step: 8 TMEM columns of bf16 P, 2 groups of 8 rows x 128 B of V
#pragma unroll
            for (int k = 0; k < 8; ++k) umma_f16_ts_lohi(to, ts + k * 8, av + k * 128, kHi, idesc_o, (kb > 0 || k > 0) ? 1u : 0u);
            umma_commit(&pv_done[w]);                       // O is stable: a lazy rescale may touch it
            if (kb == nkb - 1) umma_commit(&o_full[w]);
          }
          __syncwarp();
          cnt_v += 1u << (8 * kb);
          if (((cnt_v >> (8 * kb)) & 0xFF) == (uint32_t)nkb) {
            cnt_v &= ~(0xFFu << (8 * kb));
            if (elect_one()) umma_commit(&v_free[kb]);
            __syncwarp();
          }
          if (kb < nkb - 1) {
            issue_s(w, kb + 1);                             // overwrites P(kb): issued behind the MMAs that read it
            cnt_k += 1u << (8 * (kb + 1));
            if (((cnt_k >> (8 * (kb + 1))) & 0xFF) == (uint32_t)nkb) {
              cnt_k &= ~(0xFFu << (8 * (kb + 1)));
              if (elect_one()) umma_commit(&k_free[kb + 1]);
              __syncwarp();
            }
            ev[w] = e + 1;
          } else {
            // block finished: next block of this warpgroup
            ev[w] = 0;
            ++uq[w];
            qb[w] += 2;
            if (qb[w] >= nkb) {
              qb[w] = w;
              if (++bi[w] >= my_bh) done[w] = true;
            }
          }
        }
        progressed = true;
      }
      if (!progressed) __nanosleep(20);
    }
  } else if (warp == 3) {
    // ===================== key-padding bias of the batch row, one (batch, head) ahead =====================
    if (HAS_MASK) {
      int bc = 0;
      for (int bh = blockIdx.x; bh < total_bh; bh += gridDim.x, ++bc) {
        const int b = bh / H, st = bc & 1;
        mbar_wait(&bias_free[st], ((bc >> 1) & 1) ^ 1);
        float* bias = reinterpret_cast<float*>(sBias + st * kAttBiasBytes);
        uint32_t pad_bits = 0;
        for (int k = lane; k < S; k += 32) {
          const bool ok = p.mask[(size_t)b * S + k] != 0;
          bias[k] = ok ? 0.f : -INFINITY;
          if (!ok) pad_bits |= 1u << (k >> 7);
        }
        pad_bits = __reduce_or_sync(0xffffffffu, pad_bits);
        if (lane < 4) blk_pad[st * 4 + lane] = (pad_bits >> lane) & 1u;   // blocks without padding skip the bias adds
        __syncwarp();
        if (lane == 0) mbar_arrive(&bias_full[st]);
      }
    }
  } else if (warp >= 4) {
    // ===================== softmax + epilogue: one thread per query row, one warpgroup per query block =====================
    const int w = (warp - 4) >> 2;
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t t_s = tmem_base + w * 192 + ((uint32_t)(quad * 32) << 16);
    const uint32_t t_o = t_s + 128;
    const bool store_leader = (quad == 0 && lane == 0);
    int n_sfull = 0, uq = 0, bc = 0, pending_slot = -1;
    for (int bh = blockIdx.x; bh < total_bh; bh += gridDim.x, ++bc) {
      const int b = bh / H, h = bh % H;
      const uint32_t bias_a = smem_u32(sBias) + (bc & 1) * kAttBiasBytes;
      if (HAS_MASK) mbar_wait(&bias_full[bc & 1], (bc >> 1) & 1);
      for (int qb = w; qb < nkb; qb += 2, ++uq) {
        const int slot = w * 2 + (uq & 1);
        // the TMA store of the previous block read its staging buffer (= that block's Q stage): hand it back
        if (store_leader && pending_slot >= 0) {
          tma_store_wait_read<0>();
          mbar_arrive(&q_free[pending_slot]);
          pending_slot = -1;
        }
        // ---------------- online softmax over the key blocks, lazy rescale ----------------
        // m_use = exponent offset in use (raw-score units). It follows the running row maximum only when that grew by
        // more than 8 in the log2 domain (p <= 256: harmless in bf16 P / fp32 sum and O), so the read-modify-write of
        // the O row in tensor memory is rare after the first block. The final O / sum and lse are exact either way.
        float m_use = -INFINITY, sum = 0.f;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&s_full[w], n_sfull & 1);
          ++n_sfull;
          tc_fence_after();
          const bool padded = HAS_MASK && blk_pad[(bc & 1) * 4 + kb] != 0;
          uint32_t ra[32], rb[32];
          // (a) block maximum
          float bm = -INFINITY;
          tmem_ld_32x32b_x32(t_s, ra);
          tmem_ld_32x32b_x32(t_s + 32, rb);
          auto max_pass = [&](auto PAD) {
            constexpr bool kPad = decltype(PAD)::value;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              tmem_ld_wait();       // waits for every outstanding load: chunk c and the prefetched chunk c+1
              uint32_t (&cur)[32] = (c & 1) ? rb : ra;
#pragma unroll
              for (int i = 0; i < 32; i += 4) {
                float v0 = __uint_as_float(cur[i]), v1 = __uint_as_float(cur[i + 1]);
                float v2 = __uint_as_float(cur[i + 2]), v3 = __uint_as_float(cur[i + 3]);
                if (kPad) {
                  const float4 bb = lds_f4(bias_a + (kb * 128 + c * 32 + i) * 4);
                  v0 += bb.x; v1 += bb.y; v2 += bb.z; v3 += bb.w;
                }
                bm = fmaxf(fmaxf(bm, v0), fmaxf(v1, fmaxf(v2, v3)));
              }
              if (c + 2 < 4) tmem_ld_32x32b_x32(t_s + (c + 2) * 32, cur);   // refill the buffer just consumed
            }
          };
          if (padded) max_pass(std::true_type{}); else max_pass(std::false_type{});
          // (b) exponent offset; rescale what has been accumulated when it has to move
          const float m_new = fmaxf(m_use, bm);
          if (kb == 0) {
            m_use = m_new;
          } else {
            const bool need = (m_new - m_use) * p.scale_log2 > 8.f;      // (-inf) - (-inf) = NaN compares false
            if (__any_sync(0xffffffffu, need)) {
              mbar_wait(&pv_done[w], (uq * nkb + kb - 1) & 1);            // O holds the products of blocks < kb
              tc_fence_after();
              const float f = need ? ex2f((m_use - m_new) * p.scale_log2) : 1.f;
              uint32_t o0[32], o1[32];
              tmem_ld_32x32b_x32(t_o, o0);
              tmem_ld_32x32b_x32(t_o + 32, o1);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                o0[i] = __float_as_uint(__uint_as_float(o0[i]) * f);
                o1[i] = __float_as_uint(__uint_as_float(o1[i]) * f);
              }
              tmem_st_32x32b_x16(t_o, reinterpret_cast<const uint32_t(&)[16]>(o0[0]));
              tmem_st_32x32b_x16(t_o + 16, reinterpret_cast<const uint32_t(&)[16]>(o0[16]));
              tmem_st_32x32b_x16(t_o + 32, reinterpret_cast<const uint32_t(&)[16]>(o1[0]));
              tmem_st_32x32b_x16(t_o + 48, reinterpret_cast<const uint32_t(&)[16]>(o1[16]));
              sum *= f;
              if (need) m_use = m_new;
            }
          }
          const float m_scaled = (m_use == -INFINITY) ? 0.f : m_use * p.scale_log2;
          // (c) p = exp2(s * scale_log2 - m_scaled), bf16 pairs in place over the scores (second read of the block)
          tmem_ld_32x32b_x32(t_s, ra);
          tmem_ld_32x32b_x32(t_s + 32, rb);
          auto exp_pass = [&](auto PAD) {
            constexpr bool kPad = decltype(PAD)::value;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              tmem_ld_wait();
              uint32_t (&cur)[32] = (c & 1) ? rb : ra;
              uint32_t pk[16];
#pragma unroll
              for (int i = 0; i < 32; i += 4) {
                float v0 = __uint_as_float(cur[i]), v1 = __uint_as_float(cur[i + 1]);
                float v2 = __uint_as_float(cur[i + 2]), v3 = __uint_as_float(cur[i + 3]);
                if (kPad) {
                  const float4 bb = lds_f4(bias_a + (kb * 128 + c * 32 + i) * 4);
                  v0 += bb.x; v1 += bb.y; v2 += bb.z; v3 += bb.w;
                }
                const float e0 = ex2f(fmaf(v0, p.scale_log2, -m_scaled));
                const float e1 = ex2f(fmaf(v1, p.scale_log2, -m_scaled));
                const float e2 = ex2f(fmaf(v2, p.scale_log2, -m_scaled));
                const float e3 = ex2f(fmaf(v3, p.scale_log2, -m_scaled));
                sum += (e0 + e1) + (e2 + e3);
                pk[i >> 1] = pack_bf16x2(e0, e1);
                pk[(i >> 1) + 1] = pack_bf16x2(e2, e3);
              }
              // in place: bf16 columns [16c, 16c+16) overlay fp32 columns of chunk c/2, which has been read; the
              // loads in flight read columns >= 32 (c + 1)
              tmem_st_32x32b_x16(t_s + c * 16, pk);
              if (c + 2 < 4) tmem_ld_32x32b_x32(t_s + (c + 2) * 32, cur);
            }
          };
          if (padded) exp_pass(std::true_type{}); else exp_pass(std::false_type{});
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&p_full[w]);
        }
        const float m = m_use;
        // ---------------- epilogue: O / sum -> bf16 -> swizzled smem (the Q stage of this block) -> TMA store ----------------
        mbar_wait(&o_full[w], uq & 1);
        tc_fence_after();
        uint32_t r0[32], r1[32];
        tmem_ld_32x32b_x32(t_o, r0);
        tmem_ld_32x32b_x32(t_o + 32, r1);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&o_free[w]);
        const float inv = sum > 0.f ? 1.f / sum : 0.f;
        const uint32_t stg = smem_u32(sQ) + slot * kAttTile + row * 128;
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          const int i = ch * 8;
          sts_u4(stg + ((ch ^ (row & 7)) << 4),
                 pack_bf16x2(__uint_as_float(r0[i]) * inv, __uint_as_float(r0[i + 1]) * inv),
                 pack_bf16x2(__uint_as_float(r0[i + 2]) * inv, __uint_as_float(r0[i + 3]) * inv),
                 pack_bf16x2(__uint_as_float(r0[i + 4]) * inv, __uint_as_float(r0[i + 5]) * inv),
                 pack_bf16x2(__uint_as_float(r0[i + 6]) * inv, __uint_as_float(r0[i + 7]) * inv));
          sts_u4(stg + (((ch + 4) ^ (row & 7)) << 4),
                 pack_bf16x2(__uint_as_float(r1[i]) * inv, __uint_as_float(r1[i + 1]) * inv),
                 pack_bf16x2(__uint_as_float(r1[i + 2]) * inv, __uint_as_float(r1[i + 3]) * inv),
                 pack_bf16x2(__uint_as_float(r1[i + 4]) * inv, __uint_as_float(r1[i + 5]) * inv),
                 pack_bf16x2(__uint_as_float(r1[i + 6]) * inv, __uint_as_float(r1[i + 7]) * inv));
        }
        if (p.lse) p.lse[((size_t)b * H + h) * S + qb * kAttBQ + row] = (sum > 0.f) ? m * p.scale + __logf(sum) : -INFINITY;
        fence_proxy_async_smem();
        asm volatile("bar.sync %0, 128;" ::"r"(1 + w) : "memory");
        if (store_leader) {
          tma_store_2d(&tmap_out, sQ + slot * kAttTile, h * kAttD, b * S + qb * kAttBQ);
          tma_store_commit();
          pending_slot = slot;
        }
      }
      if (HAS_MASK) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&bias_free[bc & 1]);
      }
    }
    if (store_leader) tma_store_wait<0>();
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
  CUtensorMap tm, to;
  // packed qkv as a 2-D matrix: rows = B*S tokens, cols = 3*H*64, box = 64 columns x 128 rows
  int rc = make_tmap_2d_16bit(&tm, qkv, (long long)B * S, 3LL * H * D, 3LL * H * D, 64, 128, 1);
  if (rc) return rc;
  rc = make_tmap_2d_16bit(&to, out, (long long)B * S, (long long)H * D, (long long)H * D, 64, 128, 1);
  if (rc) return rc;
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    DEDLOC_CUDA_CHECK(cudaGetDevice(&dev));
    DEDLOC_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    DEDLOC_CUDA_CHECK(cudaFuncSetAttribute(attention_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttSmem));
    DEDLOC_CUDA_CHECK(cudaFuncSetAttribute(attention_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttSmem));
  }
  AttFwdParams p;
  p.B = B; p.S = S; p.H = H;
  p.scale = scale;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.mask = mask;
  p.lse = lse;
  const int grid = B * H < sms ? B * H : sms;
  if (mask) attention_fwd_kernel<true><<<grid, kAttThreads, kAttSmem, (cudaStream_t)stream>>>(tm, to, p);
  else attention_fwd_kernel<false><<<grid, kAttThreads, kAttSmem, (cudaStream_t)stream>>>(tm, to, p);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}
