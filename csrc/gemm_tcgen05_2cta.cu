// 2-CTA (cta_group::2) variant of the tcgen05 GEMM: a CTA pair on one TPC computes a 256 x 256 output tile.
//
// Why: with one CTA per 128 x 256 tile every CTA streams its own A (16 KB) and B (32 KB) per 64-wide k-block --
// 85 FLOP per byte of L2->SMEM traffic, and the v1 kernel measured ~0.9-1.1 PFLOP/s, i.e. bound by the L2 fabric
// (~6300 B/clk chip-wide), not by the tensor cores. Pairing two CTAs lets each one load ITS 128 rows of A and only
// HALF of B (128 of the 256 columns); tcgen05.mma.cta_group::2 reads both halves across the pair: 128 FLOP/byte.
//
// Roles per CTA (352 threads, 5 smem stages): warp 0 TMA producer (both CTAs; transaction bytes of both land on the
// LEADER's full barrier; MN-major operands through a 3-D tensor map, one request per stage and operand), warp 1 MMA
// issuer (leader CTA only, elect.sync region with precomputed descriptor words; tcgen05.commit multicasts to both
// CTAs' barriers), warp 2 TMEM allocator (cta_group::2, both CTAs), warps 3..10 epilogue (two per TMEM lane quadrant:
// each CTA drains its own 128 accumulator rows, 128 columns per warp, and arrives on the leader's tmem_empty
// barrier). Same operand layouts / fused epilogues as gemm_tcgen05.cu; K tails are zero-filled by TMA.
#include "gemm_common.cuh"

namespace dedloc {

constexpr int k2Threads = 352;    // TMA, MMA, TMEM-alloc + 8 epilogue warps (11 warps: 184 registers per thread)
constexpr int k2BM = 128;       // rows per CTA (256 per pair)
constexpr int k2BN = 256;       // columns per pair; each CTA stages 128 of them
constexpr int k2BK = 64;
constexpr int k2Stages = 5;
constexpr int k2ABytes = k2BM * k2BK * 2;          // 16 KB
constexpr int k2BBytes = (k2BN / 2) * k2BK * 2;    // 16 KB
constexpr int k2StageBytes = k2ABytes + k2BBytes;  // 32 KB per CTA per stage
constexpr int k2EpiWarps = 8;      // two per TMEM lane quadrant: columns [0,128) and [128,256) of the tile
constexpr int k2StagingBytes = k2EpiWarps * 2 * 4096;  // per epilogue warp: two 32-row x 128-byte TMA-store buffers
constexpr int k2BiasBytes = k2BN * 4;                   // fp32 bias of the current tile's columns
constexpr int k2SmemTotal = k2Stages * k2StageBytes + k2StagingBytes + k2BiasBytes + 128 + 1024;
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;     // clears the CTA-pair bit of a shared::cluster address

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const void* tmap, uint64_t* leader_bar, int c0, int c1) {
  const uint32_t bar = smem_u32(leader_bar) & kPeerBitMask;   // transaction bytes update CTA 0's barrier
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
// descriptors as a 32-bit low word (address | LBO) plus the high word shared by all of them
__device__ __forceinline__ void tma_load_3d_2sm(void* smem_dst, const void* tmap, uint64_t* leader_bar, int c0, int c1, int c2) {
  const uint32_t bar = smem_u32(leader_bar) & kPeerBitMask;
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void umma_f16_ss_2cta(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t hi,
                                                 uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %5, 0;\n\t"
      "mov.b64 da, {%1, %3};\n\t"
      "mov.b64 db, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, p;\n\t}" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive (once MMAs retire) on the barrier at this smem offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_2cta_mc(uint64_t* bar) {
  const uint16_t mask = 0b11;
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(mask)
               : "memory");
}
// arrive on the barrier at this smem offset in CTA 0 of the pair
__device__ __forceinline__ void mbar_arrive_leader(uint64_t* bar) {
  const uint32_t addr = smem_u32(bar) & kPeerBitMask;
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(addr) : "memory");
}
__device__ __forceinline__ void tmem_alloc_2cta(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2cta(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

__global__ void __launch_bounds__(k2Threads, 1)
gemm_tcgen05_2cta_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                         const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ CUtensorMap tmap_pre,
                         const __grid_constant__ GemmKernelParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* staging = smem + k2Stages * k2StageBytes;
  float* sbias = reinterpret_cast<float*>(staging + k2StagingBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + k2StagingBytes + k2BiasBytes);
  uint64_t* full_bar = bars;                        // [k2Stages]  (leader's copy is the live one)
  uint64_t* empty_bar = bars + k2Stages;            // [k2Stages]  (live in both CTAs)
  uint64_t* tmem_full = bars + 2 * k2Stages;        // [2]         (live in both CTAs)
  uint64_t* tmem_empty = bars + 2 * k2Stages + 2;   // [2]         (leader's copy is the live one)
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(bars + 2 * k2Stages + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta_rank = cluster_ctarank();
  const bool leader = cta_rank == 0;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

  const int m_tiles = (p.M + 2 * k2BM - 1) / (2 * k2BM);
  const int n_tiles = (p.N + k2BN - 1) / k2BN;
  const int kb_total = (p.K + k2BK - 1) / k2BK;   // a K tail is zero-filled by TMA (out-of-bounds box elements)
  const int total_tiles = m_tiles * n_tiles * p.splits;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    if (p.out_kind == 0) tma_prefetch_desc(&tmap_c);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < k2Stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&tmem_full[s], 1); mbar_init(&tmem_empty[s], 2 * k2EpiWarps); }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc_2cta(tmem_base_slot, 2 * k2BN);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // the peer's barriers are initialised before any remote arrive / complete_tx targets them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < total_tiles; tile += num_clusters) {
        const int split = tile / (m_tiles * n_tiles);
        const int rem = tile - split * (m_tiles * n_tiles);
        int mt, nt;
        tile_coords(rem, m_tiles, n_tiles, mt, nt);
        const int m0 = mt * (2 * k2BM) + (int)cta_rank * k2BM;
        const int n0 = nt * k2BN + (int)cta_rank * (k2BN / 2);
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb_total, kb0 + p.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * k2StageBytes;
          uint8_t* sb = sa + k2ABytes;
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * k2StageBytes);
          const int k0 = kb * k2BK;
          if (!p.a_mn) {
            tma_load_2d_2sm(sa, &tmap_a, &full_bar[stage], k0, m0);
          } else if (p.a_mn3d) {
            tma_load_3d_2sm(sa, &tmap_a, &full_bar[stage], 0, k0, m0 >> 6);   // both 64-row groups in one request
          } else {
#pragma unroll
            for (int i = 0; i < k2BM / 64; ++i)
              tma_load_2d_2sm(sa + i * (k2BK * 128), &tmap_a, &full_bar[stage], m0 + 64 * i, k0);
          }
          if (!p.b_mn) {
            tma_load_2d_2sm(sb, &tmap_b, &full_bar[stage], k0, n0);
          } else if (p.b_mn3d) {
            tma_load_3d_2sm(sb, &tmap_b, &full_bar[stage], 0, k0, n0 >> 6);
          } else {
#pragma unroll
            for (int i = 0; i < (k2BN / 2) / 64; ++i)
              tma_load_2d_2sm(sb + i * (k2BK * 128), &tmap_b, &full_bar[stage], n0 + 64 * i, k0);
          }
          if (++stage == k2Stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader) {
      const uint32_t idesc = make_idesc_f16(2 * k2BM, k2BN, p.a_mn, p.b_mn, p.fmt);
      // Issue cost: descriptors are (precomputed low word + immediate) and the single-thread region is entered
      // through elect.sync, which keeps every MMA operand in uniform registers (a dozen instructions fewer per MMA
      // than a lane == 0 branch).
      constexpr uint32_t kHi = 0x40004040u;                 // SBO = 1024, version 1, SWIZZLE_128B
      const uint32_t a_step = p.a_mn ? 128u : 2u, b_step = p.b_mn ? 128u : 2u;     // one k16 step, in 16-byte units
      const uint32_t a0 = ((smem_u32(smem) & 0x3FFFF) >> 4) | ((p.a_mn ? (uint32_t)(k2BK * 128) >> 4 : 1u) << 16);
      const uint32_t b0 = (((smem_u32(smem) + k2ABytes) & 0x3FFFF) >> 4) | ((p.b_mn ? (uint32_t)(k2BK * 128) >> 4 : 1u) << 16);
      const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = cluster_id; tile < total_tiles; tile += num_clusters) {
        const int split = tile / (m_tiles * n_tiles);
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb_total, kb0 + p.kb_per_split);
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tb + acc * k2BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t sa = a0 + stage * (k2StageBytes >> 4), sb = b0 + stage * (k2StageBytes >> 4);
            umma_f16_ss_2cta(tmem_d, sa, sb, kHi, idesc, kb > kb0 ? 1u : 0u);
#pragma unroll
            for (int j = 1; j < k2BK / 16; ++j)
              umma_f16_ss_2cta(tmem_d, sa + j * a_step, sb + j * b_step, kHi, idesc, 1u);
            umma_commit_2cta_mc(&empty_bar[stage]);                    // frees this stage in BOTH CTAs
            if (kb == kb1 - 1) umma_commit_2cta_mc(&tmem_full[acc]);   // accumulators ready in BOTH CTAs
          }
          __syncwarp();
          if (++stage == k2Stages) { stage = 0; phase ^= 1; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= 3) {
    // ===================== epilogue (both CTAs, own 128 rows; 8 warps) =====================
    // The fused epilogues (bias, gelu, gelu', residual, second output) cost 20-40 instructions per element, which
    // four warps cannot retire within one tile's MMA time (8192 clk): warps 3-6 take columns [0,128) of the tile,
    // warps 7-10 columns [128,256), one TMEM lane quadrant (warp id % 4) each. Per tile and warp:
    //   - the bias of the warp's 128 columns is staged once in shared memory as fp32 (broadcast LDS afterwards)
    //   - the row-indexed operands (gelu' pre-activation, residual) of all four 32-column chunks are requested BEFORE
    //     waiting for the accumulator, so their latency hides behind the mainloop instead of stalling every chunk
    //   - bf16 outputs leave through swizzled smem + TMA stores (coalesced 128-byte rows, tails clipped by the tensor
    //     map); fp32 / atomic outputs use direct register stores.
    const int quad = warp & 3, half = (warp - 3) >> 2, ew = warp - 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    const bool tma_c = (p.out_kind == 0);
    const bool tma_pre = tma_c && (p.epi & DEDLOC_EPI_SAVE_PRE);
    const bool has_bias = (p.epi & DEDLOC_EPI_BIAS) != 0;
    const bool has_aux = (p.epi & (DEDLOC_EPI_DGELU | DEDLOC_EPI_MUL_AUX)) != 0, has_res = (p.epi & DEDLOC_EPI_RESIDUAL) != 0;
    const bool mul_aux = (p.epi & DEDLOC_EPI_MUL_AUX) != 0;     // aux is the saved derivative itself
    const uint32_t stg = smem_u32(staging) + ew * 8192;
    const uint32_t sbias_a = smem_u32(sbias) + half * 512;
    int sbuf = 0;
    for (int tile = cluster_id; tile < total_tiles; tile += num_clusters) {
      const int rem = tile % (m_tiles * n_tiles);
      int mt, nt;
      tile_coords(rem, m_tiles, n_tiles, mt, nt);
      const int m0 = mt * (2 * k2BM) + (int)cta_rank * k2BM;
      const int n0 = nt * k2BN + half * (k2BN / 2);          // first column of this warp
      const int row = m0 + quad * 32 + lane;
      const bool row_ok = row < p.M;
      if (has_bias) {
        // everybody in this half is done with the previous tile's bias; quadrant 0 stages the next 128 values
        asm volatile("bar.sync %0, 128;" ::"r"(1 + half) : "memory");
        if (quad == 0) {
          float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
          const int c = n0 + lane * 4;
          if (p.epi & DEDLOC_EPI_BIAS_F32) {
            const float* bp = reinterpret_cast<const float*>(p.bias);
            if (c + 3 < p.N) b4 = __ldg(reinterpret_cast<const float4*>(bp + c));
            else { if (c < p.N) b4.x = bp[c]; if (c + 1 < p.N) b4.y = bp[c + 1]; if (c + 2 < p.N) b4.z = bp[c + 2]; }
          } else {
            const __nv_bfloat16* bp = reinterpret_cast<const __nv_bfloat16*>(p.bias);
            if (c + 3 < p.N) {
              const uint2 u = __ldg(reinterpret_cast<const uint2*>(bp + c));
              const float2 lo = unpack_bf16x2(u.x), hi = unpack_bf16x2(u.y);
              b4 = make_float4(lo.x, lo.y, hi.x, hi.y);
            } else {
              if (c < p.N) b4.x = __bfloat162float(bp[c]);
              if (c + 1 < p.N) b4.y = __bfloat162float(bp[c + 1]);
              if (c + 2 < p.N) b4.z = __bfloat162float(bp[c + 2]);
            }
          }
          asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(sbias_a + lane * 16), "f"(b4.x), "f"(b4.y),
                       "f"(b4.z), "f"(b4.w) : "memory");
        }
        asm volatile("bar.sync %0, 128;" ::"r"(1 + half) : "memory");
      }
      // Row-indexed operand (gelu' input or residual; if both are requested the residual is read directly below):
      // requested BEFORE waiting for the accumulator with COALESCED loads -- 8 lanes x 16 B per 128-byte row segment,
      // 4 rows per instruction (row-per-thread loads touch 32 lines per instruction and throttle the LSU to ~4 clk
      // per request: 15 000 clk per tile) -- and transposed to row-per-thread through the warp's staging buffer.
      uint4 pf[8];                                   // one 64-column box at a time (box 1 is requested after box 0 is staged)
      const bool use_pf = tma_c && (has_aux || has_res) && (n0 + 128 <= p.N) && (m0 + quad * 32 + 32 <= p.M);
      const __nv_bfloat16* pf_src = nullptr;
      long long pf_ld = 0;
      if (use_pf) {
        pf_src = has_aux ? reinterpret_cast<const __nv_bfloat16*>(p.aux_in) : reinterpret_cast<const __nv_bfloat16*>(p.residual);
        pf_ld = has_aux ? p.ld_aux : p.ld_res;
        pf_src += (size_t)(m0 + quad * 32 + (lane >> 3)) * pf_ld + n0 + (lane & 7) * 8;
#pragma unroll
        for (int q = 0; q < 8; ++q) pf[q] = ld_stream_v4(pf_src + (size_t)(q * 4) * pf_ld);
      }
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + acc * k2BN + half * (k2BN / 2);
      // v = epilogue(accumulator chunk c); `pre` only with SAVE_PRE; `rowop_a` = smem address of this row's 32 staged
      // bf16 operand values (swizzled 16-byte pieces), read 8 at a time to keep the live set small
      auto chunk_math = [&](int c, const uint32_t (&r)[32], float (&v)[32], float (&pre)[32], uint32_t rowop_a,
                            bool have_rowop) {
        const int col0 = n0 + c * 32;
        if (p.alpha != 1.f) {
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]) * p.alpha;
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
        }
        if (has_bias) {
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float4 b4 = lds_f4(sbias_a + (c * 32 + i) * 4);
            v[i] += b4.x; v[i + 1] += b4.y; v[i + 2] += b4.z; v[i + 3] += b4.w;
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
        if (has_aux && row_ok) {
          if (have_rowop) {
#pragma unroll
            for (int pc = 0; pc < 4; ++pc) {
              const float4 f = lds_f4(rowop_a + ((((uint32_t)((c & 1) * 4 + pc)) ^ ((uint32_t)lane & 7)) << 4));
              const float2 a0 = unpack_bf16x2(__float_as_uint(f.x)), a1 = unpack_bf16x2(__float_as_uint(f.y));
              const float2 a2 = unpack_bf16x2(__float_as_uint(f.z)), a3 = unpack_bf16x2(__float_as_uint(f.w));
              if (mul_aux) {
                v[pc * 8 + 0] *= a0.x; v[pc * 8 + 1] *= a0.y; v[pc * 8 + 2] *= a1.x; v[pc * 8 + 3] *= a1.y;
                v[pc * 8 + 4] *= a2.x; v[pc * 8 + 5] *= a2.y; v[pc * 8 + 6] *= a3.x; v[pc * 8 + 7] *= a3.y;
              } else {
                v[pc * 8 + 0] *= gelu_tanh_grad(a0.x); v[pc * 8 + 1] *= gelu_tanh_grad(a0.y);
                v[pc * 8 + 2] *= gelu_tanh_grad(a1.x); v[pc * 8 + 3] *= gelu_tanh_grad(a1.y);
                v[pc * 8 + 4] *= gelu_tanh_grad(a2.x); v[pc * 8 + 5] *= gelu_tanh_grad(a2.y);
                v[pc * 8 + 6] *= gelu_tanh_grad(a3.x); v[pc * 8 + 7] *= gelu_tanh_grad(a3.y);
              }
            }
          } else {
            const __nv_bfloat16* ai = reinterpret_cast<const __nv_bfloat16*>(p.aux_in) + (size_t)row * p.ld_aux + col0;
            _Pragma("unroll") for (int i = 0; i < 32; ++i) if (col0 + i < p.N) {
              const float a = __bfloat162float(ai[i]);
              v[i] *= mul_aux ? a : gelu_tanh_grad(a);
            }
          }
        }
        if (has_res && row_ok) {
          if (have_rowop && !has_aux) {
#pragma unroll
            for (int pc = 0; pc < 4; ++pc) {
              const float4 f = lds_f4(rowop_a + ((((uint32_t)((c & 1) * 4 + pc)) ^ ((uint32_t)lane & 7)) << 4));
              const float2 a0 = unpack_bf16x2(__float_as_uint(f.x)), a1 = unpack_bf16x2(__float_as_uint(f.y));
              const float2 a2 = unpack_bf16x2(__float_as_uint(f.z)), a3 = unpack_bf16x2(__float_as_uint(f.w));
              v[pc * 8 + 0] += a0.x; v[pc * 8 + 1] += a0.y; v[pc * 8 + 2] += a1.x; v[pc * 8 + 3] += a1.y;
              v[pc * 8 + 4] += a2.x; v[pc * 8 + 5] += a2.y; v[pc * 8 + 6] += a3.x; v[pc * 8 + 7] += a3.y;
            }
          } else {
            const __nv_bfloat16* rs = reinterpret_cast<const __nv_bfloat16*>(p.residual) + (size_t)row * p.ld_res + col0;
            _Pragma("unroll") for (int i = 0; i < 32; ++i) if (col0 + i < p.N) v[i] += __bfloat162float(rs[i]);
          }
        }
      };
      if (!tma_c) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int col0 = n0 + c * 32;
          if (col0 >= p.N) break;  // warp-uniform
          uint32_t r[32];
          tmem_ld_32x32b_x32(taddr + c * 32, r);
          tmem_ld_wait();
          float v[32], pre[32];
          chunk_math(c, r, v, pre, 0u, false);
          if (!row_ok) continue;
          if (p.epi & DEDLOC_EPI_SAVE_PRE)
            store_bf16_32(reinterpret_cast<__nv_bfloat16*>(p.pre_out) + (size_t)row * p.ld_pre + col0, pre, col0 + 32 <= p.N,
                          p.N - col0);
          gemm_epilogue_store(p, row, col0, v);
        }
      } else {
        // two staging buffers per warp. Without a second output they alternate per 64-column box; with SAVE_PRE buffer
        // 0 carries the pre-activation and buffer 1 the output, each waiting only for ITS previous store
        // (wait_group.read 1 leaves the most recent store in flight).
#pragma unroll
        for (int cb = 0; cb < 2; ++cb) {
          const int col0 = n0 + cb * 64;
          if (col0 >= p.N) break;  // warp-uniform
          const int b_first = tma_pre ? 0 : sbuf;
          const uint32_t buf_first = stg + b_first * 4096;
          if (lane == 0) tma_store_wait_read<1>();      // buf_first has been read by its previous store
          __syncwarp();
          if (use_pf) {
            // coalesced registers -> swizzled smem; afterwards thread L only touches row L of the buffer (it reads the
            // operand of a chunk and later writes the result of the same chunk in place), so one __syncwarp suffices
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const uint32_t rr = i * 4 + (lane >> 3);
              sts_u4(buf_first + rr * 128 + ((((uint32_t)lane & 7) ^ (rr & 7)) << 4), pf[i].x, pf[i].y, pf[i].z, pf[i].w);
            }
            __syncwarp();
            if (cb == 0) {
#pragma unroll
              for (int q = 0; q < 8; ++q) pf[q] = ld_stream_v4(pf_src + (size_t)(q * 4) * pf_ld + 64);
            }
          }
          auto stage_chunk = [&](const float (&x)[32], uint32_t buf, int hh) {
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
              const uint32_t off = (uint32_t)lane * 128 + (((uint32_t)(hh * 4 + ch) ^ ((uint32_t)lane & 7)) << 4);
              sts_u4(buf + off, pack_bf16x2(x[8 * ch], x[8 * ch + 1]), pack_bf16x2(x[8 * ch + 2], x[8 * ch + 3]),
                     pack_bf16x2(x[8 * ch + 4], x[8 * ch + 5]), pack_bf16x2(x[8 * ch + 6], x[8 * ch + 7]));
            }
          };
          auto store_box = [&](int bsel, const CUtensorMap* tm) {
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(tm, staging + ew * 8192 + bsel * 4096, col0, m0 + quad * 32);
              tma_store_commit();
            }
          };
          if (tma_pre) {
            // second output: the saved tensor (pre-activation or gelu') is staged chunk by chunk into buffer 0; the
            // activated output waits -- packed to bf16, 32 registers -- until buffer 1's previous store has been read
            uint32_t keep[2][16];
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
              uint32_t r[32];
              tmem_ld_32x32b_x32(taddr + cb * 64 + hh * 32, r);
              tmem_ld_wait();
              // 8 columns at a time (bias, activation, both packs) to keep the live set small
              const int c = cb * 2 + hh;
              const bool both = (p.epi & DEDLOC_EPI_GELU) && (p.epi & DEDLOC_EPI_SAVE_DGELU);
#pragma unroll
              for (int g8 = 0; g8 < 4; ++g8) {
                float x[8], sv[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) x[k] = __uint_as_float(r[g8 * 8 + k]);
                if (p.alpha != 1.f) {
#pragma unroll
                  for (int k = 0; k < 8; ++k) x[k] *= p.alpha;
                }
                if (has_bias) {
                  const float4 b0 = lds_f4(sbias_a + (c * 32 + g8 * 8) * 4), b1 = lds_f4(sbias_a + (c * 32 + g8 * 8 + 4) * 4);
                  x[0] += b0.x; x[1] += b0.y; x[2] += b0.z; x[3] += b0.w;
                  x[4] += b1.x; x[5] += b1.y; x[6] += b1.z; x[7] += b1.w;
                }
                if (both) {
#pragma unroll
                  for (int k = 0; k < 8; ++k) gelu_tanh_both(x[k], x[k], sv[k]);
                } else {
#pragma unroll
                  for (int k = 0; k < 8; ++k) { sv[k] = x[k]; if (p.epi & DEDLOC_EPI_GELU) x[k] = gelu_tanh(x[k]); }
                }
                const uint32_t off = (uint32_t)lane * 128 + (((uint32_t)(hh * 4 + g8) ^ ((uint32_t)lane & 7)) << 4);
                sts_u4(stg + off, pack_bf16x2(sv[0], sv[1]), pack_bf16x2(sv[2], sv[3]), pack_bf16x2(sv[4], sv[5]),
                       pack_bf16x2(sv[6], sv[7]));
                keep[hh][g8 * 4 + 0] = pack_bf16x2(x[0], x[1]); keep[hh][g8 * 4 + 1] = pack_bf16x2(x[2], x[3]);
                keep[hh][g8 * 4 + 2] = pack_bf16x2(x[4], x[5]); keep[hh][g8 * 4 + 3] = pack_bf16x2(x[6], x[7]);
              }
            }
            store_box(0, &tmap_pre);
            if (lane == 0) tma_store_wait_read<1>();    // buffer 1's previous store (one box ago) is done
            __syncwarp();
#pragma unroll
            for (int hh = 0; hh < 2; ++hh)
#pragma unroll
              for (int ch = 0; ch < 4; ++ch) {
                const uint32_t off = (uint32_t)lane * 128 + (((uint32_t)(hh * 4 + ch) ^ ((uint32_t)lane & 7)) << 4);
                sts_u4(stg + 4096 + off, keep[hh][4 * ch], keep[hh][4 * ch + 1], keep[hh][4 * ch + 2], keep[hh][4 * ch + 3]);
              }
            store_box(1, &tmap_c);
          } else {
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
              uint32_t r[32];
              tmem_ld_32x32b_x32(taddr + cb * 64 + hh * 32, r);
              tmem_ld_wait();
              float v[32], pre[32];
              chunk_math(cb * 2 + hh, r, v, pre, buf_first + (uint32_t)lane * 128, use_pf);
              stage_chunk(v, buf_first, hh);
            }
            store_box(sbuf, &tmap_c);
            sbuf ^= 1;
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_leader(&tmem_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (tma_c && lane == 0) tma_store_wait<0>();   // staging smem must outlive the last bulk store
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // both CTAs are done with tensor memory and with each other's barriers
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_2cta(tmem_base, 2 * k2BN);
  }
}

int launch_gemm_2cta(const DedlocGemmArgs& g, cudaStream_t st) {
  GemmKernelParams kp;
  kp.M = g.M; kp.N = g.N; kp.K = g.K;
  kp.a_mn = g.a_mn_major; kp.b_mn = g.b_mn_major;
  kp.epi = g.epilogue; kp.out_kind = g.out_kind; kp.fmt = g.is_fp16 ? 0 : 1;
  kp.alpha = g.alpha;
  kp.C = g.C; kp.ldc = g.ldc;
  kp.bias = g.bias; kp.residual = g.residual; kp.ld_res = g.ld_residual;
  kp.aux_in = g.aux_in; kp.ld_aux = g.ld_aux; kp.pre_out = g.pre_out; kp.ld_pre = g.ld_pre;
  const int kb_total = (g.K + k2BK - 1) / k2BK;
  int splits = g.split_k < 1 ? 1 : g.split_k;
  if (splits > kb_total) splits = kb_total;
  const int kb_per = (kb_total + splits - 1) / splits;
  splits = (kb_total + kb_per - 1) / kb_per;
  kp.splits = splits;
  kp.kb_per_split = kb_per;
  DEDLOC_REQUIRE(splits == 1 || g.out_kind == 2, "gemm: split-K needs the fp32 atomic-add output");
  { int rc0 = bind_context_for_driver_calls(); if (rc0) return rc0; }

  CUtensorMap ta, tb;
  int rc;
  // MN-major operands whose MN extent is a multiple of 64 use a 3-D map: ONE TMA request per stage and operand, as
  // for K-major (the producer issuing twice as many requests per stage was measurable: 8 % at N x K = 1024 x 4096)
  kp.a_mn3d = g.a_mn_major && (g.M % 64 == 0);
  kp.b_mn3d = g.b_mn_major && (g.N % 64 == 0);
  if (!g.a_mn_major) rc = make_tmap_2d_16bit(&ta, g.A, g.M, g.K, g.lda, k2BK, k2BM, !g.is_fp16);
  else if (kp.a_mn3d) rc = make_tmap_mn_3d_16bit(&ta, g.A, g.K, g.M, g.lda, k2BK, k2BM / 64, !g.is_fp16);
  else rc = make_tmap_2d_16bit(&ta, g.A, g.K, g.M, g.lda, 64, k2BK, !g.is_fp16);
  if (rc) return rc;
  if (!g.b_mn_major) rc = make_tmap_2d_16bit(&tb, g.B, g.N, g.K, g.ldb, k2BK, k2BN / 2, !g.is_fp16);
  else if (kp.b_mn3d) rc = make_tmap_mn_3d_16bit(&tb, g.B, g.K, g.N, g.ldb, k2BK, (k2BN / 2) / 64, !g.is_fp16);
  else rc = make_tmap_2d_16bit(&tb, g.B, g.K, g.N, g.ldb, 64, k2BK, !g.is_fp16);
  if (rc) return rc;

  CUtensorMap tc = ta, tp = ta;   // placeholders unless the bf16 TMA-store epilogue is used
  if (g.out_kind == 0) {
    rc = make_tmap_2d_16bit(&tc, g.C, g.M, g.N, g.ldc, 64, 32, 1);
    if (rc) return rc;
    if (g.epilogue & DEDLOC_EPI_SAVE_PRE) {
      rc = make_tmap_2d_16bit(&tp, g.pre_out, g.M, g.N, g.ld_pre, 64, 32, 1);
      if (rc) return rc;
    }
  }
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    DEDLOC_CUDA_CHECK(cudaGetDevice(&dev));
    DEDLOC_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    DEDLOC_CUDA_CHECK(cudaFuncSetAttribute(gemm_tcgen05_2cta_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           k2SmemTotal));
  }
  const int m_tiles = (g.M + 2 * k2BM - 1) / (2 * k2BM), n_tiles = (g.N + k2BN - 1) / k2BN;
  const int total = m_tiles * n_tiles * splits;
  int clusters = total < sms / 2 ? total : sms / 2;
  if (g.max_ctas > 0 && g.max_ctas / 2 < clusters) clusters = g.max_ctas / 2 > 0 ? g.max_ctas / 2 : 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * clusters);
  cfg.blockDim = dim3(k2Threads);
  cfg.dynamicSmemBytes = k2SmemTotal;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  DEDLOC_CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemm_tcgen05_2cta_kernel, ta, tb, tc, tp, kp));
  count_launch();
  return 0;
}

}  // namespace dedloc
