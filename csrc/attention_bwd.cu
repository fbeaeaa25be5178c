// Shared-layer attention, backward: dQ, dK, dV written straight into the PACKED [B, S, 3, H, 64]
// gradient of the fused QKV projection -- the stack / unbind / contiguous copies of the autograd path disappear.
//
// One CTA owns one (batch, head) at a time; a step is (key block j of 128 keys) x (query block i of 128 queries),
// handled as two HALVES of 64 queries (a, b) so that the tensor core works on one half while the other half is in
// the exp / dS arithmetic:
//   warp 0      TMA: K_j, V_j (2 stages), Q_i, dO_i (2 stages)
//   warp 1      tcgen05.mma issuer
//                 1h  S^T_h  = K_j Q_h^T,  dP^T_h = V_j dO_h^T          (keys on TMEM lanes, 64 queries on columns)
//                 2h  dV_j  += P^T_h dO_h, dK_j  += dS^T_h Q_h          (A = bf16 P^T / dS^T read FROM TENSOR MEMORY)
//                 3   dQ_i(partial) = dS K_j    (A = the dS^T tile in shared memory read as an MN-major operand)
//               issue order per step t:  wait P_a -> 2a(t), 1a(t+1);  wait P_b -> 2b(t), 3(t), 1b(t+1)
//   warp 2      TMEM allocator
//   warp 3      per (batch, head): lse*log2(e), D = rowsum(dO * O) * scale and the key mask, staged in shared memory
//               one (batch, head) ahead of the compute warps
//   warps 4-7   half a, warps 8-11 half b (thread = key row): P^T = exp2(S^T c - lse_q), dS^T = P^T (dP^T c' - D_q),
//               both packed to bf16 IN PLACE over the fp32 scores (tcgen05.st); dS^T also goes to 128B-swizzled
//               shared memory for MMA 3. The same warps drain dQ (thread = query row, 32 of the 64 columns per
//               half-group) through an fp32 L2-resident scratch that accumulates over the key blocks, and dK_j / dV_j
//               after the last query block.
#include "gemm_common.cuh"

namespace dedloc {

constexpr int kAbThreads = 384;
constexpr int kAbD = 64;
constexpr int kAbBlk = 128;
constexpr int kAbMaxS = 512;
constexpr int kAbTile = kAbBlk * 128;               // 16 KB: 128 rows x 128 B
// smem: K, V x 2 stages; Q, dO x 2 stages; dS^T x 2 buffers x 2 halves (12 tiles) + staged vectors x 2 + barriers
constexpr int kAbVecBytes = kAbMaxS * (8 + 4);      // float2 {lse2, D*scale} + float key_ok per query / key
constexpr int kAbSmem = 12 * kAbTile + 2 * kAbVecBytes + 512 + 1024;

struct AttBwdParams {
  int B, S, H;
  float scale, scale_log2;
  const unsigned char* mask;       // [B, S] or null
  const __nv_bfloat16* out;        // [B, S, H*64]
  const __nv_bfloat16* dout;       // [B, S, H*64]
  const float* lse;                // [B, H, S]
  __nv_bfloat16* dqkv;             // [B, S, 3, H, 64]
  float* dq_scratch;               // [gridDim.x, S, 64] fp32, see drain_dq for the layout
};

__device__ __forceinline__ float ex2f_b(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int NBLK>
__global__ void __launch_bounds__(kAbThreads, 1)
attention_bwd_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_do,
                     const __grid_constant__ CUtensorMap tmap_dqkv, const __grid_constant__ AttBwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;                               // 2 stages
  uint8_t* sV = smem + 2 * kAbTile;                 // 2 stages
  uint8_t* sQ = smem + 4 * kAbTile;                 // 2 stages
  uint8_t* sdO = smem + 6 * kAbTile;                // 2 stages
  uint8_t* sdS = smem + 8 * kAbTile;                // [2 buffers][2 halves of 64 queries][128 keys][128 B]
  uint8_t* sVecBase = smem + 12 * kAbTile;          // [2 stages]{ float2 vec[512]; float key_ok[512]; }
  uint64_t* bars = reinterpret_cast<uint64_t*>(sVecBase + 2 * kAbVecBytes);
  uint64_t* kv_full = bars + 0;     // [2]
  uint64_t* kv_empty = bars + 2;    // [2]
  uint64_t* qdo_full = bars + 4;    // [2]
  uint64_t* qdo_empty = bars + 6;   // [2]
  uint64_t* s_full = bars + 8;      // [2] per half
  uint64_t* p_full = bars + 10;     // [2] per half
  uint64_t* dq_full = bars + 12;    // [2] dQ partial buffers (step parity)
  uint64_t* dq_empty = bars + 14;   // [2]
  uint64_t* dkv_full = bars + 16;
  uint64_t* dkv_empty = bars + 17;
  uint64_t* vec_full = bars + 18;   // [2]
  uint64_t* vec_empty = bars + 20;  // [2]
  uint64_t* stage_free = bars + 22; // [2] per half-group: the dK / dV TMA store has read its staging tile
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 24);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int nblk = NBLK;                      // key / query blocks of 128 (compile time: no integer divisions)
  const int S = p.S, H = p.H;
  const int total_bh = p.B * H;
  const int my_bh = (total_bh - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;   // (batch, head)s of this CTA
  constexpr int steps_per_bh = nblk * nblk;
  const int T = my_bh * steps_per_bh;                                                     // steps of this CTA

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmap_qkv); tma_prefetch_desc(&tmap_do); tma_prefetch_desc(&tmap_dqkv); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&kv_full[i], 1); mbar_init(&kv_empty[i], 1);
      mbar_init(&qdo_full[i], 1); mbar_init(&qdo_empty[i], 1);
      mbar_init(&s_full[i], 1); mbar_init(&p_full[i], 4);
      mbar_init(&vec_full[i], 1); mbar_init(&vec_empty[i], 8);
      mbar_init(&dq_full[i], 1); mbar_init(&dq_empty[i], 8);
      mbar_init(&stage_free[i], 1);
    }
    mbar_init(dkv_full, 1); mbar_init(dkv_empty, 8);
    mbar_fence_init();
  }
  if (warp == 2) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // columns: S^T_a | dP^T_a | S^T_b | dP^T_b (64 each; the bf16 P^T / dS^T reuse the first 32 of each) | dV | dK | dQ x 2
  const uint32_t t_dV = tmem_base + 256, t_dK = tmem_base + 320, t_dQ = tmem_base + 384;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int t = 0, jc = 0;
      for (int bh = blockIdx.x; bh < total_bh; bh += gridDim.x) {
        const int b = bh / H, h = bh % H;
        for (int j = 0; j < nblk; ++j, ++jc) {
          const int ks = jc & 1;
          mbar_wait(&kv_empty[ks], ((jc >> 1) & 1) ^ 1);
          mbar_expect_tx(&kv_full[ks], 2 * kAbTile);
          tma_load_2d(sK + ks * kAbTile, &tmap_qkv, &kv_full[ks], (1 * H + h) * kAbD, b * S + j * kAbBlk);
          tma_load_2d(sV + ks * kAbTile, &tmap_qkv, &kv_full[ks], (2 * H + h) * kAbD, b * S + j * kAbBlk);
          for (int i = 0; i < nblk; ++i, ++t) {
            const int qs = t & 1;
            mbar_wait(&qdo_empty[qs], ((t >> 1) & 1) ^ 1);
            mbar_expect_tx(&qdo_full[qs], 2 * kAbTile);
            tma_load_2d(sQ + qs * kAbTile, &tmap_qkv, &qdo_full[qs], (0 * H + h) * kAbD, b * S + i * kAbBlk);
            tma_load_2d(sdO + qs * kAbTile, &tmap_do, &qdo_full[qs], h * kAbD, b * S + i * kAbBlk);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // Issue cost matters here (a 128 x 64 x 16 MMA runs for 32 clocks): descriptors are a precomputed low word plus
    // an immediate -- the high word (SBO = 1024, version 1, SWIZZLE_128B) is the same for all of them -- and the
    // single-thread regions are entered through elect.sync so that the operands stay in uniform registers.
    constexpr uint32_t kHi = 0x40004040u;
    constexpr uint32_t kLboK = 1u << 16;                         // K-major: LBO = 16 B (ignored under swizzle)
    constexpr uint32_t kLboMN = (uint32_t)(kAbTile >> 4) << 16;  // MN-major: next 64-element group = next tile
    constexpr uint32_t kStage = kAbTile >> 4;                    // one tile, in descriptor address units
    const uint32_t id_s = make_idesc_f16(128, 64, 0, 0, 1);     // S^T_h, dP^T_h: A K-major, B K-major, N = 64
    const uint32_t id_ts = make_idesc_f16(128, 64, 0, 1, 1);    // dV, dK: A from TMEM, B MN-major (dO / Q)
    const uint32_t id_mnmn = make_idesc_f16(128, 64, 1, 1, 1);  // dQ: A MN-major (dS^T read as dS), B MN-major (K)
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);  // provably warp-uniform copy
    const uint32_t u_dV = tb + 256, u_dK = tb + 320, u_dQ = tb + 384;
    const uint32_t bK = (smem_u32(sK) & 0x3FFFF) >> 4, bV = (smem_u32(sV) & 0x3FFFF) >> 4;
    const uint32_t bQ = (smem_u32(sQ) & 0x3FFFF) >> 4, bO = (smem_u32(sdO) & 0x3FFFF) >> 4;
    const uint32_t bS = (smem_u32(sdS) & 0x3FFFF) >> 4;
    // 1h: scores of half hf for the step whose operands sit in K/V stage ks and Q/dO stage qs
    auto issue_scores = [&](int ks, int qs, int hf) {
      if (elect_one()) {
        const uint32_t aK = (bK + ks * kStage) | kLboK, aV = (bV + ks * kStage) | kLboK;
        const uint32_t aQ = (bQ + qs * kStage + hf * 512) | kLboK, aO = (bO + qs * kStage + hf * 512) | kLboK;
        const uint32_t tS = tb + hf * 128, tP = tS + 64;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) umma_f16_ss_lohi(tS, aK + kk * 2, aQ + kk * 2, kHi, id_s, kk > 0);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) umma_f16_ss_lohi(tP, aV + kk * 2, aO + kk * 2, kHi, id_s, kk > 0);
        umma_commit(&s_full[hf]);
      }
      __syncwarp();
    };
    // 2h: dV += P^T_h dO_h, dK += dS^T_h Q_h with A in tensor memory (16 queries = 8 columns per k-step)
    auto issue_dkv = [&](int qs, int hf, uint32_t first) {
      if (elect_one()) {
        const uint32_t aQ = (bQ + qs * kStage + hf * 512) | kLboMN, aO = (bO + qs * kStage + hf * 512) | kLboMN;
        const uint32_t tS = tb + hf * 128, tP = tS + 64;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) umma_f16_ts_lohi(u_dV, tS + kk * 8, aO + kk * 128, kHi, id_ts, kk > 0 ? 1u : first);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) umma_f16_ts_lohi(u_dK, tP + kk * 8, aQ + kk * 128, kHi, id_ts, kk > 0 ? 1u : first);
      }
      __syncwarp();
    };
    if (T > 0) {
      mbar_wait(&kv_full[0], 0);
      mbar_wait(&qdo_full[0], 0);
      tc_fence_after();
      issue_scores(0, 0, 0);
      issue_scores(0, 0, 1);
    }
    int i = 0, jc = 0;                       // query block of step t, running key-block counter
    for (int t = 0; t < T; ++t) {
      const int ks = jc & 1, qs = t & 1;
      const bool last_i = (i == nblk - 1);
      // operands of step t+1
      const int n_jc = last_i ? jc + 1 : jc, n_ks = n_jc & 1, n_qs = qs ^ 1;
      mbar_wait(&p_full[0], t & 1);
      if (i == 0 && jc > 0) mbar_wait(dkv_empty, (jc - 1) & 1);   // dK / dV of the previous key block are drained
      tc_fence_after();
      issue_dkv(qs, 0, i == 0 ? 0u : 1u);
      if (t + 1 < T) {
        if (last_i) mbar_wait(&kv_full[n_ks], (n_jc >> 1) & 1);
        mbar_wait(&qdo_full[n_qs], ((t + 1) >> 1) & 1);
        tc_fence_after();
        issue_scores(n_ks, n_qs, 0);
      }
      mbar_wait(&p_full[1], t & 1);
      tc_fence_after();
      issue_dkv(qs, 1, 1u);
      if (elect_one()) umma_commit(&qdo_empty[qs]);               // Q_i / dO_i are not needed by MMA 3
      __syncwarp();
      if (t + 1 < T) issue_scores(n_ks, n_qs, 1);                 // before MMA 3: half b waits for these scores
      if (t > 1) mbar_wait(&dq_empty[t & 1], ((t - 2) >> 1) & 1); // dQ partial of step t-2 (same buffer) was read out
      tc_fence_after();
      if (elect_one()) {
        const uint32_t aS = (bS + (t & 1) * 2 * kStage) | kLboMN, aK = (bK + ks * kStage) | kLboMN;
        const uint32_t tQ = u_dQ + (t & 1) * 64;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)       // reduction over the 128 keys
          umma_f16_ss_lohi(tQ, aS + kk * 128, aK + kk * 128, kHi, id_mnmn, kk > 0);
        umma_commit(&dq_full[t & 1]);
        if (last_i) { umma_commit(dkv_full); umma_commit(&kv_empty[ks]); }
      }
      __syncwarp();
      if (last_i) { i = 0; ++jc; } else { ++i; }
    }
  } else if (warp == 3) {
    // ===================== per-(batch, head) vectors, one (batch, head) ahead =====================
    int bc = 0;
    for (int bh = blockIdx.x; bh < total_bh; bh += gridDim.x, ++bc) {
      const int b = bh / H, h = bh % H, st = bc & 1;
      float2* vec = reinterpret_cast<float2*>(sVecBase + st * kAbVecBytes);
      float* key_ok = reinterpret_cast<float*>(sVecBase + st * kAbVecBytes + kAbMaxS * 8);
      mbar_wait(&vec_empty[st], ((bc >> 1) & 1) ^ 1);
      // lse and the key mask: one row per lane
      for (int r = lane; r < S; r += 32) {
        const float l = p.lse[((size_t)b * H + h) * S + r];
        // fully masked row: lse = -inf -> +inf here -> P = exp2(-inf) = 0
        vec[r].x = (l == -INFINITY) ? INFINITY : l * 1.4426950408889634f;
        key_ok[r] = (p.mask == nullptr || p.mask[(size_t)b * S + r]) ? 1.f : 0.f;
      }
      // D = rowsum(dO * O) * scale: 8 lanes x 16 B cover one 128-byte row (4 rows per warp instruction -> 4 cache
      // lines instead of 32), 16 row groups in flight
      {
        const int sub = lane >> 3, ch = lane & 7;
        for (int r0 = 0; r0 < S; r0 += 64) {
          uint4 a[16], g[16];
#pragma unroll
          for (int u = 0; u < 16; ++u) {
            const size_t off = ((size_t)(b * S + r0 + u * 4 + sub) * H + h) * kAbD + ch * 8;
            a[u] = *reinterpret_cast<const uint4*>(p.out + off);
            g[u] = *reinterpret_cast<const uint4*>(p.dout + off);
          }
#pragma unroll
          for (int u = 0; u < 16; ++u) {
            float2 x, y;
            float acc = 0.f;
            x = unpack_bf16x2(a[u].x); y = unpack_bf16x2(g[u].x); acc = fmaf(x.x, y.x, acc); acc = fmaf(x.y, y.y, acc);
            x = unpack_bf16x2(a[u].y); y = unpack_bf16x2(g[u].y); acc = fmaf(x.x, y.x, acc); acc = fmaf(x.y, y.y, acc);
            x = unpack_bf16x2(a[u].z); y = unpack_bf16x2(g[u].z); acc = fmaf(x.x, y.x, acc); acc = fmaf(x.y, y.y, acc);
            x = unpack_bf16x2(a[u].w); y = unpack_bf16x2(g[u].w); acc = fmaf(x.x, y.x, acc); acc = fmaf(x.y, y.y, acc);
            acc += __shfl_xor_sync(0xffffffffu, acc, 1);
            acc += __shfl_xor_sync(0xffffffffu, acc, 2);
            acc += __shfl_xor_sync(0xffffffffu, acc, 4);
            if (ch == 0) vec[r0 + u * 4 + sub].y = acc * p.scale;
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&vec_full[st]);
    }
  } else if (warp >= 4) {
    // ===================== exp / dS arithmetic + read-out: warps 4-7 = half a, warps 8-11 = half b =====================
    const int hf = (warp - 4) >> 2;                // 0: queries [0, 64) of the block, 1: queries [64, 128)
    const int quad = warp & 3;
    const int row = quad * 32 + lane;              // TMEM lane: key row for S^T/dP^T/dK/dV, query row for dQ
    const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
    const uint32_t tS = tmem_base + hf * 128 + lane_off, tP = tS + 64;
    float* dq_acc = p.dq_scratch + (size_t)blockIdx.x * S * kAbD;

    // dQ partial of step t (query row `row`, columns [hf*32, hf*32+32)) accumulated over the key blocks in an fp32
    // scratch private to this CTA. Scratch layout [query block][16-byte column chunk (16)][row (128)][4 floats]: the
    // 32 lanes of a warp touch 512 contiguous bytes per access. The running sum needed by the NEXT drain is fetched
    // right after the current one, a whole arithmetic phase ahead of its use.
    float4 prev[8];
    auto scratch_of = [&](int t) {
      const int i = t % nblk;
      return dq_acc + ((size_t)(i * 16 + hf * 8) * kAbBlk + row) * 4;
    };
    auto prefetch_dq = [&](int t) {
      if ((t % steps_per_bh) / nblk > 0) {
        const float* acc = scratch_of(t);
#pragma unroll
        for (int q = 0; q < 8; ++q) prev[q] = *reinterpret_cast<const float4*>(acc + (size_t)q * kAbBlk * 4);
      }
    };
    auto drain_dq = [&](int t) {
      const int bh = blockIdx.x + (t / steps_per_bh) * gridDim.x;
      const int b = bh / H, h = bh % H;
      const int j = (t % steps_per_bh) / nblk, i = t % nblk;
      mbar_wait(&dq_full[t & 1], (t >> 1) & 1);
      tc_fence_after();
      uint32_t r[32];
      tmem_ld_32x32b_x32(t_dQ + (t & 1) * 64 + lane_off + hf * 32, r);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&dq_empty[t & 1]);   // the values are in registers: this dQ buffer may be rewritten
      float v[32];
#pragma unroll
      for (int q = 0; q < 32; ++q) v[q] = __uint_as_float(r[q]);
      if (j > 0) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          v[4 * q] += prev[q].x; v[4 * q + 1] += prev[q].y; v[4 * q + 2] += prev[q].z; v[4 * q + 3] += prev[q].w;
        }
      }
      if (j < nblk - 1) {
        float* acc = scratch_of(t);
#pragma unroll
        for (int q = 0; q < 8; ++q)
          *reinterpret_cast<float4*>(acc + (size_t)q * kAbBlk * 4) = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
      } else {
        __nv_bfloat16* dst = p.dqkv + (((size_t)(b * S + i * kAbBlk + row) * 3 + 0) * H + h) * kAbD + hf * 32;
#pragma unroll
        for (int q = 0; q < 32; q += 8)
          *reinterpret_cast<uint4*>(dst + q) =
              make_uint4(pack_bf16x2(v[q], v[q + 1]), pack_bf16x2(v[q + 2], v[q + 3]),
                         pack_bf16x2(v[q + 4], v[q + 5]), pack_bf16x2(v[q + 6], v[q + 7]));
      }
    };
    // dV_j (half-group a) / dK_j (half-group b) after the last query block of key block jc: tensor memory -> bf16 ->
    // this half-group's own half of the dS^T buffer of the step that just finished (free: its MMA 3 has retired and
    // nobody but this half-group writes it) -> ONE TMA store of the [128 keys x 64] tile. Row-per-thread global
    // stores touch 32 lines per instruction and kept the tensor core waiting for dkv_empty at every key block.
    auto drain_dkv = [&](int jc) {
      const int bh = blockIdx.x + (jc / nblk) * gridDim.x;
      const int b = bh / H, h = bh % H, j = jc % nblk;
      const int t_last = jc * nblk + nblk - 1;
      mbar_wait(dkv_full, jc & 1);
      tc_fence_after();
      const uint32_t tacc = (hf == 0 ? t_dV : t_dK) + lane_off;
      uint32_t r0[32], r1[32];
      tmem_ld_32x32b_x32(tacc, r0);
      tmem_ld_32x32b_x32(tacc + 32, r1);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(dkv_empty);
      uint8_t* stage = sdS + ((t_last & 1) * 2 + hf) * kAbTile;
      const uint32_t st_row = smem_u32(stage) + row * 128;
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) {
        const int q = ch * 8;
        sts_u4(st_row + (((uint32_t)ch ^ ((uint32_t)row & 7)) << 4),
               pack_bf16x2(__uint_as_float(r0[q]), __uint_as_float(r0[q + 1])),
               pack_bf16x2(__uint_as_float(r0[q + 2]), __uint_as_float(r0[q + 3])),
               pack_bf16x2(__uint_as_float(r0[q + 4]), __uint_as_float(r0[q + 5])),
               pack_bf16x2(__uint_as_float(r0[q + 6]), __uint_as_float(r0[q + 7])));
        sts_u4(st_row + (((uint32_t)(ch + 4) ^ ((uint32_t)row & 7)) << 4),
               pack_bf16x2(__uint_as_float(r1[q]), __uint_as_float(r1[q + 1])),
               pack_bf16x2(__uint_as_float(r1[q + 2]), __uint_as_float(r1[q + 3])),
               pack_bf16x2(__uint_as_float(r1[q + 4]), __uint_as_float(r1[q + 5])),
               pack_bf16x2(__uint_as_float(r1[q + 6]), __uint_as_float(r1[q + 7])));
      }
      fence_proxy_async_smem();
      asm volatile("bar.sync %0, 128;" ::"r"(1 + hf) : "memory");
      if (quad == 0 && lane == 0) {
        tma_store_2d(&tmap_dqkv, stage, ((hf == 0 ? 2 : 1) * H + h) * kAbD, b * S + j * kAbBlk);
        tma_store_commit();
        tma_store_wait_read<0>();
        mbar_arrive(&stage_free[hf]);              // compute_half two steps later overwrites this tile
      }
    };
    // P^T and dS^T of key `row` against the 64 queries of this half
    auto compute_half = [&](int t) {
      const int bc = t / steps_per_bh, st = bc & 1;
      const int j = (t % steps_per_bh) / nblk, i = t % nblk;
      const uint32_t vec_a = smem_u32(sVecBase) + st * kAbVecBytes + (i * kAbBlk + hf * 64) * 8;
      const uint32_t kok_a = smem_u32(sVecBase) + st * kAbVecBytes + kAbMaxS * 8 + (j * kAbBlk + row) * 4;
      if (t % steps_per_bh == 0) mbar_wait(&vec_full[st], (bc >> 1) & 1);
      const bool kok = lds_f1(kok_a) != 0.f;
      const uint32_t ds_row = smem_u32(sdS) + ((t & 1) * 2 + hf) * kAbTile + row * 128;
      if (t >= 2 && (t - 2) % nblk == nblk - 1) mbar_wait(&stage_free[hf], ((t - 2) / nblk) & 1);
      mbar_wait(&s_full[hf], t & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t rs[32], rp[32];
        tmem_ld_32x32b_x32(tS + c * 32, rs);
        tmem_ld_32x32b_x32(tP + c * 32, rp);
        tmem_ld_wait();
        uint32_t pk[16], dk[16];
#pragma unroll
        for (int q = 0; q < 32; q += 2) {
          const float4 v = lds_f4(vec_a + (c * 16 + (q >> 1)) * 16);   // {lse2, D*scale} of queries q, q+1 (broadcast)
          const float p0 = ex2f_b(fmaf(__uint_as_float(rs[q]), p.scale_log2, -v.x));
          const float p1 = ex2f_b(fmaf(__uint_as_float(rs[q + 1]), p.scale_log2, -v.z));
          const float d0 = p0 * fmaf(__uint_as_float(rp[q]), p.scale, -v.y);
          const float d1 = p1 * fmaf(__uint_as_float(rp[q + 1]), p.scale, -v.w);
          pk[q >> 1] = pack_bf16x2(p0, p1);
          dk[q >> 1] = pack_bf16x2(d0, d1);
        }
        if (!kok) {                                          // padded key: contributes nothing
#pragma unroll
          for (int q = 0; q < 16; ++q) { pk[q] = 0u; dk[q] = 0u; }
        }
        // in place: chunk c occupies columns [16c, 16c+16) of the bf16 view; the fp32 columns it overwrites
        // ([0,16) / [16,32)) belong to chunk 0, which has been read
        tmem_st_32x32b_x16(tS + c * 16, pk);
        tmem_st_32x32b_x16(tP + c * 16, dk);
        // 32 queries = 64 B = four 16-byte chunks of this key's 128-byte row (SWIZZLE_128B)
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          const uint32_t cidx = c * 4 + ch;
          sts_u4(ds_row + ((cidx ^ (row & 7)) << 4), dk[4 * ch], dk[4 * ch + 1], dk[4 * ch + 2], dk[4 * ch + 3]);
        }
      }
      tmem_st_wait();
      fence_proxy_async_smem();     // generic-proxy smem writes -> visible to the tensor-core (async) proxy
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[hf]);
      if (t % steps_per_bh == steps_per_bh - 1) {            // last use of this (batch, head)'s staged vectors
        __syncwarp();
        if (lane == 0) mbar_arrive(&vec_empty[st]);
      }
    };

    // both half-groups: arithmetic of step t first, then the read-out of step t-1 (its MMA 3 finished long ago: the dQ
    // partials are double-buffered in tensor memory, so neither side waits for the other here)
    for (int t = 0; t < T; ++t) {
      compute_half(t);
      if (t > 0) {
        if (t % nblk == 0) drain_dkv(t / nblk - 1);   // first: the tensor core waits for these accumulators
        drain_dq(t - 1);
      }
      prefetch_dq(t);
    }
    if (T > 0) {
      drain_dq(T - 1);
      drain_dkv(T / nblk - 1);
    }
    if (quad == 0 && lane == 0) tma_store_wait<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

}  // namespace dedloc

using namespace dedloc;

extern "C" long long dedloc_attention_bwd_scratch_floats(int S) {
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return (long long)sms * S * kAbD;
}

extern "C" int dedloc_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse,
                                    const unsigned char* mask, void* dqkv, float* scratch, int B, int S, int H, int D,
                                    float scale, void* stream) {
  DEDLOC_REQUIRE(D == kAbD, "attention: head_dim must be 64 (got %d)", D);
  DEDLOC_REQUIRE(S % 128 == 0 && S >= 128 && S <= kAbMaxS, "attention: S (%d) must be a multiple of 128 up to %d", S, kAbMaxS);
  DEDLOC_REQUIRE(scratch != nullptr, "attention_bwd: scratch of dedloc_attention_bwd_scratch_floats(S) floats is required");
  { int rc0 = bind_context_for_driver_calls(); if (rc0) return rc0; }
  CUtensorMap tq, td, tg;
  int rc = make_tmap_2d_16bit(&tq, qkv, (long long)B * S, 3LL * H * D, 3LL * H * D, 64, 128, 1);
  if (rc) return rc;
  rc = make_tmap_2d_16bit(&td, dout, (long long)B * S, (long long)H * D, (long long)H * D, 64, 128, 1);
  if (rc) return rc;
  rc = make_tmap_2d_16bit(&tg, dqkv, (long long)B * S, 3LL * H * D, 3LL * H * D, 64, 128, 1);
  if (rc) return rc;
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    DEDLOC_CUDA_CHECK(cudaGetDevice(&dev));
    DEDLOC_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    DEDLOC_CUDA_CHECK(cudaFuncSetAttribute(attention_bwd_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAbSmem));
    DEDLOC_CUDA_CHECK(cudaFuncSetAttribute(attention_bwd_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAbSmem));
    DEDLOC_CUDA_CHECK(cudaFuncSetAttribute(attention_bwd_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAbSmem));
    DEDLOC_CUDA_CHECK(cudaFuncSetAttribute(attention_bwd_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAbSmem));
  }
  AttBwdParams p;
  p.B = B; p.S = S; p.H = H;
  p.scale = scale;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.mask = mask;
  p.out = (const __nv_bfloat16*)out;
  p.dout = (const __nv_bfloat16*)dout;
  p.lse = lse;
  p.dqkv = (__nv_bfloat16*)dqkv;
  p.dq_scratch = scratch;
  const int grid = B * H < sms ? B * H : sms;
  switch (S / kAbBlk) {
    case 1: attention_bwd_kernel<1><<<grid, kAbThreads, kAbSmem, (cudaStream_t)stream>>>(tq, td, tg, p); break;
    case 2: attention_bwd_kernel<2><<<grid, kAbThreads, kAbSmem, (cudaStream_t)stream>>>(tq, td, tg, p); break;
    case 3: attention_bwd_kernel<3><<<grid, kAbThreads, kAbSmem, (cudaStream_t)stream>>>(tq, td, tg, p); break;
    default: attention_bwd_kernel<4><<<grid, kAbThreads, kAbSmem, (cudaStream_t)stream>>>(tq, td, tg, p); break;
  }
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}
