// Shared-layer attention, backward: dQ, dK, dV written straight into the PACKED [B, S, 3, H, 64]
// gradient of the fused QKV projection -- the stack / unbind / contiguous copies of the autograd path disappear.
//
// One CTA owns one (batch, head). For every key block j (128 keys) and query block i (128 queries):
//   MMA1  S^T  = K_j Q_i^T              (keys on TMEM lanes, queries on columns)
//   MMA2  dP^T = V_j dO_i^T
//   warps 4-7 (thread = key row): P^T = exp2(S^T*c - lse_q), dS^T = P^T (dP^T - D_q) * scale, both written as
//         bf16 into 128B-swizzled shared memory [key][query]
//   MMA3  dV_j += P^T dO_i      MMA4  dK_j += dS^T Q_i      (A = the swizzled tiles, K-major; accumulators stay
//                                                            in tensor memory across the i loop)
//   MMA5  dQ_i(partial) = dS K_j   (A = the SAME dS^T tile read as an MN-major operand -> no transposed copy)
//   warps 4-7 (thread = query row): dQ partial += into an fp32 L2-resident scratch; on the last key block the
//         row is converted to bf16 and stored into dqkv
// D_q = rowsum(dO * O) and lse_q are staged in shared memory once per (batch, head). MMA1/2 of the next step are
// issued right behind MMA3-5, so they overlap with the dQ read-out.
#include "gemm_common.cuh"

namespace dedloc {

constexpr int kAbThreads = 256;
constexpr int kAbD = 64;
constexpr int kAbBlk = 128;
constexpr int kAbMaxS = 512;
constexpr int kAbTile = kAbBlk * 128;               // 16 KB: 128 rows x 128 B
// smem: K_j, V_j, Q_i x2, dO_i x2 (6 tiles) + P^T, dS^T (2 x 2 tiles) + vectors + barriers
constexpr int kAbSmem = 6 * kAbTile + 4 * kAbTile + 3 * kAbMaxS * 4 + 256 + 1024;

struct AttBwdParams {
  int B, S, H;
  float scale, scale_log2;
  const unsigned char* mask;       // [B, S] or null
  const __nv_bfloat16* out;        // [B, S, H*64]
  const __nv_bfloat16* dout;       // [B, S, H*64]
  const float* lse;                // [B, H, S]
  __nv_bfloat16* dqkv;             // [B, S, 3, H, 64]
  float* dq_scratch;               // [gridDim.x, S, 64] fp32
};

__device__ __forceinline__ float ex2f_b(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(kAbThreads, 1)
attention_bwd_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_do,
                     const __grid_constant__ AttBwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = smem + kAbTile;
  uint8_t* sQ = smem + 2 * kAbTile;                 // 2 stages
  uint8_t* sdO = smem + 4 * kAbTile;                // 2 stages
  uint8_t* sP = smem + 6 * kAbTile;                 // [2 blocks of 64 queries][128 keys][128 B]
  uint8_t* sdS = smem + 8 * kAbTile;
  float* sLse = reinterpret_cast<float*>(smem + 10 * kAbTile);   // lse * log2(e), per query
  float* sD = sLse + kAbMaxS;                                     // rowsum(dO * O), per query
  float* sKeyOk = sD + kAbMaxS;                                   // 1 / 0 per key
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKeyOk + kAbMaxS);
  uint64_t* kv_full = bars + 0;
  uint64_t* kv_empty = bars + 1;
  uint64_t* qdo_full = bars + 2;    // [2]
  uint64_t* qdo_empty = bars + 4;   // [2]
  uint64_t* s_full = bars + 6;
  uint64_t* pds_full = bars + 7;
  uint64_t* dq_full = bars + 8;
  uint64_t* dq_empty = bars + 9;
  uint64_t* dkv_full = bars + 10;
  uint64_t* dkv_empty = bars + 11;
  uint64_t* vec_ready = bars + 12;  // lse / D / key mask of this (b, h) are staged
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = p.S, H = p.H, nblk = S / kAbBlk;
  const int total_bh = p.B * H;

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmap_qkv); tma_prefetch_desc(&tmap_do); }
  if (warp == 1 && lane == 0) {
    mbar_init(kv_full, 1); mbar_init(kv_empty, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&qdo_full[i], 1); mbar_init(&qdo_empty[i], 1); }
    mbar_init(s_full, 1); mbar_init(pds_full, 4);
    mbar_init(dq_full, 1); mbar_init(dq_empty, 4);
    mbar_init(dkv_full, 1); mbar_init(dkv_empty, 4);
    mbar_init(vec_ready, 4);
    mbar_fence_init();
  }
  if (warp == 2) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t t_S = tmem_base, t_dP = tmem_base + 128, t_dV = tmem_base + 256, t_dK = tmem_base + 320,
                 t_dQ = tmem_base + 384;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t kv_phase = 0, qphase = 0;
      int qs = 0;
      for (int bh = blockIdx.x; bh < total_bh; bh += gridDim.x) {
        const int b = bh / H, h = bh % H;
        for (int j = 0; j < nblk; ++j) {
          mbar_wait(kv_empty, kv_phase ^ 1);
          mbar_expect_tx(kv_full, 2 * kAbTile);
          tma_load_2d(sK, &tmap_qkv, kv_full, (1 * H + h) * kAbD, b * S + j * kAbBlk);
          tma_load_2d(sV, &tmap_qkv, kv_full, (2 * H + h) * kAbD, b * S + j * kAbBlk);
          kv_phase ^= 1;
          for (int i = 0; i < nblk; ++i) {
            mbar_wait(&qdo_empty[qs], qphase ^ 1);
            mbar_expect_tx(&qdo_full[qs], 2 * kAbTile);
            tma_load_2d(sQ + qs * kAbTile, &tmap_qkv, &qdo_full[qs], (0 * H + h) * kAbD, b * S + i * kAbBlk);
            tma_load_2d(sdO + qs * kAbTile, &tmap_do, &qdo_full[qs], h * kAbD, b * S + i * kAbBlk);
            if (++qs == 2) { qs = 0; qphase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t id_kk = make_idesc_f16(128, 128, 0, 0, 1);   // S^T, dP^T: A K-major, B K-major
    const uint32_t id_kmn = make_idesc_f16(128, 64, 0, 1, 1);   // dV, dK: A K-major (P^T / dS^T), B MN-major (dO / Q)
    const uint32_t id_mnmn = make_idesc_f16(128, 64, 1, 1, 1);  // dQ: A MN-major (dS^T read as dS), B MN-major (K)
    uint32_t kv_phase = 0, qphase = 0, step_phase = 0, j_phase = 0;
    int qs = 0;
    for (int bh = blockIdx.x; bh < total_bh; bh += gridDim.x) {
      for (int j = 0; j < nblk; ++j) {
        mbar_wait(kv_full, kv_phase);
        kv_phase ^= 1;
        mbar_wait(dkv_empty, j_phase ^ 1);       // previous dK_j / dV_j accumulators have been drained
        for (int i = 0; i < nblk; ++i) {
          mbar_wait(&qdo_full[qs], qphase);
          tc_fence_after();
          const uint32_t aK = smem_u32(sK), aV = smem_u32(sV);
          const uint32_t aQ = smem_u32(sQ + qs * kAbTile), aO = smem_u32(sdO + qs * kAbTile);
          if (lane == 0) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {      // reduction over d = 64
              umma_f16_ss(t_S, make_smem_desc_sw128(aK + kk * 32, 16, 1024), make_smem_desc_sw128(aQ + kk * 32, 16, 1024),
                          id_kk, kk > 0);
            }
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              umma_f16_ss(t_dP, make_smem_desc_sw128(aV + kk * 32, 16, 1024), make_smem_desc_sw128(aO + kk * 32, 16, 1024),
                          id_kk, kk > 0);
            }
            umma_commit(s_full);
          }
          __syncwarp();
          mbar_wait(pds_full, step_phase);        // P^T / dS^T are in shared memory
          mbar_wait(dq_empty, step_phase ^ 1);    // previous dQ partial has been read out
          tc_fence_after();
          if (lane == 0) {
            const uint32_t aP = smem_u32(sP), aS = smem_u32(sdS);
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {      // reduction over the 128 queries
              const uint32_t aoff = (kk >> 2) * kAbTile + (kk & 3) * 32;
              umma_f16_ss(t_dV, make_smem_desc_sw128(aP + aoff, 16, 1024),
                          make_smem_desc_sw128(aO + kk * 2048, kAbTile, 1024), id_kmn, (i > 0 || kk > 0));
            }
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
              const uint32_t aoff = (kk >> 2) * kAbTile + (kk & 3) * 32;
              umma_f16_ss(t_dK, make_smem_desc_sw128(aS + aoff, 16, 1024),
                          make_smem_desc_sw128(aQ + kk * 2048, kAbTile, 1024), id_kmn, (i > 0 || kk > 0));
            }
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {      // reduction over the 128 keys
              umma_f16_ss(t_dQ, make_smem_desc_sw128(aS + kk * 2048, kAbTile, 1024),
                          make_smem_desc_sw128(aK + kk * 2048, kAbTile, 1024), id_mnmn, kk > 0);
            }
            umma_commit(&qdo_empty[qs]);
            umma_commit(dq_full);
            if (i == nblk - 1) { umma_commit(dkv_full); umma_commit(kv_empty); }
          }
          __syncwarp();
          step_phase ^= 1;
          if (++qs == 2) { qs = 0; qphase ^= 1; }
        }
        j_phase ^= 1;
      }
    }
  } else if (warp >= 4) {
    // ===================== softmax-backward + read-out warps =====================
    const int quad = warp & 3;
    const int row = quad * 32 + lane;              // TMEM lane: key row for S^T/dP^T/dK/dV, query row for dQ
    const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
    const int tid = threadIdx.x - 128;
    uint32_t step_phase = 0, j_phase = 0;
    float* dq_acc = p.dq_scratch + (size_t)blockIdx.x * S * kAbD;
    for (int bh = blockIdx.x; bh < total_bh; bh += gridDim.x) {
      const int b = bh / H, h = bh % H;
      // ---- stage lse, D = rowsum(dO * O) and the key mask of this (b, h) ----
      asm volatile("bar.sync 1, 128;");
      for (int r = tid; r < S; r += 128) {
        const size_t off = ((size_t)(b * S + r) * H + h) * kAbD;
        float acc = 0.f;
#pragma unroll
        for (int c = 0; c < kAbD; c += 8) {
          const uint4 a = *reinterpret_cast<const uint4*>(p.out + off + c);
          const uint4 g = *reinterpret_cast<const uint4*>(p.dout + off + c);
          float2 x, y;
          x = unpack_bf16x2(a.x); y = unpack_bf16x2(g.x); acc = fmaf(x.x, y.x, acc); acc = fmaf(x.y, y.y, acc);
          x = unpack_bf16x2(a.y); y = unpack_bf16x2(g.y); acc = fmaf(x.x, y.x, acc); acc = fmaf(x.y, y.y, acc);
          x = unpack_bf16x2(a.z); y = unpack_bf16x2(g.z); acc = fmaf(x.x, y.x, acc); acc = fmaf(x.y, y.y, acc);
          x = unpack_bf16x2(a.w); y = unpack_bf16x2(g.w); acc = fmaf(x.x, y.x, acc); acc = fmaf(x.y, y.y, acc);
        }
        sD[r] = acc;
        const float l = p.lse[((size_t)b * H + h) * S + r];
        sLse[r] = (l == -INFINITY) ? INFINITY : l * 1.4426950408889634f;   // fully masked row -> P = 0
        sKeyOk[r] = (p.mask == nullptr || p.mask[(size_t)b * S + r]) ? 1.f : 0.f;
      }
      asm volatile("bar.sync 1, 128;");
      for (int j = 0; j < nblk; ++j) {
        const float key_ok = sKeyOk[j * kAbBlk + row];
        for (int i = 0; i < nblk; ++i) {
          mbar_wait(s_full, step_phase);
          tc_fence_after();
          // ---- P^T and dS^T for key `row` against the 128 queries of block i ----
          uint32_t rsA[32], rpA[32], rsB[32], rpB[32];
          tmem_ld_32x32b_x32(t_S + lane_off, rsA);
          tmem_ld_32x32b_x32(t_dP + lane_off, rpA);
#pragma unroll
          for (int cidx4 = 0; cidx4 < 4; ++cidx4) {
            const int c = cidx4 * 32;
            tmem_ld_wait();
            uint32_t (&rs)[32] = (cidx4 & 1) ? rsB : rsA;
            uint32_t (&rp)[32] = (cidx4 & 1) ? rpB : rpA;
            if (cidx4 < 3) {   // next chunk in flight while this one is processed
              tmem_ld_32x32b_x32(t_S + lane_off + c + 32, (cidx4 & 1) ? rsA : rsB);
              tmem_ld_32x32b_x32(t_dP + lane_off + c + 32, (cidx4 & 1) ? rpA : rpB);
            }
            uint32_t pk[16], dk[16];
#pragma unroll
            for (int q = 0; q < 32; q += 2) {
              const int qi = i * kAbBlk + c + q;
              const float p0 = key_ok * ex2f_b(fmaf(__uint_as_float(rs[q]), p.scale_log2, -sLse[qi]));
              const float p1 = key_ok * ex2f_b(fmaf(__uint_as_float(rs[q + 1]), p.scale_log2, -sLse[qi + 1]));
              const float d0 = p0 * (__uint_as_float(rp[q]) - sD[qi]) * p.scale;
              const float d1 = p1 * (__uint_as_float(rp[q + 1]) - sD[qi + 1]) * p.scale;
              pk[q >> 1] = pack_bf16x2(p0, p1);
              dk[q >> 1] = pack_bf16x2(d0, d1);
            }
            // 32 queries = 64 B = four 16-byte chunks of this key's 128-byte row in query-block (c / 64)
            const uint32_t blk = (c >> 6) * kAbTile + row * 128;
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
              const uint32_t cidx = ((c & 63) >> 3) + ch;                 // 16-byte chunk index inside the row
              const uint32_t off = blk + ((cidx ^ (row & 7)) << 4);      // SWIZZLE_128B
              *reinterpret_cast<uint4*>(sP + off) = make_uint4(pk[4 * ch], pk[4 * ch + 1], pk[4 * ch + 2], pk[4 * ch + 3]);
              *reinterpret_cast<uint4*>(sdS + off) = make_uint4(dk[4 * ch], dk[4 * ch + 1], dk[4 * ch + 2], dk[4 * ch + 3]);
            }
          }
          fence_proxy_async_smem();     // generic-proxy smem writes -> visible to the tensor-core (async) proxy
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(pds_full);
          // ---- dQ partial of query `row` of block i ----
          {
            float* acc = dq_acc + (size_t)(i * kAbBlk + row) * kAbD;
            __nv_bfloat16* dst = p.dqkv + (((size_t)(b * S + i * kAbBlk + row) * 3 + 0) * H + h) * kAbD;
            // the running fp32 sum of this row (L2-resident scratch) is fetched while MMA3-5 are still running
            float4 prev[16];
            if (j > 0) {
#pragma unroll
              for (int q = 0; q < 16; ++q) prev[q] = *reinterpret_cast<const float4*>(acc + 4 * q);
            }
            mbar_wait(dq_full, step_phase);
            tc_fence_after();
            uint32_t r0[32], r1[32];
            tmem_ld_32x32b_x32(t_dQ + lane_off, r0);
            tmem_ld_32x32b_x32(t_dQ + lane_off + 32, r1);
            tmem_ld_wait();
            float v[64];
#pragma unroll
            for (int q = 0; q < 32; ++q) { v[q] = __uint_as_float(r0[q]); v[32 + q] = __uint_as_float(r1[q]); }
            if (j > 0) {
#pragma unroll
              for (int q = 0; q < 16; ++q) {
                v[4 * q] += prev[q].x; v[4 * q + 1] += prev[q].y; v[4 * q + 2] += prev[q].z; v[4 * q + 3] += prev[q].w;
              }
            }
            if (j < nblk - 1) {
#pragma unroll
              for (int q = 0; q < 64; q += 4)
                *reinterpret_cast<float4*>(acc + q) = make_float4(v[q], v[q + 1], v[q + 2], v[q + 3]);
            } else {
#pragma unroll
              for (int q = 0; q < 64; q += 8)
                *reinterpret_cast<uint4*>(dst + q) =
                    make_uint4(pack_bf16x2(v[q], v[q + 1]), pack_bf16x2(v[q + 2], v[q + 3]),
                               pack_bf16x2(v[q + 4], v[q + 5]), pack_bf16x2(v[q + 6], v[q + 7]));
            }
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(dq_empty);
          step_phase ^= 1;
        }
        // ---- dK_j, dV_j of key `row` ----
        mbar_wait(dkv_full, j_phase);
        tc_fence_after();
        {
          const size_t tok = (size_t)(b * S + j * kAbBlk + row);
          __nv_bfloat16* dk_dst = p.dqkv + ((tok * 3 + 1) * H + h) * kAbD;
          __nv_bfloat16* dv_dst = p.dqkv + ((tok * 3 + 2) * H + h) * kAbD;
#pragma unroll
          for (int which = 0; which < 2; ++which) {
            __nv_bfloat16* dst = which == 0 ? dk_dst : dv_dst;
            const uint32_t t = (which == 0 ? t_dK : t_dV) + lane_off;
#pragma unroll
            for (int c = 0; c < kAbD; c += 32) {
              uint32_t r[32];
              tmem_ld_32x32b_x32(t + c, r);
              tmem_ld_wait();
#pragma unroll
              for (int q = 0; q < 32; q += 8)
                *reinterpret_cast<uint4*>(dst + c + q) =
                    make_uint4(pack_bf16x2(__uint_as_float(r[q]), __uint_as_float(r[q + 1])),
                               pack_bf16x2(__uint_as_float(r[q + 2]), __uint_as_float(r[q + 3])),
                               pack_bf16x2(__uint_as_float(r[q + 4]), __uint_as_float(r[q + 5])),
                               pack_bf16x2(__uint_as_float(r[q + 6]), __uint_as_float(r[q + 7])));
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(dkv_empty);
        j_phase ^= 1;
      }
    }
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
  CUtensorMap tq, td;
  int rc = make_tmap_2d_16bit(&tq, qkv, (long long)B * S, 3LL * H * D, 3LL * H * D, 64, 128, 1);
  if (rc) return rc;
  rc = make_tmap_2d_16bit(&td, dout, (long long)B * S, (long long)H * D, (long long)H * D, 64, 128, 1);
  if (rc) return rc;
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    DEDLOC_CUDA_CHECK(cudaGetDevice(&dev));
    DEDLOC_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    DEDLOC_CUDA_CHECK(cudaFuncSetAttribute(attention_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAbSmem));
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
  attention_bwd_kernel<<<grid, kAbThreads, kAbSmem, (cudaStream_t)stream>>>(tq, td, p);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}
