// tcgen05 / TMEM / TMA GEMM with fused epilogues for the ALBERT + SwAV-head hot ops
// (BASELINE.json north-star "no library GEMM").
//
//   C[M,N] = epilogue( sum_k A[m,k] * B[n,k] )        bf16 (or fp16) inputs, fp32 accumulation in TMEM
//
// Operand layouts (runtime flags, so one kernel serves forward, dgrad and wgrad without transposes):
//   A K-major : stored [M,K] row-major          A MN-major : stored [K,M] row-major
//   B K-major : stored [N,K] row-major (nn.Linear weight)   B MN-major : stored [K,N] row-major
//
// Structure (one persistent CTA per SM, 256 threads):
//   warp 0      TMA producer: cp.async.bulk.tensor (SWIZZLE_128B) into a 4-stage smem ring
//   warp 1      MMA issuer: one elected lane issues tcgen05.mma.cta_group::1.kind::f16 (128 x BN x 16),
//               tcgen05.commit releases smem stages and publishes finished accumulators
//   warp 2      TMEM allocator (2 accumulator stages x BN columns = up to all 512 columns)
//   warps 4..7  epilogue: tcgen05.ld (32x32b) -> registers -> fused bias / gelu_new / gelu' / residual /
//               scaling -> bf16 or fp32 stores, or fp32 red.add for split-K weight-gradient accumulation
// so the epilogue of tile i overlaps the mainloop of tile i+1 (double-buffered accumulators).
#include <cuda.h>
#include "gemm_common.cuh"

namespace dedloc {

constexpr int kGemmThreads = 256;
constexpr int BM = 128;
constexpr int BK = 64;  // 64 x 2 bytes = 128 B = one SWIZZLE_128B row
constexpr int kStagesBN256 = 4;
constexpr int kStagesBN128 = 6;

template <int BN>
struct GemmSmem {
  static constexpr int kStages = (BN == 256) ? kStagesBN256 : kStagesBN128;
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarBytes = 256;
  static constexpr int kTotal = kStages * kStageBytes + kBarBytes + 1024;  // +1024 alignment slack
};

template <int BN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const __grid_constant__ GemmKernelParams p) {
  using S = GemmSmem<BN>;
  constexpr int kStages = S::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * S::kStageBytes);
  uint64_t* full_bar = bars;                   // [kStages]
  uint64_t* empty_bar = bars + kStages;        // [kStages]
  uint64_t* tmem_full = bars + 2 * kStages;    // [2]
  uint64_t* tmem_empty = bars + 2 * kStages + 2;  // [2]
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int m_tiles = (p.M + BM - 1) / BM;
  const int n_tiles = (p.N + BN - 1) / BN;
  const int kb_total = (p.K + BK - 1) / BK;   // a K tail is zero-filled by TMA (out-of-bounds box elements)
  const int total_tiles = m_tiles * n_tiles * p.splits;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&tmem_full[s], 1); mbar_init(&tmem_empty[s], 4); }
    mbar_fence_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_base_slot, 2 * BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int split = tile / (m_tiles * n_tiles);
        const int rem = tile - split * (m_tiles * n_tiles);
        int mt, nt;
        tile_coords(rem, m_tiles, n_tiles, mt, nt);
        const int m0 = mt * BM;
        const int n0 = nt * BN;
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb_total, kb0 + p.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * S::kStageBytes;
          uint8_t* sb = sa + S::kABytes;
          mbar_expect_tx(&full_bar[stage], S::kStageBytes);
          const int k0 = kb * BK;
          if (!p.a_mn) {
            tma_load_2d(sa, &tmap_a, &full_bar[stage], k0, m0);
          } else {
#pragma unroll
            for (int i = 0; i < BM / 64; ++i)
              tma_load_2d(sa + i * (BK * 128), &tmap_a, &full_bar[stage], m0 + 64 * i, k0);
          }
          if (!p.b_mn) {
            tma_load_2d(sb, &tmap_b, &full_bar[stage], k0, n0);
          } else {
#pragma unroll
            for (int i = 0; i < BN / 64; ++i)
              tma_load_2d(sb + i * (BK * 128), &tmap_b, &full_bar[stage], n0 + 64 * i, k0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t idesc = make_idesc_f16(BM, BN, p.a_mn, p.b_mn, p.fmt);
    // Descriptors: precomputed low word (address | LBO) + immediate per k16 step; the high word (SBO = 1024,
    // version 1, SWIZZLE_128B) is shared. K-major: 16 elements = 32 B further along the swizzled 128 B row;
    // MN-major: 16 k-rows = 2 groups of 8 rows x 128 B = 2048 B further. elect.sync keeps the operands uniform.
    constexpr uint32_t kHi = 0x40004040u;
    const uint32_t a_step = p.a_mn ? 128u : 2u, b_step = p.b_mn ? 128u : 2u;
    const uint32_t a0 = ((smem_u32(smem) & 0x3FFFF) >> 4) | ((p.a_mn ? (uint32_t)(BK * 128) >> 4 : 1u) << 16);
    const uint32_t b0 = (((smem_u32(smem) + S::kABytes) & 0x3FFFF) >> 4) | ((p.b_mn ? (uint32_t)(BK * 128) >> 4 : 1u) << 16);
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int split = tile / (m_tiles * n_tiles);
      const int kb0 = split * p.kb_per_split;
      const int kb1 = min(kb_total, kb0 + p.kb_per_split);
      mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tb + acc * BN;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = a0 + stage * (S::kStageBytes >> 4), sb = b0 + stage * (S::kStageBytes >> 4);
          umma_f16_ss_lohi(tmem_d, sa, sb, kHi, idesc, kb > kb0 ? 1u : 0u);
#pragma unroll
          for (int j = 1; j < BK / 16; ++j) umma_f16_ss_lohi(tmem_d, sa + j * a_step, sb + j * b_step, kHi, idesc, 1u);
          umma_commit(&empty_bar[stage]);                   // smem stage reusable once these MMAs retire
          if (kb == kb1 - 1) umma_commit(&tmem_full[acc]);  // accumulator complete
        }
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int quad = warp & 3;  // TMEM lane quadrant this warp may read
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int rem = tile % (m_tiles * n_tiles);
      int mt, nt;
      tile_coords(rem, m_tiles, n_tiles, mt, nt);
      const int m0 = mt * BM;
      const int n0 = nt * BN;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const int row = m0 + quad * 32 + lane;
      const bool row_ok = row < p.M;
      const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int cb = 0; cb < BN; cb += 32) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(taddr + cb, r);
        tmem_ld_wait();
        const int col0 = n0 + cb;
        if (col0 >= p.N) continue;  // warp-uniform
        gemm_epilogue_32cols(p, row, row_ok, col0, r);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN);
  }
}

// ------------------------------------------------------------------------------------------------
// Host side: tensor maps through the driver entry point (no link-time libcuda dependency).
// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)f;
  }
  return fn;
}

// 2-D tensor map over a row-major [rows, cols] 16-bit matrix with row stride ld (elements).
int make_tmap_2d_16bit(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld,
                       int box_cols, int box_rows, int is_bf16) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_last_error("cuTensorMapEncodeTiled driver entry point not available");
    return DEDLOC_EUNSUPPORTED;
  }
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                  const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled failed with %d (base %p rows %lld cols %lld ld %lld box %dx%d)", (int)r,
                   base, rows, cols, ld, box_cols, box_rows);
    return DEDLOC_EINVAL;
  }
  return 0;
}

int make_tmap_mn_3d_16bit(CUtensorMap* map, const void* base, long long k_rows, long long mn_cols, long long ld,
                          int box_k, int groups, int is_bf16) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_last_error("cuTensorMapEncodeTiled driver entry point not available");
    return DEDLOC_EUNSUPPORTED;
  }
  cuuint64_t gdim[3] = {64, (cuuint64_t)k_rows, (cuuint64_t)(mn_cols / 64)};
  cuuint64_t gstride[2] = {(cuuint64_t)ld * 2, 128};
  cuuint32_t box[3] = {64, (cuuint32_t)box_k, (cuuint32_t)groups};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(map, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3,
                  const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled (3-D MN-major) failed with %d (base %p k %lld mn %lld ld %lld)", (int)r, base,
                   k_rows, mn_cols, ld);
    return DEDLOC_EINVAL;
  }
  return 0;
}

// cuTensorMapEncodeTiled is a driver call: it needs a context bound to THIS thread (autograd's backward threads
// have none until a runtime call binds the primary context).
int bind_context_for_driver_calls() {
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    int dev = 0;
    DEDLOC_CUDA_CHECK(cudaGetDevice(&dev));
    DEDLOC_CUDA_CHECK(cudaSetDevice(dev));
    DEDLOC_CUDA_CHECK(cudaFree(0));
    ctx_bound = true;
  }
  return 0;
}

template <int BN>
static int launch_gemm(const DedlocGemmArgs& g, cudaStream_t st) {
  GemmKernelParams kp;
  kp.M = g.M; kp.N = g.N; kp.K = g.K;
  kp.a_mn = g.a_mn_major; kp.b_mn = g.b_mn_major;
  kp.epi = g.epilogue; kp.out_kind = g.out_kind; kp.fmt = g.is_fp16 ? 0 : 1;
  kp.a_mn3d = 0; kp.b_mn3d = 0;
  kp.alpha = g.alpha;
  kp.C = g.C; kp.ldc = g.ldc;
  kp.bias = g.bias; kp.residual = g.residual; kp.ld_res = g.ld_residual;
  kp.aux_in = g.aux_in; kp.ld_aux = g.ld_aux; kp.pre_out = g.pre_out; kp.ld_pre = g.ld_pre;
  const int kb_total = (g.K + BK - 1) / BK;
  int splits = g.split_k < 1 ? 1 : g.split_k;
  if (splits > kb_total) splits = kb_total;
  int kb_per = (kb_total + splits - 1) / splits;
  splits = (kb_total + kb_per - 1) / kb_per;
  kp.splits = splits;
  kp.kb_per_split = kb_per;
  DEDLOC_REQUIRE(splits == 1 || g.out_kind == 2, "gemm: split-K needs the fp32 atomic-add output");

  { int rc0 = bind_context_for_driver_calls(); if (rc0) return rc0; }
  CUtensorMap ta, tb;
  int rc;
  if (!g.a_mn_major) rc = make_tmap_2d_16bit(&ta, g.A, g.M, g.K, g.lda, BK, BM, !g.is_fp16);
  else rc = make_tmap_2d_16bit(&ta, g.A, g.K, g.M, g.lda, 64, BK, !g.is_fp16);
  if (rc) return rc;
  if (!g.b_mn_major) rc = make_tmap_2d_16bit(&tb, g.B, g.N, g.K, g.ldb, BK, BN, !g.is_fp16);
  else rc = make_tmap_2d_16bit(&tb, g.B, g.K, g.N, g.ldb, 64, BK, !g.is_fp16);
  if (rc) return rc;

  static int sms = 0;
  static bool attr_set = false;
  if (!attr_set) {
    int dev = 0;
    DEDLOC_CUDA_CHECK(cudaGetDevice(&dev));
    DEDLOC_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    DEDLOC_CUDA_CHECK(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           GemmSmem<BN>::kTotal));
    attr_set = true;
  }
  const int m_tiles = (g.M + BM - 1) / BM, n_tiles = (g.N + BN - 1) / BN;
  const int total = m_tiles * n_tiles * splits;
  int grid = total < sms ? total : sms;
  if (g.max_ctas > 0 && g.max_ctas < grid) grid = g.max_ctas;
  gemm_tcgen05_kernel<BN><<<grid, kGemmThreads, GemmSmem<BN>::kTotal, st>>>(ta, tb, kp);
  count_launch();
  DEDLOC_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace dedloc

namespace dedloc { int launch_gemm_2cta(const DedlocGemmArgs& g, cudaStream_t st); }
using namespace dedloc;

extern "C" int dedloc_gemm(const DedlocGemmArgs* gp, void* stream) {
  const DedlocGemmArgs& g = *gp;
  DEDLOC_REQUIRE(g.M > 0 && g.N > 0 && g.K > 0, "gemm: empty problem %dx%dx%d", g.M, g.N, g.K);
  DEDLOC_REQUIRE(g.K % 8 == 0, "gemm: K (%d) must be a multiple of 8 (16-byte rows); tails of the 64-wide K block are zero-filled", g.K);
  DEDLOC_REQUIRE(g.N % 8 == 0, "gemm: N (%d) must be a multiple of 8", g.N);
  DEDLOC_REQUIRE(g.lda % 8 == 0 && g.ldb % 8 == 0, "gemm: lda/ldb must be multiples of 8 elements (16 B)");
  DEDLOC_REQUIRE(((uintptr_t)g.A & 15) == 0 && ((uintptr_t)g.B & 15) == 0 && ((uintptr_t)g.C & 15) == 0,
                 "gemm: A, B, C must be 16-byte aligned");
  DEDLOC_REQUIRE(g.ldc % 8 == 0, "gemm: ldc must be a multiple of 8");
  if (g.a_mn_major) DEDLOC_REQUIRE(g.M % 8 == 0, "gemm: MN-major A needs M %% 8 == 0");
  cudaStream_t st = (cudaStream_t)stream;
  // CTA pairs halve the L2->SMEM traffic of the B operand; worth it once the problem fills 74 pairs
  // (the pair kernel's two-output epilogue does not combine with row-indexed operands)
  const bool two_out_rowop = (g.epilogue & DEDLOC_EPI_SAVE_PRE) && g.out_kind == 0 &&
                             (g.epilogue & (DEDLOC_EPI_RESIDUAL | DEDLOC_EPI_DGELU | DEDLOC_EPI_MUL_AUX));
  const bool pair_ok = g.block_n != 128 && g.N > 128 && !two_out_rowop;
  const long long tiles_2cta = (long long)((g.M + 255) / 256) * ((g.N + 255) / 256) * (g.split_k > 1 ? g.split_k : 1);
  if (g.cta_pair == 2 || (g.cta_pair == 0 && pair_ok && tiles_2cta >= 64)) {
    DEDLOC_REQUIRE(pair_ok, "gemm: the CTA-pair kernel needs N > 128 and block_n != 128");
    return launch_gemm_2cta(g, st);
  }
  const int bn = g.block_n ? g.block_n : (g.N <= 128 ? 128 : 256);
  if (bn == 128) return launch_gemm<128>(g, st);
  if (bn == 256) return launch_gemm<256>(g, st);
  set_last_error("gemm: block_n must be 128 or 256");
  return DEDLOC_EINVAL;
}
