// K1 / K2 / K3 -- the averaging hot path (BASELINE.json north-star).
//
// ONE cooperative, persistent launch per global step does, over NVLink peer memory (no NCCL):
//
//   barrier 1   every member's encoded gradient payload is in its symmetric send buffer
//   phase A     reduce-scatter: each member reads ITS shard of every member's payload with P2P loads,
//               decodes in registers (fp16 / bf16 / mean-std / 8-bit codebook / block fp8), applies the
//               per-peer sample-count weight, and produces the averaged gradient of the shard.
//                 MEAN (K3): the average is pushed straight into every member's output (all-gather by
//                            P2P stores) -> weighted butterfly all-reduce.
//                 LAMB (K1) / LARS (K2): per-tensor partial sums  sum p^2, sum u^2  (u = Adam direction for
//                            LAMB, averaged gradient for LARS) go to a per-chunk table in fixed order.
//   table       per-tensor partials of this member are pushed to every member's norm table
//   barrier 2   all partial tables have landed
//   phase B     trust ratio per tensor from the summed tables; LAMB / LARC+SGD-momentum update of the
//               OWNED shard; new fp32 master params, their bf16 mirror and (optionally) the optimizer
//               moments are pushed to every member (all-gather by P2P stores). Optional per-row L2
//               re-normalisation (SwAV prototypes, C17) and frozen tensors (FreezeParametersHook).
//   barrier 3   all pushes have landed
//
// Elasticity: the group (members, weights, shard ranges) is an argument of every launch; shard ranges of
// client-mode members are empty; auxiliary members have weight 0. Cross-GPU waits are bounded by
// `timeout_ns` of %globaltimer: block 0 decides, every other block follows its verdict, and the launch
// returns a status word instead of hanging.
//
// Single-GPU testing: `phases` selects A and/or B and `xgpu_barriers = 0` disables the flag waits so that
// P emulated ranks (P buffer sets on one device) can be run as sequential launches (phase A for every
// rank, then phase B for every rank), as the profiling guide requires for dependent ranks on one GPU.
#include <cooperative_groups.h>
#include "codec.cuh"
#include "../include/dedloc_b200.h"

namespace cg = cooperative_groups;

namespace dedloc {

constexpr int kARThreads = 256;

enum ARStatus : uint32_t {
  AR_OK = 0,
  AR_ABORT_B1 = 1,       // timed out before anything was modified: fall back to a local step
  AR_ABORT_B2 = 2,       // timed out waiting for norm tables: nothing committed locally
  AR_ABORT_B3 = 3,       // timed out after pushes: state may be torn -> caller must resync from peers
  AR_SKIPPED_NONFINITE = 4,
};

__device__ __forceinline__ bool any_nonfinite8(const float (&x)[8]) {
  bool bad = false;
#pragma unroll
  for (int i = 0; i < 8; ++i) bad |= !isfinite(x[i]);
  return bad;
}

// Cross-GPU barrier `bar` (1..3). Block 0 signals every member and waits for every member; the verdict is
// published in a local word that all blocks follow. Returns false on abort.
__device__ bool xgpu_barrier(const DedlocARParams& p, int bar, cg::grid_group& grid) {
  __threadfence_system();
  grid.sync();
  const uint32_t tag = (p.tag & ~3u) | (uint32_t)bar;
  uint32_t* go = p.status + 4 + bar;
  if (!p.xgpu_barriers) return true;
  if (blockIdx.x == 0) {
    if (threadIdx.x < p.P) {
      uint32_t* remote = p.flags[threadIdx.x] + bar * DEDLOC_MAX_PEERS + p.slot[p.me];
      st_release_sys(remote, tag);
    }
    bool ok = true;
    if (threadIdx.x < p.P) {
      const uint32_t* mine = p.flags[p.me] + bar * DEDLOC_MAX_PEERS + p.slot[threadIdx.x];
      const uint64_t t0 = globaltimer_ns();
      while (ld_acquire_sys(mine) != tag) {
        if (globaltimer_ns() - t0 > p.timeout_ns) { ok = false; break; }
        __nanosleep(64);
      }
    }
    const int all_ok = __syncthreads_and((int)ok);
    if (threadIdx.x == 0) {
      if (!all_ok) atomicCAS(p.status, AR_OK, (uint32_t)bar);
      __threadfence();
      st_release_gpu(go, all_ok ? tag : ~tag);
    }
  }
  uint32_t v;
  if (threadIdx.x == 0) {
    while (true) {
      v = ld_acquire_gpu(go);
      if (v == tag || v == ~tag) break;
      __nanosleep(32);
    }
  }
  __shared__ uint32_t s_v;
  if (threadIdx.x == 0) s_v = v;
  __syncthreads();
  v = s_v;
  __syncthreads();
  return v == tag;
}

// PB = peers per load batch (2, 4 or 8, the smallest >= the group size): the number of chunks processed per iteration is
// chosen so that every thread keeps ~16 independent 16-byte loads in flight whatever the group size (with two peers and
// two chunks only 32 remote bytes per thread were in flight -- 2.4 MB chip-wide, below NVLink's latency x bandwidth).
template <int CODEC, int MODE, int PB>
__global__ void __launch_bounds__(kARThreads, 2) fused_allreduce_kernel(const __grid_constant__ DedlocARParams p) {
  cg::grid_group grid = cg::this_grid();
  __shared__ float red[64];
  constexpr bool kIsLut = (CODEC == CODEC_UNIFORM_8BIT || CODEC == CODEC_QUANTILE_8BIT);
  __shared__ float lut[kIsLut ? DEDLOC_MAX_PEERS * 256 : 1];

  const int me = p.me;
  const long long cbeg = p.cbeg[me], cend = p.cend[me];
  const long long nmy = cend - cbeg;
  const long long per_cta = (nmy + gridDim.x - 1) / gridDim.x;
  const long long c0 = cbeg + (long long)blockIdx.x * per_cta;
  const long long c1 = (c0 + per_cta < cend) ? c0 + per_cta : cend;
  const int tid = threadIdx.x;

  if (p.phases & 1) {
    if (!xgpu_barrier(p, 1, grid)) return;

    long long cur_cb = -1;
    unsigned int bad_count = 0;
    // Loads of ALL chunks of an iteration from every peer are issued before any is consumed.
    constexpr int kBatch = (CODEC == CODEC_NONE) ? (PB < 4 ? PB : 4) : PB;   // fp32 payloads are 2x uint4 per peer
    constexpr int kUnroll = (CODEC == CODEC_NONE) ? 4 / kBatch : 16 / PB;
    for (long long cbase = c0; cbase < c1;) {
      int nch = (int)((c1 - cbase) < kUnroll ? (c1 - cbase) : kUnroll);
      if constexpr (kIsLut) {
        // all chunks of an iteration must share one codebook block
        const long long blk_end = (cbase / kCodecBlockChunks + 1) * kCodecBlockChunks;
        if (cbase + nch > blk_end) nch = (int)(blk_end - cbase);
        const long long cb = cbase / kCodecBlockChunks;
        if (cb != cur_cb) {
          __syncthreads();
          for (int q = 0; q < p.P; ++q)
            if (p.w[q] != 0.f) lut[q * 256 + tid] = p.aux[q][cb * 256 + tid];
          __syncthreads();
          cur_cb = cb;
        }
      }
      float accs[kUnroll][8];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u)
#pragma unroll
        for (int i = 0; i < 8; ++i) accs[u][i] = 0.f;
      for (int q0 = 0; q0 < p.P; q0 += kBatch) {
        Raw8 raw[kUnroll][kBatch];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
#pragma unroll
          for (int k = 0; k < kBatch; ++k)
            if (u < nch && q0 + k < p.P && p.w[q0 + k] != 0.f)
              load_raw8<CODEC>(p.send[q0 + k], (cbase + u) * kChunk + tid * 8, raw[u][k]);
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
#pragma unroll
          for (int k = 0; k < kBatch; ++k) {
            if (u < nch && q0 + k < p.P && p.w[q0 + k] != 0.f) {
              float x[8];
              decode8<CODEC>(raw[u][k], p.aux[q0 + k], kIsLut ? &lut[(q0 + k) * 256] : nullptr,
                             (cbase + u) * kChunk + tid * 8, cbase + u, x);
              const float w = p.w[q0 + k];
#pragma unroll
              for (int i = 0; i < 8; ++i) accs[u][i] = fmaf(w, x[i], accs[u][i]);
            }
          }
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        if (u >= nch) break;
        const long long c = cbase + u;
        const long long e = c * kChunk + tid * 8;
        float (&acc)[8] = accs[u];
      if (any_nonfinite8(acc)) ++bad_count;
  
        if constexpr (MODE == DEDLOC_AR_MEAN) {
          const uint4 lo = make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3]));
          const uint4 hi = make_uint4(__float_as_uint(acc[4]), __float_as_uint(acc[5]), __float_as_uint(acc[6]), __float_as_uint(acc[7]));
          const uint4 h16 = make_uint4(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]),
                                       pack_bf16x2(acc[4], acc[5]), pack_bf16x2(acc[6], acc[7]));
          for (int q = 0; q < p.P; ++q) {
            if (p.out32[q]) {
              uint4* o = reinterpret_cast<uint4*>(p.out32[q] + e);
              st_stream_v4(o, lo);
              st_stream_v4(o + 1, hi);
            }
            if (p.out16[q]) st_stream_v4(p.out16[q] + e, h16);
          }
        } else {
          float* gs = p.gavg + (e - cbeg * kChunk);
          reinterpret_cast<float4*>(gs)[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
          reinterpret_cast<float4*>(gs)[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
          const float4* pp = reinterpret_cast<const float4*>(p.out32[me] + e);
          const float4 pa = pp[0], pb = pp[1];
          const float pv[8] = {pa.x, pa.y, pa.z, pa.w, pb.x, pb.y, pb.z, pb.w};
          float s1 = 0.f, s2 = 0.f;
          if constexpr (MODE == DEDLOC_AR_LAMB) {
            const float4* mp = reinterpret_cast<const float4*>(p.m[me] + e);
            const float4* vp = reinterpret_cast<const float4*>(p.v[me] + e);
            const float4 ma = mp[0], mb = mp[1], va = vp[0], vb = vp[1];
            const float mv[8] = {ma.x, ma.y, ma.z, ma.w, mb.x, mb.y, mb.z, mb.w};
            const float vv[8] = {va.x, va.y, va.z, va.w, vb.x, vb.y, vb.z, vb.w};
            const float wd = p.t_wd[p.chunk_tid[c]];
  #pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float m1 = p.beta1 * mv[i] + (1.f - p.beta1) * acc[i];
              const float v1 = p.beta2 * vv[i] + (1.f - p.beta2) * acc[i] * acc[i];
              const float u = m1 / (sqrtf(v1) + p.eps) + wd * pv[i];
              s1 = fmaf(pv[i], pv[i], s1);
              s2 = fmaf(u, u, s2);
            }
          } else {
  #pragma unroll
            for (int i = 0; i < 8; ++i) { s1 = fmaf(pv[i], pv[i], s1); s2 = fmaf(acc[i], acc[i], s2); }
          }
          const float2 r = block_sum2(s1, s2, red);
          if (tid == 0) reinterpret_cast<float2*>(p.chunk_stats)[c - cbeg] = r;
        }
      }
      cbase += nch;
    }
    if (bad_count) atomicAdd(p.status + 2, bad_count);

    if constexpr (MODE != DEDLOC_AR_MEAN) {
      __threadfence();
      grid.sync();
      // per-tensor partial sums over the chunks of my shard, one warp per tensor, fixed order
      const int lane = tid & 31;
      const int gw = (blockIdx.x * kARThreads + tid) >> 5, nw = (gridDim.x * kARThreads) >> 5;
      const int stride = 2 * (p.T + 1);
      for (int t = gw; t <= p.T; t += nw) {
        float a = 0.f, b = 0.f;
        if (t < p.T) {
          long long tb = p.tensor_cbeg[t], te = p.tensor_cbeg[t + 1];
          tb = tb > cbeg ? tb : cbeg;
          te = te < cend ? te : cend;
          for (long long c = tb + lane; c < te; c += 32) {
            const float2 s = reinterpret_cast<const float2*>(p.chunk_stats)[c - cbeg];
            a += s.x;
            b += s.y;
          }
          a = warp_sum(a);
          b = warp_sum(b);
        } else {
          a = (float)(*(volatile uint32_t*)(p.status + 2));  // non-finite chunk count of this member
        }
        if (lane == 0)
          for (int q = 0; q < p.P; ++q) {
            float* dst = p.table[q] + (size_t)me * stride + 2 * t;
            dst[0] = a;
            dst[1] = b;
          }
      }
    }
  }

  if constexpr (MODE != DEDLOC_AR_MEAN) {
    if (p.phases & 2) {
      if ((p.phases & 1) || p.xgpu_barriers) {
        if (!xgpu_barrier(p, 2, grid)) return;
      }
      const int stride = 2 * (p.T + 1);
      const float* tab = p.table[me];
      float bad = 0.f;
      for (int q = 0; q < p.P; ++q) bad += tab[(size_t)q * stride + 2 * p.T];
      if (bad > 0.f) {
        if (blockIdx.x == 0 && tid == 0) atomicCAS(p.status, AR_OK, AR_SKIPPED_NONFINITE);
        return;  // uniform across the whole group: every member sees the same tables
      }
      for (long long c = c0; c < c1; ++c) {
        const long long e = c * kChunk + tid * 8;
        const int t = p.chunk_tid[c];
        float s1 = 0.f, s2 = 0.f;
        for (int q = 0; q < p.P; ++q) {
          s1 += tab[(size_t)q * stride + 2 * t];
          s2 += tab[(size_t)q * stride + 2 * t + 1];
        }
        const int tf = p.t_flags[t];
        const float wd = p.t_wd[t];
        const float lr = p.lr * p.t_lr_scale[t];
        const float* gs = p.gavg + (e - cbeg * kChunk);
        const float4 ga = reinterpret_cast<const float4*>(gs)[0], gb = reinterpret_cast<const float4*>(gs)[1];
        const float g[8] = {ga.x, ga.y, ga.z, ga.w, gb.x, gb.y, gb.z, gb.w};
        const float4* pp = reinterpret_cast<const float4*>(p.out32[me] + e);
        const float4 pa = pp[0], pb = pp[1];
        float pv[8] = {pa.x, pa.y, pa.z, pa.w, pb.x, pb.y, pb.z, pb.w};
        float m1[8], v1[8];
        const bool frozen = (tf & DEDLOC_TF_FROZEN) != 0;

        if constexpr (MODE == DEDLOC_AR_LAMB) {
          const float4* mp = reinterpret_cast<const float4*>(p.m[me] + e);
          const float4* vp = reinterpret_cast<const float4*>(p.v[me] + e);
          const float4 ma = mp[0], mb = mp[1], va = vp[0], vb = vp[1];
          const float mv[8] = {ma.x, ma.y, ma.z, ma.w, mb.x, mb.y, mb.z, mb.w};
          const float vv[8] = {va.x, va.y, va.z, va.w, vb.x, vb.y, vb.z, vb.w};
          const float wn = fminf(sqrtf(s1), p.clamp_value);
          const float un = sqrtf(s2);
          const float ratio = (wn == 0.f || un == 0.f) ? 1.f : wn / un;
          const float step = lr * p.bias_correction * ratio;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            m1[i] = p.beta1 * mv[i] + (1.f - p.beta1) * g[i];
            v1[i] = p.beta2 * vv[i] + (1.f - p.beta2) * g[i] * g[i];
            const float u = m1[i] / (sqrtf(v1[i]) + p.eps) + wd * pv[i];
            if (!frozen) pv[i] -= step * u;
            else { m1[i] = mv[i]; v1[i] = vv[i]; }
          }
        } else {  // LARS: apex LARC (clip=False) in front of SGD with momentum
          const float4* mp = reinterpret_cast<const float4*>(p.m[me] + e);
          const float4 ma = mp[0], mb = mp[1];
          const float mv[8] = {ma.x, ma.y, ma.z, ma.w, mb.x, mb.y, mb.z, mb.w};
          const float pn = sqrtf(s1), gn = sqrtf(s2);
          const bool adapt = !(tf & DEDLOC_TF_NO_LARS) && pn != 0.f && gn != 0.f;
          const float alr = adapt ? p.trust * pn / (gn + wd * pn + p.eps) : 1.f;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float gg = adapt ? (g[i] + wd * pv[i]) * alr : g[i];
            m1[i] = p.momentum * mv[i] + gg;
            v1[i] = 0.f;
            if (!frozen) pv[i] -= lr * m1[i];
            else m1[i] = mv[i];
          }
        }
        if (tf & DEDLOC_TF_ROWNORM128) {
          // rows of 128 elements = 16 consecutive threads: L2-normalise each row (SwAV prototypes, C17)
          float ss = 0.f;
#pragma unroll
          for (int i = 0; i < 8; ++i) ss = fmaf(pv[i], pv[i], ss);
#pragma unroll
          for (int o = 8; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
          const float inv = 1.f / fmaxf(sqrtf(ss), 1e-12f);
#pragma unroll
          for (int i = 0; i < 8; ++i) pv[i] *= inv;
        }
        const uint4 lo = make_uint4(__float_as_uint(pv[0]), __float_as_uint(pv[1]), __float_as_uint(pv[2]), __float_as_uint(pv[3]));
        const uint4 hi = make_uint4(__float_as_uint(pv[4]), __float_as_uint(pv[5]), __float_as_uint(pv[6]), __float_as_uint(pv[7]));
        const uint4 h16 = make_uint4(pack_bf16x2(pv[0], pv[1]), pack_bf16x2(pv[2], pv[3]),
                                     pack_bf16x2(pv[4], pv[5]), pack_bf16x2(pv[6], pv[7]));
        const uint4 mlo = make_uint4(__float_as_uint(m1[0]), __float_as_uint(m1[1]), __float_as_uint(m1[2]), __float_as_uint(m1[3]));
        const uint4 mhi = make_uint4(__float_as_uint(m1[4]), __float_as_uint(m1[5]), __float_as_uint(m1[6]), __float_as_uint(m1[7]));
        const uint4 vlo = make_uint4(__float_as_uint(v1[0]), __float_as_uint(v1[1]), __float_as_uint(v1[2]), __float_as_uint(v1[3]));
        const uint4 vhi = make_uint4(__float_as_uint(v1[4]), __float_as_uint(v1[5]), __float_as_uint(v1[6]), __float_as_uint(v1[7]));
        for (int q = 0; q < p.P; ++q) {
          const bool self = (q == me);
          uint4* o = reinterpret_cast<uint4*>(p.out32[q] + e);
          st_stream_v4(o, lo);
          st_stream_v4(o + 1, hi);
          if (p.out16[q]) st_stream_v4(p.out16[q] + e, h16);
          if (self || p.replicate_state) {
            if (p.m[q]) {
              uint4* mo = reinterpret_cast<uint4*>(p.m[q] + e);
              st_stream_v4(mo, mlo);
              st_stream_v4(mo + 1, mhi);
            }
            if (MODE == DEDLOC_AR_LAMB && p.v[q]) {
              uint4* vo = reinterpret_cast<uint4*>(p.v[q] + e);
              st_stream_v4(vo, vlo);
              st_stream_v4(vo + 1, vhi);
            }
          }
        }
      }
    }
  }
  if ((MODE == DEDLOC_AR_MEAN && (p.phases & 1)) || (MODE != DEDLOC_AR_MEAN && (p.phases & 2))) {
    if (p.xgpu_barriers) xgpu_barrier(p, 3, grid);
  }
}

template <int CODEC, int MODE, int PB>
static int launch_ar_pb(const DedlocARParams& p, cudaStream_t st) {
  static int max_blocks = 0;  // per (CODEC, MODE) instantiation; all devices on the box are identical
  if (max_blocks == 0) {
    int dev = 0, sms = 0, occ = 0;
    DEDLOC_CUDA_CHECK(cudaGetDevice(&dev));
    DEDLOC_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    DEDLOC_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fused_allreduce_kernel<CODEC, MODE, PB>,
                                                                    kARThreads, 0));
    DEDLOC_REQUIRE(occ >= 1, "fused_allreduce: kernel does not fit on an SM");
    max_blocks = sms * (occ > 2 ? 2 : occ);
  }
  long long nmy = p.cend[p.me] - p.cbeg[p.me];
  int blocks = max_blocks;
  if (p.max_blocks > 0 && p.max_blocks < blocks) blocks = p.max_blocks;
  if (nmy < blocks) blocks = nmy < 1 ? 1 : (int)nmy;
  void* args[] = {(void*)&p};
  DEDLOC_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)fused_allreduce_kernel<CODEC, MODE, PB>, dim3(blocks),
                                                dim3(kARThreads), args, 0, st));
  count_launch();
  return 0;
}

template <int CODEC, int MODE>
static int launch_ar(const DedlocARParams& p, cudaStream_t st) {
  if (p.P <= 2) return launch_ar_pb<CODEC, MODE, 2>(p, st);
  if (p.P <= 4) return launch_ar_pb<CODEC, MODE, 4>(p, st);
  return launch_ar_pb<CODEC, MODE, 8>(p, st);
}

template <int MODE>
static int dispatch_codec(const DedlocARParams& p, cudaStream_t st) {
  switch (p.codec) {
    case CODEC_NONE: return launch_ar<CODEC_NONE, MODE>(p, st);
    case CODEC_FLOAT16: return launch_ar<CODEC_FLOAT16, MODE>(p, st);
    case CODEC_BFLOAT16: return launch_ar<CODEC_BFLOAT16, MODE>(p, st);
    case CODEC_MEANSTD_16BIT: return launch_ar<CODEC_MEANSTD_16BIT, MODE>(p, st);
    case CODEC_UNIFORM_8BIT: return launch_ar<CODEC_UNIFORM_8BIT, MODE>(p, st);
    case CODEC_QUANTILE_8BIT: return launch_ar<CODEC_QUANTILE_8BIT, MODE>(p, st);
    case CODEC_BLOCKWISE_FP8: return launch_ar<CODEC_BLOCKWISE_FP8, MODE>(p, st);
  }
  set_last_error("fused_allreduce: unknown codec %d", p.codec);
  return DEDLOC_EINVAL;
}

}  // namespace dedloc

using namespace dedloc;

extern "C" int dedloc_fused_allreduce(const DedlocARParams* pp, void* stream) {
  const DedlocARParams& p = *pp;
  DEDLOC_REQUIRE(p.P >= 1 && p.P <= DEDLOC_MAX_PEERS, "fused_allreduce: group size %d out of range", p.P);
  DEDLOC_REQUIRE(p.me >= 0 && p.me < p.P, "fused_allreduce: my index %d not in group of %d", p.me, p.P);
  DEDLOC_REQUIRE(p.status != nullptr, "fused_allreduce: status buffer is required");
  DEDLOC_REQUIRE((p.phases & 3) != 0, "fused_allreduce: no phase selected");
  for (int q = 0; q < p.P; ++q) {
    DEDLOC_REQUIRE(p.cbeg[q] <= p.cend[q], "fused_allreduce: bad shard of member %d", q);
    DEDLOC_REQUIRE(p.send[q] != nullptr || p.w[q] == 0.f, "fused_allreduce: member %d has weight but no payload", q);
  }
  if (p.mode != DEDLOC_AR_MEAN) {
    DEDLOC_REQUIRE(p.gavg && p.chunk_stats && p.chunk_tid && p.tensor_cbeg && p.t_wd && p.t_lr_scale && p.t_flags,
                   "fused_allreduce: optimizer modes need gavg/chunk_stats/chunk_tid/tensor tables");
    DEDLOC_REQUIRE(p.m[p.me] != nullptr, "fused_allreduce: optimizer modes need the local moment buffer");
  }
  cudaStream_t st = (cudaStream_t)stream;
  // words [0..3] of the status pad (verdict, non-finite counter) restart at zero with every phase A
  if (p.phases & 1) DEDLOC_CUDA_CHECK(cudaMemsetAsync(p.status, 0, 4 * sizeof(uint32_t), st));
  switch (p.mode) {
    case DEDLOC_AR_MEAN: return dispatch_codec<DEDLOC_AR_MEAN>(p, st);
    case DEDLOC_AR_LAMB: return dispatch_codec<DEDLOC_AR_LAMB>(p, st);
    case DEDLOC_AR_LARS: return dispatch_codec<DEDLOC_AR_LARS>(p, st);
  }
  set_last_error("fused_allreduce: unknown mode %d", p.mode);
  return DEDLOC_EINVAL;
}

extern "C" int dedloc_ar_params_size(void) { return (int)sizeof(DedlocARParams); }
