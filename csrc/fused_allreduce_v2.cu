// K1 / K2 / K3, second generation: the averaging hot path as a push-based, stage-pipelined dataflow kernel
// (BASELINE.json north-star). Same reference call sites as
// fused_allreduce.cu: hivemind's averager.step() + torch_optimizer.Lamb / apex LARC+SGD
// (albert/run_trainer.py:86-94,251-264; swav/ClassyVision/classy_vision/optim/sgd_collaborative.py:135-157).
//
// What changed against the first kernel, and why:
//   * reduce-scatter by PUSH. Every member stores its encoded gradient into the owner's staging slot with posted
//     NVLink writes; the owner then reads local memory only. (P2P loads pay a round trip per request.)
//   * interleaved ownership: chunk c (2048 elements) belongs to owner c % K. Any contiguous range of chunks is
//     spread evenly over the owners, so the vector can be cut into NS stages that every member works on at once.
//   * four CTA roles run concurrently on different stages and hand stages to each other, and to the other GPUs,
//     through round-tagged flags -- no grid-wide or cross-GPU barrier anywhere:
//        PUSH(s)    encode + store to the owners; the last CTA of the stage raises inflag[s][me] at every owner
//        REDUCE(s)  wait inflag[s][*]; decode + weighted sum (+ MEAN: store the average to every member); LAMB/LARS:
//                   per-chunk partial norms; the last CTA of the stage publishes the per-tensor sums of every tensor
//                   that ENDS in the stage to every member as self-validating {value, tag} 64-bit words
//        UPDATE(s)  wait for the words of the tensors in the stage from every owner; moments; store the update
//                   direction (fp16 u for LAMB, bf16 momentum for LARS) into every member's delta buffer; the last
//                   CTA raises outflag[s][me] at every member
//        APPLY(s)   wait outflag[s][*]; p32 -= scale_t * direction; bf16 mirror; optional row normalisation
//     so the NVLink stores of PUSH and UPDATE, the local reduction and the HBM-bound APPLY overlap stage by stage,
//     and a tensor's update starts as soon as ITS norms are complete (not after a barrier over all tensors).
//     Two schedules of the same roles (Params::sequential): CONCURRENT splits the CTAs into four role groups that
//     work on different stages at once; SEQUENTIAL (the default) lets every CTA run PUSH, REDUCE, UPDATE, APPLY in
//     turn over ONE stage, so each phase has the whole chip. Measured (profiles/traces/, profiles/README.md): the
//     thread that completes a stage pays a system-scope fence that waits for ALL traffic in flight on the GPU
//     (7-25 us under load, even when its own stores were local), so every additional stage adds latency that the
//     overlap does not win back at these sizes, and a role group that owns a quarter of the SMs cannot saturate
//     NVLink or HBM on its own. The hand-over between phases stays flag-based either way (no grid.sync, no barrier
//     across GPUs): a CTA that has finished its share of a phase polls the flags of the next one (raised by the last
//     CTA of every sender for that stage) while other CTAs, here or on other GPUs, are still working.
//   * 2 bytes per element leave the owner instead of 6 (fp32 master + bf16 mirror) or 14 (with moments): every
//     member applies the same rounded direction to its own replica of the fp32 master, which therefore stays
//     bit-identical on all members; moments live with the owner (replicate_state pushes them as well).
//   * a tensor whose averaged gradient is not finite is skipped by every member alike (its norm words are NaN
//     everywhere); the other tensors step. Status word 4 reports it.
//
// Failure model: every wait is bounded by timeout_ns of %globaltimer. The member that gives up raises the abort word
// of every member; all CTAs leave at their next wait. status[0] = 1 when nothing local was modified (the caller steps
// on local gradients), 3 when UPDATE/APPLY had begun (the caller resynchronises from a peer).
//
// `n_ranks` > 1 runs that many members as ONE cooperative launch over buffers on one GPU (ranks that wait on each
// other must not be separate launches on one device): the single-GPU tests run the production protocol unchanged.
#include "codec.cuh"
#include "../include/dedloc_b200.h"

namespace dedloc {
namespace ar2 {

using Params = DedlocAR2Params;
constexpr int kThreads = 256;
constexpr int kMaxS = DEDLOC_AR2_MAX_STAGES;
constexpr int kUnrollPush = 4;
constexpr int kUnrollApply = 2;
constexpr int kPeerBatch = 4;     // staged payloads in flight per thread in REDUCE
constexpr int kCtasPerSm = 4;     // <= 64 registers per thread: 32 warps per SM hide the load latency of the streaming roles
constexpr int kFlagStride = DEDLOC_MAX_PEERS;

template <int CODEC>
struct CodecTraits {
  static constexpr int kBpe = (CODEC == CODEC_NONE) ? 4
                              : (CODEC == CODEC_FLOAT16 || CODEC == CODEC_BFLOAT16 || CODEC == CODEC_MEANSTD_16BIT) ? 2 : 1;
  static constexpr bool kLut = (CODEC == CODEC_UNIFORM_8BIT || CODEC == CODEC_QUANTILE_8BIT);
  static constexpr bool kFly = (CODEC == CODEC_NONE || CODEC == CODEC_FLOAT16 || CODEC == CODEC_BFLOAT16);
};

// ---- memory-model helpers ------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 ld_cg_v4(const void* p) {  // L2 only: data produced by other SMs / GPUs in this launch
  uint4 r;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ld_cg_v2(const void* p) {
  uint2 r;
  asm volatile("ld.global.cg.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ float ld_cg_f32(const float* p) {
  float r;
  asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// ---- bounded waits ---------------------------------------------------------------------------------------------
// status[1] holds the tag of the last round that modified this member's state (no reset between rounds needed)
__device__ __noinline__ void note_abort(const Params& p) {
  const uint32_t modified = *(volatile uint32_t*)(p.status + 1) == p.tag;
  atomicCAS(p.status, 0u, modified ? 3u : 1u);
}
__device__ __noinline__ void raise_abort(const Params& p) {
  note_abort(p);
  for (int q = 0; q < p.P; ++q) st_release_sys(p.abort_word[q], p.tag);
}
__device__ __forceinline__ bool saw_abort(const Params& p) {
  if (ld_relaxed_sys_u32(p.abort_word[p.me]) != p.tag) return false;
  note_abort(p);
  return true;
}
// One thread waits until *f == tag. false = this member gives up (timeout here, or somebody raised the abort word).
// Polls with relaxed loads (an acquire per poll is a system-scope fence per poll, which slows every fence on the
// chip while CTAs idle); one acquire load orders the data reads after the flag has been seen.
__device__ bool wait_flag(const uint32_t* f, const Params& p) {
  if (ld_relaxed_sys_u32(f) != p.tag) {
    const uint64_t t0 = globaltimer_ns();
    for (int spins = 1;; ++spins) {
      if (ld_relaxed_sys_u32(f) == p.tag) break;
      if ((spins & 31) == 0) {
        if (saw_abort(p)) return false;
        if (globaltimer_ns() - t0 > p.timeout_ns) { raise_abort(p); return false; }
      }
      __nanosleep(100);
    }
  }
  return ld_acquire_sys(f) == p.tag;
}
// {float value | tag << 32}: the word validates itself, so no fence orders it against anything.
__device__ bool wait_word(const unsigned long long* w, const Params& p, float& val) {
  const uint64_t t0 = globaltimer_ns();
  for (int spins = 1;; ++spins) {
    const unsigned long long x = ld_relaxed_sys_u64(w);
    if ((uint32_t)(x >> 32) == p.tag) { val = __uint_as_float((uint32_t)x); return true; }
    if ((spins & 31) == 0) {
      if (saw_abort(p)) return false;
      if (globaltimer_ns() - t0 > p.timeout_ns) { raise_abort(p); return false; }
    }
    __nanosleep(40);
  }
}
// optional timeline: CTA 0 of every role stamps %globaltimer at up to four points of every stage
__device__ __forceinline__ void stamp(const Params& p, int role, int b, int s, int slot) {
  if (p.trace != nullptr && b == 0 && threadIdx.x == 0) p.trace[((size_t)role * kMaxS + s) * 4 + slot] = globaltimer_ns();
}
__device__ __forceinline__ void fence_sys(const Params& p) {
  if (!(p.debug_flags & 1)) __threadfence_system();
}
__device__ __forceinline__ unsigned long long pack_word(float v, uint32_t tag) {
  return (unsigned long long)__float_as_uint(v) | ((unsigned long long)tag << 32);
}

// Summed norms (sum p^2, sum u^2) of tensor t from the words every owner sent to this member. CTA-wide; `sm` holds 32 floats.
__device__ bool tensor_norms(const Params& p, int t, float* sm, float& s1, float& s2) {
  __syncthreads();
  bool ok = true;
  const int tid = threadIdx.x;
  if (tid < p.K) {
    const unsigned long long* w = p.table[p.me] + ((size_t)tid * p.T + t) * 2;
    float a = 0.f, b = 0.f;
    ok = wait_word(w, p, a) && wait_word(w + 1, p, b);
    sm[tid] = a;
    sm[16 + tid] = b;
  }
  if (!__syncthreads_and((int)ok)) return false;
  float a = 0.f, b = 0.f;
  for (int k = 0; k < p.K; ++k) { a += sm[k]; b += sm[16 + k]; }   // same order on every member
  s1 = a;
  s2 = b;
  return true;
}

// ---- payload helpers -------------------------------------------------------------------------------------------
template <int CODEC>
__device__ __forceinline__ void load_raw8_cg(const uint8_t* payload, long long e, Raw8& r) {
  if constexpr (CodecTraits<CODEC>::kBpe == 4) {
    r.a = ld_cg_v4(payload + e * 4);
    r.b = ld_cg_v4(payload + e * 4 + 16);
  } else if constexpr (CodecTraits<CODEC>::kBpe == 2) {
    r.a = ld_cg_v4(payload + e * 2);
  } else {
    const uint2 t = ld_cg_v2(payload + e);
    r.a.x = t.x;
    r.a.y = t.y;
  }
}
template <int CODEC>
__device__ __forceinline__ void store_raw8(uint8_t* payload, long long e, const Raw8& r) {
  if constexpr (CodecTraits<CODEC>::kBpe == 4) {
    st_stream_v4(payload + e * 4, r.a);
    st_stream_v4(payload + e * 4 + 16, r.b);
  } else if constexpr (CodecTraits<CODEC>::kBpe == 2) {
    st_stream_v4(payload + e * 2, r.a);
  } else {
    st_stream_v2(payload + e, make_uint2(r.a.x, r.a.y));
  }
}
// Raw8 holding 8 fp32 values -> encoded in place (on-the-fly codecs); bit-identical to encode_chunk_kernel.
template <int CODEC>
__device__ __forceinline__ void encode_fly(Raw8& r, float scale) {
  float x[8] = {__uint_as_float(r.a.x) * scale, __uint_as_float(r.a.y) * scale, __uint_as_float(r.a.z) * scale,
                __uint_as_float(r.a.w) * scale, __uint_as_float(r.b.x) * scale, __uint_as_float(r.b.y) * scale,
                __uint_as_float(r.b.z) * scale, __uint_as_float(r.b.w) * scale};
  if constexpr (CODEC == CODEC_NONE) {
    r.a = make_uint4(__float_as_uint(x[0]), __float_as_uint(x[1]), __float_as_uint(x[2]), __float_as_uint(x[3]));
    r.b = make_uint4(__float_as_uint(x[4]), __float_as_uint(x[5]), __float_as_uint(x[6]), __float_as_uint(x[7]));
  } else if constexpr (CODEC == CODEC_FLOAT16) {
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = x[i] > 65504.f ? 65504.f : (x[i] < -65504.f ? -65504.f : x[i]);
    r.a = make_uint4(pack_f16x2(x[0], x[1]), pack_f16x2(x[2], x[3]), pack_f16x2(x[4], x[5]), pack_f16x2(x[6], x[7]));
  } else {
    r.a = make_uint4(pack_bf16x2(x[0], x[1]), pack_bf16x2(x[2], x[3]), pack_bf16x2(x[4], x[5]), pack_bf16x2(x[6], x[7]));
  }
}
// decode8 of codec.cuh with the side information passed by value (it was written by a peer during this launch, so
// it must not travel through the non-coherent path).
template <int CODEC>
__device__ __forceinline__ void decode8_v2(const Raw8& r, float a0, float a1, const float* lut, float (&x)[8]) {
  if constexpr (CODEC == CODEC_MEANSTD_16BIT) {
    decode8<CODEC_FLOAT16>(r, nullptr, nullptr, 0, 0, x);
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = fmaf(x[i], a1, a0);
  } else if constexpr (CODEC == CODEC_BLOCKWISE_FP8) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      x[i] = fp8_e4m3_to_float((r.a.x >> (8 * i)) & 0xFF) * a0;
      x[4 + i] = fp8_e4m3_to_float((r.a.y >> (8 * i)) & 0xFF) * a0;
    }
  } else {
    decode8<CODEC>(r, nullptr, lut, 0, 0, x);
  }
}

__device__ __forceinline__ int first_owned(int lo, int kme, int K) {  // first chunk >= lo owned by owner kme
  return lo + (((kme - lo % K) % K) + K) % K;
}

// ---- PUSH ------------------------------------------------------------------------------------------------------
template <int CODEC>
__device__ void role_push(const Params& p, int b) {
  using CT = CodecTraits<CODEC>;
  const int tid = threadIdx.x, me = p.me, K = p.K;
  if (p.w[me] == 0.f) return;   // auxiliary members contribute no gradient; REDUCE does not wait for them
  const bool fly = CT::kFly && p.src_f32;
  for (int s = 0; s < p.NS; ++s) {
    const int s0 = (int)(s * p.Ls);
    const int s1 = (int)((s0 + p.Ls < p.C) ? s0 + p.Ls : p.C);
    stamp(p, 0, b, s, 0);
    for (int c0 = s0 + b; c0 < s1; c0 += p.nP * kUnrollPush) {
      Raw8 raw[kUnrollPush];
#pragma unroll
      for (int j = 0; j < kUnrollPush; ++j) {
        const int c = c0 + j * p.nP;
        if (c < s1) {
          const long long e = (long long)c * kChunk + tid * 8;
          if (fly) {
            raw[j].a = ld_stream_v4(p.acc + e);
            raw[j].b = ld_stream_v4(p.acc + e + 4);
          } else {
            load_raw8<CODEC>(p.send, e, raw[j]);
          }
        }
      }
#pragma unroll
      for (int j = 0; j < kUnrollPush; ++j) {
        const int c = c0 + j * p.nP;
        if (c < s1) {
          if constexpr (CT::kFly) {
            if (fly) encode_fly<CODEC>(raw[j], p.scale);
          }
          const int owner = (p.debug_flags & 4) ? me : p.owner_of_k[c % K];   // 4: keep every store local (diagnostic)
          const long long el = (long long)(c / K) * kChunk + tid * 8;
          store_raw8<CODEC>(p.stage_in[owner] + (size_t)me * p.slot_bytes, el, raw[j]);
          float* ax = p.aux_in[owner] + (size_t)me * p.aux_slot_floats;
          if constexpr (CODEC == CODEC_MEANSTD_16BIT) {
            if (tid < 2) ax[2 * (size_t)c + tid] = p.aux[2 * (size_t)c + tid];
          } else if constexpr (CODEC == CODEC_BLOCKWISE_FP8) {
            if (tid < kChunk / kFp8Block) ax[(size_t)c * (kChunk / kFp8Block) + tid] = p.aux[(size_t)c * (kChunk / kFp8Block) + tid];
          } else if constexpr (CT::kLut) {
            if (c % kCodecBlockChunks == 0) {   // stages begin at multiples of 32 chunks: every codebook start is seen
              const size_t o = (size_t)(c / kCodecBlockChunks) * 256 + tid;
              const float val = p.aux[o];
              for (int k = 0; k < K; ++k) p.aux_in[p.owner_of_k[k]][(size_t)me * p.aux_slot_floats + o] = val;
            }
          }
        }
      }
    }
    // The warps of a PUSH CTA never wait for each other: each one moves its 256-element slice of the CTA's chunks
    // and reports the stage on its own. gpu-scope release by every warp, ONE system-scope fence by the warp that
    // completes the stage: causality is transitive across scopes, and a system fence per CTA and stage costs
    // 10-40 us when 147 of them run at once.
    __syncwarp();
    stamp(p, 0, b, s, 1);
    if (p.trace != nullptr && (tid & 31) == 0 && (tid >> 5) == 0 && b < 1024 && s < 4) {
      // per-CTA view (warp 0): [SM id, stored s0..s3, fenced s0..s3] after the role traces
      unsigned smid;
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
      unsigned long long* t2 = p.trace + 4 * kMaxS * 4 + (size_t)b * 10;
      t2[0] = smid;
      t2[1 + s] = globaltimer_ns();
    }
    if ((tid & 31) == 0) {
      if (!(p.debug_flags & 2)) __threadfence();      // 2: omit the per-warp release fence (diagnostic)
      if (p.trace != nullptr && (tid >> 5) == 0 && b < 1024 && s < 4) p.trace[4 * kMaxS * 4 + (size_t)b * 10 + 5 + s] = globaltimer_ns();
      stamp(p, 0, b, s, 2);
      const unsigned old = atomicAdd(p.counters + 0 * kMaxS + s, 1u);
      if (old == (unsigned)p.nP * (kThreads / 32) - 1) {
        p.counters[0 * kMaxS + s] = 0;     // the last warp of the stage leaves the counter ready for the next round
        fence_sys(p);
        for (int k = 0; k < K; ++k) st_release_sys(p.inflag[p.owner_of_k[k]] + s * kFlagStride + me, p.tag);
        if (p.trace) p.trace[((size_t)0 * kMaxS + s) * 4 + 3] = globaltimer_ns();
      }
    }
  }
}

// ---- REDUCE ----------------------------------------------------------------------------------------------------
template <int CODEC, int MODE>
__device__ void role_reduce(const Params& p, int b, float* red, float* lut, int* s_last) {
  using CT = CodecTraits<CODEC>;
  const int tid = threadIdx.x, me = p.me, K = p.K, kme = p.kidx[me];
  if (kme < 0) return;
  unsigned bad = 0;
  int cur_cb = -1;
  const uint8_t* stage = p.stage_in[me];
  const float* auxs = p.aux_in[me];
  // LAMB / LARS: every thread keeps running sums (sum p^2, sum u^2) over the chunks of ONE tensor; they are folded
  // into this CTA's slot {a, b, tag} of that tensor when the tensor changes and at the end of every stage. The CTA
  // that completes a stage adds the slots of all REDUCE CTAs, in CTA order, for the tensors that end in it.
  float run1 = 0.f, run2 = 0.f;
  int run_t = -1;
  auto flush = [&]() {
    if (run_t < 0) return;
    const float2 r = block_sum2(run1, run2, red);
    if (tid == 0) {
      float4* slot = reinterpret_cast<float4*>(p.chunk_stats) + (size_t)run_t * p.nR + b;
      float4 cur = *slot;
      if (__float_as_uint(cur.z) != p.tag) cur = make_float4(0.f, 0.f, __uint_as_float(p.tag), 0.f);
      cur.x += r.x;
      cur.y += r.y;
      *slot = cur;
    }
    run1 = 0.f;
    run2 = 0.f;
  };
  for (int s = 0; s < p.NS; ++s) {
    const int s0 = (int)(s * p.Ls);
    const int s1 = (int)((s0 + p.Ls < p.C) ? s0 + p.Ls : p.C);
    bool ok = true;
    stamp(p, 1, b, s, 0);
    if (tid < p.P && p.w[tid] != 0.f) ok = wait_flag(p.inflag[me] + s * kFlagStride + tid, p);
    if (!__syncthreads_and((int)ok)) return;
    stamp(p, 1, b, s, 1);
    for (int c = first_owned(s0, kme, K) + b * K; c < s1; c += p.nR * K) {
      const long long el = (long long)(c / K) * kChunk + tid * 8;
      const long long e = (long long)c * kChunk + tid * 8;
      if constexpr (MODE != DEDLOC_AR_MEAN) {
        const int t = p.chunk_tid[c];
        if (t != run_t) {
          flush();
          run_t = t;
        }
      }
      if constexpr (CT::kLut) {
        const int cb = c / kCodecBlockChunks;
        if (cb != cur_cb) {
          __syncthreads();
          for (int q = 0; q < p.P; ++q)
            if (p.w[q] != 0.f) lut[q * 256 + tid] = ld_cg_f32(auxs + (size_t)q * p.aux_slot_floats + (size_t)cb * 256 + tid);
          __syncthreads();
          cur_cb = cb;
        }
      }
      float acc[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = 0.f;
      for (int q0 = 0; q0 < p.P; q0 += kPeerBatch) {
        Raw8 raw[kPeerBatch];
        float a0[kPeerBatch], a1[kPeerBatch];
#pragma unroll
        for (int k = 0; k < kPeerBatch; ++k) {
          const int q = q0 + k;
          a0[k] = 0.f;
          a1[k] = 0.f;
          if (q < p.P && p.w[q] != 0.f) {
            load_raw8_cg<CODEC>(stage + (size_t)q * p.slot_bytes, el, raw[k]);
            if constexpr (CODEC == CODEC_MEANSTD_16BIT) {
              a0[k] = ld_cg_f32(auxs + (size_t)q * p.aux_slot_floats + 2 * (size_t)c);
              a1[k] = ld_cg_f32(auxs + (size_t)q * p.aux_slot_floats + 2 * (size_t)c + 1);
            } else if constexpr (CODEC == CODEC_BLOCKWISE_FP8) {
              a0[k] = ld_cg_f32(auxs + (size_t)q * p.aux_slot_floats + e / kFp8Block);
            }
          }
        }
#pragma unroll
        for (int k = 0; k < kPeerBatch; ++k) {
          const int q = q0 + k;
          if (q < p.P && p.w[q] != 0.f) {
            float x[8];
            decode8_v2<CODEC>(raw[k], a0[k], a1[k], CT::kLut ? &lut[q * 256] : nullptr, x);
            const float w = p.w[q];
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[i] = fmaf(w, x[i], acc[i]);
          }
        }
      }
      bool nf = false;
#pragma unroll
      for (int i = 0; i < 8; ++i) nf |= !isfinite(acc[i]);
      if (nf) ++bad;
      if constexpr (MODE == DEDLOC_AR_MEAN) {
        const uint4 lo = make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3]));
        const uint4 hi = make_uint4(__float_as_uint(acc[4]), __float_as_uint(acc[5]), __float_as_uint(acc[6]), __float_as_uint(acc[7]));
        const uint4 h16 = make_uint4(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]),
                                     pack_bf16x2(acc[4], acc[5]), pack_bf16x2(acc[6], acc[7]));
        for (int q = 0; q < p.P; ++q) {
          if (p.out32[q]) {
            st_stream_v4(p.out32[q] + e, lo);
            st_stream_v4(p.out32[q] + e + 4, hi);
          }
          if (p.out16[q]) st_stream_v4(p.out16[q] + e, h16);
        }
      } else {
        reinterpret_cast<float4*>(p.gavg + e)[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
        reinterpret_cast<float4*>(p.gavg + e)[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
        const float4 pa = reinterpret_cast<const float4*>(p.p32 + e)[0], pb = reinterpret_cast<const float4*>(p.p32 + e)[1];
        const float pv[8] = {pa.x, pa.y, pa.z, pa.w, pb.x, pb.y, pb.z, pb.w};
        float n1 = 0.f, n2 = 0.f;
        if constexpr (MODE == DEDLOC_AR_LAMB) {
          const float4 ma = reinterpret_cast<const float4*>(p.m + e)[0], mb = reinterpret_cast<const float4*>(p.m + e)[1];
          const float4 va = reinterpret_cast<const float4*>(p.v + e)[0], vb = reinterpret_cast<const float4*>(p.v + e)[1];
          const float mv[8] = {ma.x, ma.y, ma.z, ma.w, mb.x, mb.y, mb.z, mb.w};
          const float vv[8] = {va.x, va.y, va.z, va.w, vb.x, vb.y, vb.z, vb.w};
          const float wd = p.t_wd[p.chunk_tid[c]];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float m1 = p.beta1 * mv[i] + (1.f - p.beta1) * acc[i];
            const float v1 = p.beta2 * vv[i] + (1.f - p.beta2) * acc[i] * acc[i];
            const float u = m1 / (sqrtf(v1) + p.eps) + wd * pv[i];
            n1 = fmaf(pv[i], pv[i], n1);
            n2 = fmaf(u, u, n2);
          }
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) { n1 = fmaf(pv[i], pv[i], n1); n2 = fmaf(acc[i], acc[i], n2); }
        }
        run1 += n1;
        run2 += n2;
      }
    }
    // hand the stage over
    if constexpr (MODE != DEDLOC_AR_MEAN) flush();
    __syncthreads();
    stamp(p, 1, b, s, 2);
    if (tid == 0) {
      __threadfence();
      const unsigned old = atomicAdd(p.counters + 1 * kMaxS + s, 1u);
      *s_last = (old == (unsigned)p.nR - 1);
      if (*s_last) {
        p.counters[1 * kMaxS + s] = 0;
        if (p.trace) p.trace[((size_t)1 * kMaxS + s) * 4 + 3] = globaltimer_ns();
        if (MODE == DEDLOC_AR_MEAN) {
          fence_sys(p);
          for (int q = 0; q < p.P; ++q) st_release_sys(p.outflag[q] + s * kFlagStride + kme, p.tag);
        } else {
          __threadfence();
        }
      }
    }
    __syncthreads();
    if constexpr (MODE != DEDLOC_AR_MEAN) {
      if (*s_last) {
        // every REDUCE CTA of this member has finished stages <= s: tensors that END in (s0, s1] are complete here
        const int warp = tid >> 5, lane = tid & 31;
        for (int t = warp; t < p.T; t += kThreads / 32) {
          const int te = (int)p.tensor_cbeg[t + 1];
          if (te > s0 && te <= s1) {
            const int tb = (int)p.tensor_cbeg[t];
            float a = 0.f, bsum = 0.f;
            (void)tb;
            for (int i = lane; i < p.nR; i += 32) {
              const float4 st = __ldcg(reinterpret_cast<const float4*>(p.chunk_stats) + (size_t)t * p.nR + i);
              if (__float_as_uint(st.z) == p.tag) {   // CTAs that saw no chunk of this tensor have a stale slot
                a += st.x;
                bsum += st.y;
              }
            }
            a = warp_sum(a);
            bsum = warp_sum(bsum);
            if (lane < p.P) {
              unsigned long long* dst = p.table[lane] + ((size_t)kme * p.T + t) * 2;
              st_relaxed_sys_u64(dst, pack_word(a, p.tag));
              st_relaxed_sys_u64(dst + 1, pack_word(bsum, p.tag));
            }
          }
        }
      }
      __syncthreads();   // *s_last is rewritten in the next stage
    }
  }
  if (bad && MODE == DEDLOC_AR_MEAN) atomicAdd(p.status + 2, bad);
}

// ---- UPDATE ----------------------------------------------------------------------------------------------------
template <int MODE>
__device__ void role_update(const Params& p, int b, float* red) {
  const int tid = threadIdx.x, me = p.me, K = p.K, kme = p.kidx[me];
  if (kme < 0) return;
  int cur_t = -1;
  float s1 = 0.f, s2 = 0.f;
  bool marked = false;
  for (int s = 0; s < p.NS; ++s) {
    const int s0 = (int)(s * p.Ls);
    const int s1c = (int)((s0 + p.Ls < p.C) ? s0 + p.Ls : p.C);
    stamp(p, 2, b, s, 0);
    for (int c = first_owned(s0, kme, K) + b * K; c < s1c; c += p.nU * K) {
      const int t = p.chunk_tid[c];
      if (t != cur_t) {
        if (!tensor_norms(p, t, red, s1, s2)) return;
        cur_t = t;
      }
      const int tf = p.t_flags[t];
      if ((tf & DEDLOC_TF_FROZEN) || !isfinite(s1) || !isfinite(s2)) continue;   // APPLY uses scale 0 for this tensor
      if (!marked) {
        if (tid == 0) atomicExch(p.status + 1, p.tag);
        marked = true;
      }
      const long long e = (long long)c * kChunk + tid * 8;
      const float wd = p.t_wd[t];
      const uint4 g0 = ld_cg_v4(p.gavg + e), g1 = ld_cg_v4(p.gavg + e + 4);
      const float g[8] = {__uint_as_float(g0.x), __uint_as_float(g0.y), __uint_as_float(g0.z), __uint_as_float(g0.w),
                          __uint_as_float(g1.x), __uint_as_float(g1.y), __uint_as_float(g1.z), __uint_as_float(g1.w)};
      const float4 pa = reinterpret_cast<const float4*>(p.p32 + e)[0], pb = reinterpret_cast<const float4*>(p.p32 + e)[1];
      const float pv[8] = {pa.x, pa.y, pa.z, pa.w, pb.x, pb.y, pb.z, pb.w};
      const float4 ma = reinterpret_cast<const float4*>(p.m + e)[0], mb = reinterpret_cast<const float4*>(p.m + e)[1];
      const float mv[8] = {ma.x, ma.y, ma.z, ma.w, mb.x, mb.y, mb.z, mb.w};
      float m1[8], v1[8], d[8];
      uint4 d16;
      if constexpr (MODE == DEDLOC_AR_LAMB) {
        const float4 va = reinterpret_cast<const float4*>(p.v + e)[0], vb = reinterpret_cast<const float4*>(p.v + e)[1];
        const float vv[8] = {va.x, va.y, va.z, va.w, vb.x, vb.y, vb.z, vb.w};
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          m1[i] = p.beta1 * mv[i] + (1.f - p.beta1) * g[i];
          v1[i] = p.beta2 * vv[i] + (1.f - p.beta2) * g[i] * g[i];
          const float u = m1[i] / (sqrtf(v1[i]) + p.eps) + wd * pv[i];
          d[i] = u > 65504.f ? 65504.f : (u < -65504.f ? -65504.f : u);
        }
        d16 = make_uint4(pack_f16x2(d[0], d[1]), pack_f16x2(d[2], d[3]), pack_f16x2(d[4], d[5]), pack_f16x2(d[6], d[7]));
      } else {  // LARS: apex LARC (clip=False) in front of SGD with momentum
        const float pn = sqrtf(s1), gn = sqrtf(s2);
        const bool adapt = !(tf & DEDLOC_TF_NO_LARS) && pn != 0.f && gn != 0.f;
        const float alr = adapt ? p.trust * pn / (gn + wd * pn + p.eps) : 1.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float gg = adapt ? (g[i] + wd * pv[i]) * alr : g[i];
          m1[i] = p.momentum * mv[i] + gg;
          v1[i] = 0.f;
          d[i] = m1[i];
        }
        d16 = make_uint4(pack_bf16x2(d[0], d[1]), pack_bf16x2(d[2], d[3]), pack_bf16x2(d[4], d[5]), pack_bf16x2(d[6], d[7]));
      }
      for (int q = 0; q < p.P; ++q) st_stream_v4(p.delta[q] + e, d16);
      const uint4 mlo = make_uint4(__float_as_uint(m1[0]), __float_as_uint(m1[1]), __float_as_uint(m1[2]), __float_as_uint(m1[3]));
      const uint4 mhi = make_uint4(__float_as_uint(m1[4]), __float_as_uint(m1[5]), __float_as_uint(m1[6]), __float_as_uint(m1[7]));
      const uint4 vlo = make_uint4(__float_as_uint(v1[0]), __float_as_uint(v1[1]), __float_as_uint(v1[2]), __float_as_uint(v1[3]));
      const uint4 vhi = make_uint4(__float_as_uint(v1[4]), __float_as_uint(v1[5]), __float_as_uint(v1[6]), __float_as_uint(v1[7]));
      for (int q = 0; q < p.P; ++q) {
        if (q != me && !p.replicate_state) continue;
        float* mq = (q == me) ? p.m : p.m_all[q];
        float* vq = (q == me) ? p.v : p.v_all[q];
        if (mq) {
          st_stream_v4(mq + e, mlo);
          st_stream_v4(mq + e + 4, mhi);
        }
        if (MODE == DEDLOC_AR_LAMB && vq) {
          st_stream_v4(vq + e, vlo);
          st_stream_v4(vq + e + 4, vhi);
        }
      }
    }
    __syncthreads();
    stamp(p, 2, b, s, 1);
    if (tid == 0) {
      __threadfence();
      stamp(p, 2, b, s, 2);
      const unsigned old = atomicAdd(p.counters + 2 * kMaxS + s, 1u);
      if (old == (unsigned)p.nU - 1) {
        p.counters[2 * kMaxS + s] = 0;
        fence_sys(p);
        for (int q = 0; q < p.P; ++q) st_release_sys(p.outflag[q] + s * kFlagStride + kme, p.tag);
        if (p.trace) p.trace[((size_t)2 * kMaxS + s) * 4 + 3] = globaltimer_ns();
      }
    }
  }
}

// ---- APPLY -----------------------------------------------------------------------------------------------------
template <int MODE>
__device__ void role_apply(const Params& p, int b, float* red) {
  const int tid = threadIdx.x, me = p.me, K = p.K;
  if constexpr (MODE == DEDLOC_AR_MEAN) {
    // nothing to apply: the averages were stored straight into the outputs; wait until all of them have landed
    for (int s = 0; s < p.NS; ++s) {
      bool ok = true;
      stamp(p, 3, b, s, 0);
      if (tid < K) ok = wait_flag(p.outflag[me] + s * kFlagStride + tid, p);
      if (!__syncthreads_and((int)ok)) return;
      stamp(p, 3, b, s, 1);
    }
    return;
  } else {
    int cur_t = -1, cur_tf = 0;
    float scale = 0.f;
    unsigned skipped = 0;
    bool marked = false;
    for (int s = 0; s < p.NS; ++s) {
      const int s0 = (int)(s * p.Ls);
      const int s1 = (int)((s0 + p.Ls < p.C) ? s0 + p.Ls : p.C);
      bool ok = true;
      stamp(p, 3, b, s, 0);
      if (tid < K) ok = wait_flag(p.outflag[me] + s * kFlagStride + tid, p);
      if (!__syncthreads_and((int)ok)) return;
      stamp(p, 3, b, s, 1);
      for (int c0 = s0 + b; c0 < s1; c0 += p.nA * kUnrollApply) {
        float sc[kUnrollApply];
        int fl[kUnrollApply];
#pragma unroll
        for (int j = 0; j < kUnrollApply; ++j) {
          const int c = c0 + j * p.nA;
          sc[j] = 0.f;
          fl[j] = 0;
          if (c < s1) {
            const int t = p.chunk_tid[c];
            if (t != cur_t) {
              float n1, n2;
              if (!tensor_norms(p, t, red, n1, n2)) return;
              cur_t = t;
              cur_tf = p.t_flags[t];
              const float lr = p.lr * p.t_lr_scale[t];
              if (!isfinite(n1) || !isfinite(n2)) {
                scale = 0.f;
                if (c == (int)p.tensor_cbeg[t]) ++skipped;   // counted by exactly one CTA of this member
              } else if (cur_tf & DEDLOC_TF_FROZEN) {
                scale = 0.f;
              } else if (MODE == DEDLOC_AR_LAMB) {
                const float wn = fminf(sqrtf(n1), p.clamp_value), un = sqrtf(n2);
                const float ratio = (wn == 0.f || un == 0.f) ? 1.f : wn / un;
                scale = lr * p.bias_correction * ratio;
              } else {
                scale = lr;
              }
            }
            sc[j] = scale;
            fl[j] = cur_tf;
          }
        }
        uint4 dd[kUnrollApply];
        float4 pa[kUnrollApply], pb[kUnrollApply];
#pragma unroll
        for (int j = 0; j < kUnrollApply; ++j) {
          if (sc[j] != 0.f) {
            const long long e = (long long)(c0 + j * p.nA) * kChunk + tid * 8;
            dd[j] = ld_cg_v4(p.delta[me] + e);
            pa[j] = reinterpret_cast<const float4*>(p.p32 + e)[0];
            pb[j] = reinterpret_cast<const float4*>(p.p32 + e)[1];
          }
        }
#pragma unroll
        for (int j = 0; j < kUnrollApply; ++j) {
          if (sc[j] != 0.f) {
            if (!marked) {
              if (tid == 0) atomicExch(p.status + 1, p.tag);
              marked = true;
            }
            const long long e = (long long)(c0 + j * p.nA) * kChunk + tid * 8;
            float d[8];
            float2 t2;
            if constexpr (MODE == DEDLOC_AR_LAMB) {
              t2 = unpack_f16x2(dd[j].x); d[0] = t2.x; d[1] = t2.y;
              t2 = unpack_f16x2(dd[j].y); d[2] = t2.x; d[3] = t2.y;
              t2 = unpack_f16x2(dd[j].z); d[4] = t2.x; d[5] = t2.y;
              t2 = unpack_f16x2(dd[j].w); d[6] = t2.x; d[7] = t2.y;
            } else {
              t2 = unpack_bf16x2(dd[j].x); d[0] = t2.x; d[1] = t2.y;
              t2 = unpack_bf16x2(dd[j].y); d[2] = t2.x; d[3] = t2.y;
              t2 = unpack_bf16x2(dd[j].z); d[4] = t2.x; d[5] = t2.y;
              t2 = unpack_bf16x2(dd[j].w); d[6] = t2.x; d[7] = t2.y;
            }
            float pv[8] = {pa[j].x, pa[j].y, pa[j].z, pa[j].w, pb[j].x, pb[j].y, pb[j].z, pb[j].w};
#pragma unroll
            for (int i = 0; i < 8; ++i) pv[i] = fmaf(-sc[j], d[i], pv[i]);
            if (fl[j] & DEDLOC_TF_ROWNORM128) {
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
            reinterpret_cast<float4*>(p.p32 + e)[0] = make_float4(pv[0], pv[1], pv[2], pv[3]);
            reinterpret_cast<float4*>(p.p32 + e)[1] = make_float4(pv[4], pv[5], pv[6], pv[7]);
            if (p.p16)
              *reinterpret_cast<uint4*>(p.p16 + e) = make_uint4(pack_bf16x2(pv[0], pv[1]), pack_bf16x2(pv[2], pv[3]),
                                                                pack_bf16x2(pv[4], pv[5]), pack_bf16x2(pv[6], pv[7]));
          }
        }
      }
      stamp(p, 3, b, s, 2);
    }
    if (skipped && tid == 0) {
      atomicAdd(p.status + 2, skipped);
      atomicCAS(p.status, 0u, 4u);
    }
  }
}

template <int CODEC, int MODE>
__device__ __forceinline__ void run_role(const Params& p, int b, float* red, float* lut, int* s_last) {
  if (p.debug_flags >> 8) {
    // test hook: this member starts late, by a pseudo-random time below (debug_flags >> 8) * 16 us that depends on
    // the round tag (bit 7 set: by exactly that time) -- arrival skew for the flag soak, a parked kernel for the
    // "peer dies inside the kernel" test
    if (threadIdx.x == 0) {
      const unsigned long long span = (unsigned long long)((unsigned)p.debug_flags >> 8) * 16000ull;
      const unsigned long long wait = (p.debug_flags & 128) ? span : ((p.tag * 2654435761u + (unsigned)p.me * 40503u) % 1000u) * span / 1000u;
      const unsigned long long t0 = globaltimer_ns();
      while (globaltimer_ns() - t0 < wait) __nanosleep(200);
    }
    __syncthreads();
  }
  if (p.sequential) {
    // Every CTA runs the four roles one after the other over ALL stages (nP = nR = nU = nA = grid): each phase has
    // the whole chip, which hides load latency far better than a quarter of the CTAs per role, and both the gradient
    // and the update stores leave through the same NVLink direction anyway, so running them side by side gains no
    // bandwidth. Waits stay stage-granular: a phase starts on stage 0 while the peers are still finishing the last
    // stages of their previous phase. No role waits for a later role, so the order cannot deadlock.
    role_push<CODEC>(p, b);
    role_reduce<CODEC, MODE>(p, b, red, lut, s_last);
    if (saw_abort(p)) return;
    if constexpr (MODE != DEDLOC_AR_MEAN) {
      role_update<MODE>(p, b, red);
      if (saw_abort(p)) return;
      role_apply<MODE>(p, b, red);
    } else if (b == 0) {
      role_apply<MODE>(p, b, red);
    }
    return;
  }
  if (b < p.nP) {
    role_push<CODEC>(p, b);
  } else if (b < p.nP + p.nR) {
    role_reduce<CODEC, MODE>(p, b - p.nP, red, lut, s_last);
  } else if (b < p.nP + p.nR + p.nU) {
    if constexpr (MODE != DEDLOC_AR_MEAN) role_update<MODE>(p, b - p.nP - p.nR, red);
  } else {
    role_apply<MODE>(p, b - p.nP - p.nR - p.nU, red);
  }
}

// Production: one member per launch, its parameters by value (no copy, no memset precedes the launch).
template <int CODEC, int MODE>
__global__ void __launch_bounds__(kThreads, kCtasPerSm) ar2_kernel(const __grid_constant__ Params p) {
  __shared__ float red[64];
  __shared__ float lut[CodecTraits<CODEC>::kLut ? DEDLOC_MAX_PEERS * 256 : 1];
  __shared__ int s_last;
  run_role<CODEC, MODE>(p, blockIdx.x, red, lut, &s_last);
}

// Tests: several members in one cooperative launch on one GPU; every CTA copies its member's parameters to smem.
template <int CODEC, int MODE>
__global__ void __launch_bounds__(kThreads, kCtasPerSm) ar2_kernel_ranks(const Params* __restrict__ pa, int ctas_per_rank) {
  __shared__ Params sp;
  __shared__ float red[64];
  __shared__ float lut[CodecTraits<CODEC>::kLut ? DEDLOC_MAX_PEERS * 256 : 1];
  __shared__ int s_last;
  const int rank = blockIdx.x / ctas_per_rank;
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(pa + rank);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&sp);
    for (int i = threadIdx.x; i < (int)(sizeof(Params) / 4); i += kThreads) dst[i] = src[i];
  }
  __syncthreads();
  run_role<CODEC, MODE>(sp, blockIdx.x % ctas_per_rank, red, lut, &s_last);
}

template <int CODEC, int MODE>
static int max_ctas() {
  int dev = 0, sms = 0, occ = 0, occ_r = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ar2_kernel<CODEC, MODE>, kThreads, 0) != cudaSuccess) return 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_r, ar2_kernel_ranks<CODEC, MODE>, kThreads, 0) != cudaSuccess) return 0;
  if (occ_r < occ) occ = occ_r;
  return sms * (occ > kCtasPerSm ? kCtasPerSm : occ);
}

template <int CODEC, int MODE>
static int launch(const Params* host_params, const Params* dev_params, int n_ranks, int ctas_per_rank, cudaStream_t st) {
  const int cap = max_ctas<CODEC, MODE>();
  DEDLOC_REQUIRE(n_ranks * ctas_per_rank <= cap, "fused_allreduce_v2: %d CTAs requested, %d can be co-resident",
                 n_ranks * ctas_per_rank, cap);
  // a cooperative launch only for its co-residency guarantee (CTAs of different roles wait on each other)
  if (n_ranks == 1) {
    void* args[] = {(void*)host_params};
    DEDLOC_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)ar2_kernel<CODEC, MODE>, dim3(ctas_per_rank),
                                                  dim3(kThreads), args, 0, st));
  } else {
    DEDLOC_CUDA_CHECK(cudaMemcpyAsync((void*)dev_params, host_params, (size_t)n_ranks * sizeof(Params),
                                      cudaMemcpyHostToDevice, st));
    void* args[] = {(void*)&dev_params, (void*)&ctas_per_rank};
    DEDLOC_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)ar2_kernel_ranks<CODEC, MODE>,
                                                  dim3(n_ranks * ctas_per_rank), dim3(kThreads), args, 0, st));
  }
  count_launch();
  return 0;
}

template <int MODE>
static int dispatch(int codec, const Params* host_params, const Params* dev_params, int n_ranks, int ctas,
                    cudaStream_t st, bool query, int* out) {
#define DEDLOC_AR2_CASE(C)                                                   \
  case C:                                                                    \
    if (query) { *out = max_ctas<C, MODE>(); return 0; }                     \
    return launch<C, MODE>(host_params, dev_params, n_ranks, ctas, st);
  switch (codec) {
    DEDLOC_AR2_CASE(CODEC_NONE)
    DEDLOC_AR2_CASE(CODEC_FLOAT16)
    DEDLOC_AR2_CASE(CODEC_BFLOAT16)
    DEDLOC_AR2_CASE(CODEC_MEANSTD_16BIT)
    DEDLOC_AR2_CASE(CODEC_UNIFORM_8BIT)
    DEDLOC_AR2_CASE(CODEC_QUANTILE_8BIT)
    DEDLOC_AR2_CASE(CODEC_BLOCKWISE_FP8)
  }
#undef DEDLOC_AR2_CASE
  set_last_error("fused_allreduce_v2: unknown codec %d", codec);
  return DEDLOC_EINVAL;
}

static int dispatch_mode(int mode, int codec, const Params* host_params, const Params* dev_params, int n_ranks,
                         int ctas, cudaStream_t st, bool query, int* out) {
  switch (mode) {
    case DEDLOC_AR_MEAN: return dispatch<DEDLOC_AR_MEAN>(codec, host_params, dev_params, n_ranks, ctas, st, query, out);
    case DEDLOC_AR_LAMB: return dispatch<DEDLOC_AR_LAMB>(codec, host_params, dev_params, n_ranks, ctas, st, query, out);
    case DEDLOC_AR_LARS: return dispatch<DEDLOC_AR_LARS>(codec, host_params, dev_params, n_ranks, ctas, st, query, out);
  }
  set_last_error("fused_allreduce_v2: unknown mode %d", mode);
  return DEDLOC_EINVAL;
}

}  // namespace ar2
}  // namespace dedloc

using namespace dedloc;

extern "C" int dedloc_fused_allreduce_v2(const DedlocAR2Params* params, int n_ranks, void* params_dev, void* stream) {
  DEDLOC_REQUIRE(params != nullptr && (params_dev != nullptr || n_ranks == 1) && n_ranks >= 1 &&
                     n_ranks <= DEDLOC_MAX_PEERS,
                 "fused_allreduce_v2: bad arguments");
  const DedlocAR2Params& p0 = params[0];
  const int ctas = p0.sequential ? p0.nP : p0.nP + p0.nR + p0.nU + p0.nA;
  cudaStream_t st = (cudaStream_t)stream;
  for (int r = 0; r < n_ranks; ++r) {
    const DedlocAR2Params& p = params[r];
    DEDLOC_REQUIRE(p.P >= 2 && p.P <= DEDLOC_MAX_PEERS, "fused_allreduce_v2: group size %d out of range", p.P);
    DEDLOC_REQUIRE(p.me >= 0 && p.me < p.P, "fused_allreduce_v2: my index %d not in a group of %d", p.me, p.P);
    DEDLOC_REQUIRE(p.K >= 1 && p.K <= p.P, "fused_allreduce_v2: %d owners in a group of %d", p.K, p.P);
    DEDLOC_REQUIRE(p.tag != 0, "fused_allreduce_v2: the round tag must not be zero");
    DEDLOC_REQUIRE(p.mode == p0.mode && p.codec == p0.codec, "fused_allreduce_v2: members disagree on mode / codec");
    if (p.sequential) {
      DEDLOC_REQUIRE(p0.sequential && p.nP == p0.nP && p.nP >= 1 && p.nA == p.nP &&
                         (p.nR == p.nP || p.nR == 0) && (p.nU == p.nP || p.nU == 0),
                     "fused_allreduce_v2: sequential phases need the same CTA count for every role");
    } else {
      DEDLOC_REQUIRE(p.nP + p.nR + p.nU + p.nA == ctas && p.nP >= 1 && p.nA >= 1,
                     "fused_allreduce_v2: role split (%d, %d, %d, %d) does not add up to %d", p.nP, p.nR, p.nU, p.nA, ctas);
    }
    DEDLOC_REQUIRE(p.NS >= 1 && p.NS <= DEDLOC_AR2_MAX_STAGES && p.Ls % kCodecBlockChunks == 0 && p.Ls * p.NS >= p.C &&
                       p.C >= 1,
                   "fused_allreduce_v2: bad stage geometry (NS %d, Ls %lld, C %lld)", p.NS, p.Ls, p.C);
    const bool owner = p.kidx[p.me] >= 0;
    DEDLOC_REQUIRE(!owner || p.nR >= 1, "fused_allreduce_v2: an owner needs REDUCE CTAs");
    DEDLOC_REQUIRE(owner || p.nR == 0 || p.sequential, "fused_allreduce_v2: REDUCE CTAs on a member that owns nothing");
    if (p.mode != DEDLOC_AR_MEAN) {
      DEDLOC_REQUIRE(!owner || p.nU >= 1, "fused_allreduce_v2: an owner needs UPDATE CTAs");
      DEDLOC_REQUIRE(p.p32 && p.gavg && p.chunk_stats && p.chunk_tid && p.tensor_cbeg && p.t_wd && p.t_lr_scale && p.t_flags,
                     "fused_allreduce_v2: optimizer modes need the parameter / tensor tables");
      DEDLOC_REQUIRE(!owner || (p.m != nullptr && (p.mode != DEDLOC_AR_LAMB || p.v != nullptr)),
                     "fused_allreduce_v2: an owner needs its moment buffers");
    }
    DEDLOC_REQUIRE(p.status && p.counters, "fused_allreduce_v2: status / counters are required");
    DEDLOC_REQUIRE(p.w[p.me] == 0.f || (p.src_f32 ? p.acc != nullptr : p.send != nullptr),
                   "fused_allreduce_v2: a contributing member needs its gradient source");
  }
  // No memset precedes the launch: the stage counters reset themselves (the CTA that completes a stage zeroes its
  // counter) and the status words are only written on the rare paths (abort, skipped tensors). The caller keeps both
  // zero otherwise: it clears them after reading a non-zero status (PeerBuffers.read_status).
  return ar2::dispatch_mode(p0.mode, p0.codec, params, (const DedlocAR2Params*)params_dev, n_ranks, ctas, st, false,
                            nullptr);
}

extern "C" int dedloc_ar2_params_size(void) { return (int)sizeof(DedlocAR2Params); }

extern "C" int dedloc_ar2_max_ctas(int mode, int codec) {
  int out = 0;
  if (ar2::dispatch_mode(mode, codec, nullptr, nullptr, 0, 0, nullptr, true, &out) != 0) return 0;
  return out;
}
