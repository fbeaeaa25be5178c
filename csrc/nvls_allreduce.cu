// NVLS variant of the weighted-mean all-reduce (K3) -- SURVEY 5.8 / X1: the reduction happens INSIDE the NVSwitch.
//
// Every member's encoded (fp16 or bf16, already multiplied by its normalised sample-count weight) gradient sits in
// its NvlsWindow (dedloc_b200/parallel/nvls.py), and all windows are bound to one multicast object. Then
//   multimem.ld_reduce  on the multicast address of a chunk returns the SUM over all members' copies: one load
//                       brings N/P reduced elements into this GPU instead of (P-1) N/P raw ones;
//   multimem.st         on the multicast address of the result stores to ALL members at once: N/P elements leave this
//                       GPU instead of (P-1) N/P;
//   multimem.red        increments a barrier counter in every member's flag page with one instruction.
// Per GPU and direction about N payload bytes cross NVLink, against 2 (P-1)/P N for the peer-to-peer kernels.
// Same reference call site as the other averaging kernels (hivemind averager.step, albert/run_trainer.py:251-264).
// Only multimem loads / stores / reductions touch multicast addresses (no TMA).
//
// Protocol of one launch (cooperative: the CTAs wait on counters that other GPUs increment):
//   arrive A, wait A     every member's payload is in its window (written by the encoder that precedes this launch)
//   reduce               member `me` handles chunks c with c % P == me
//   arrive B, wait B     every member's stores have landed everywhere
//   (optional) copy the bf16 result out of the window into out16 / out32
// The counters only grow: target = round * P * gridDim, so nothing is reset between rounds. Waits are bounded by
// timeout_ns; status[0] = 1 on timeout (nothing outside the window was modified).
#include "codec.cuh"
#include "../include/dedloc_b200.h"

namespace dedloc {
namespace nvls {

constexpr int kThreads = 256;

__device__ __forceinline__ void multimem_red_add_release(uint32_t* mc_addr, uint32_t v) {
  asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(mc_addr), "r"(v) : "memory");
}
// sum over all members of 8 halves (4 x f16x2 / bf16x2), accumulated in fp32 inside the switch
template <bool BF16>
__device__ __forceinline__ uint4 multimem_ld_reduce8(const void* mc_addr) {
  uint4 r;
  if constexpr (BF16) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(mc_addr)
                 : "memory");
  } else {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(mc_addr)
                 : "memory");
  }
  return r;
}
__device__ __forceinline__ void multimem_st8_bf16(void* mc_addr, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(mc_addr), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

// One thread: add 1 to the counter on every member, then wait until the local copy has reached `target`.
__device__ bool barrier_all(uint32_t* mc_counter, const uint32_t* local_counter, uint32_t target,
                            unsigned long long timeout_ns) {
  multimem_red_add_release(mc_counter, 1u);
  const uint64_t t0 = globaltimer_ns();
  for (int spins = 1;; ++spins) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(local_counter) : "memory");
    if ((int32_t)(v - target) >= 0) break;
    if ((spins & 31) == 0 && globaltimer_ns() - t0 > timeout_ns) return false;
    __nanosleep(100);
  }
  return ld_acquire_sys(local_counter) != 0xFFFFFFFFu || true;   // one acquire orders what follows after the counter
}

template <bool BF16_IN>
__global__ void __launch_bounds__(kThreads, 4) nvls_mean_kernel(const __grid_constant__ DedlocNvlsParams p) {
  const int tid = threadIdx.x;
  uint8_t* mc = reinterpret_cast<uint8_t*>(p.mc_base);
  uint8_t* uc = reinterpret_cast<uint8_t*>(p.uc_base);
  const uint32_t target = p.round * (uint32_t)p.P * gridDim.x;
  __shared__ int s_ok;
  if (tid == 0)
    s_ok = barrier_all(reinterpret_cast<uint32_t*>(mc + p.off_flags), reinterpret_cast<const uint32_t*>(uc + p.off_flags),
                       target, p.timeout_ns);
  __syncthreads();
  if (!s_ok) {
    if (tid == 0) atomicCAS(p.status, 0u, 1u);
    return;
  }
  const long long n_chunks = p.n / kChunk;
  constexpr int kU = 4;   // switch round trips in flight per thread
  const long long stride = (long long)gridDim.x * p.P;
  for (long long c0 = p.me + (long long)blockIdx.x * p.P; c0 < n_chunks; c0 += stride * kU) {
    uint4 sum[kU];
#pragma unroll
    for (int j = 0; j < kU; ++j) {
      const long long c = c0 + j * stride;
      if (c < n_chunks) sum[j] = multimem_ld_reduce8<BF16_IN>(mc + p.off_payload + (c * kChunk + tid * 8) * 2);
    }
#pragma unroll
    for (int j = 0; j < kU; ++j) {
      const long long c = c0 + j * stride;
      if (c < n_chunks) {
        uint4 out = sum[j];
        if constexpr (!BF16_IN) {   // fp16 payload -> bf16 result
          float2 a = unpack_f16x2(out.x), b = unpack_f16x2(out.y), cc = unpack_f16x2(out.z), d = unpack_f16x2(out.w);
          out = make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(b.x, b.y), pack_bf16x2(cc.x, cc.y), pack_bf16x2(d.x, d.y));
        }
        multimem_st8_bf16(mc + p.off_result + (c * kChunk + tid * 8) * 2, out);
      }
    }
  }
  __syncthreads();
  if (tid == 0) {
    __threadfence_system();
    s_ok = barrier_all(reinterpret_cast<uint32_t*>(mc + p.off_flags + 128),
                       reinterpret_cast<const uint32_t*>(uc + p.off_flags + 128), target, p.timeout_ns);
  }
  __syncthreads();
  if (!s_ok) {
    if (tid == 0) atomicCAS(p.status, 0u, 1u);
    return;
  }
  if (p.out16 == nullptr && p.out32 == nullptr) return;
  // the complete result is in my window now: hand it to the caller's buffers
  for (long long c = blockIdx.x; c < n_chunks; c += gridDim.x) {
    const long long e = c * kChunk + tid * 8;
    uint4 r;
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(uc + p.off_result + e * 2));
    if (p.out16) *reinterpret_cast<uint4*>(p.out16 + e) = r;
    if (p.out32) {
      const float2 a = unpack_bf16x2(r.x), b = unpack_bf16x2(r.y), cc = unpack_bf16x2(r.z), d = unpack_bf16x2(r.w);
      reinterpret_cast<float4*>(p.out32 + e)[0] = make_float4(a.x, a.y, b.x, b.y);
      reinterpret_cast<float4*>(p.out32 + e)[1] = make_float4(cc.x, cc.y, d.x, d.y);
    }
  }
}

}  // namespace nvls
}  // namespace dedloc

using namespace dedloc;

extern "C" int dedloc_nvls_allreduce_mean(const DedlocNvlsParams* pp, void* stream) {
  const DedlocNvlsParams& p = *pp;
  DEDLOC_REQUIRE(p.mc_base && p.uc_base && p.status, "nvls_allreduce: window / status pointers are required");
  DEDLOC_REQUIRE(p.P >= 1 && p.P <= DEDLOC_MAX_PEERS && p.me >= 0 && p.me < p.P, "nvls_allreduce: bad group (%d of %d)",
                 p.me, p.P);
  DEDLOC_REQUIRE(p.n > 0 && p.n % kChunk == 0, "nvls_allreduce: n (%lld) must be a multiple of %d", p.n, kChunk);
  DEDLOC_REQUIRE(p.round >= 1 && p.ctas >= 1, "nvls_allreduce: round and ctas must be positive");
  int dev = 0, sms = 0, occ = 0;
  DEDLOC_CUDA_CHECK(cudaGetDevice(&dev));
  DEDLOC_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const void* fn = p.payload_is_bf16 ? (const void*)nvls::nvls_mean_kernel<true> : (const void*)nvls::nvls_mean_kernel<false>;
  DEDLOC_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, nvls::kThreads, 0));
  DEDLOC_REQUIRE(p.ctas <= sms * occ, "nvls_allreduce: %d CTAs requested, %d can be co-resident", p.ctas, sms * occ);
  void* args[] = {(void*)pp};
  DEDLOC_CUDA_CHECK(cudaLaunchCooperativeKernel(fn, dim3(p.ctas), dim3(nvls::kThreads), args, 0, (cudaStream_t)stream));
  count_launch();
  return 0;
}

extern "C" int dedloc_nvls_params_size(void) { return (int)sizeof(DedlocNvlsParams); }
