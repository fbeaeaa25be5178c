/* dedloc_b200 -- C ABI of the native sm_100a library (libdedloc_b200.so).
 *
 * Every entry point returns 0 on success, a positive cudaError_t, or a negative DEDLOC_E* code;
 * dedloc_last_error() gives the message for the calling thread. `stream` is a cudaStream_t.
 * Python binds these with ctypes (dedloc_b200/lib/__init__.py); nothing here depends on PyTorch.
 * Citations name the part of the reference (/root/reference) whose behaviour the
 * entry point provides.
 */
#ifndef DEDLOC_B200_H_
#define DEDLOC_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DEDLOC_MAX_PEERS 16
#define DEDLOC_IPC_HANDLE_BYTES 64

/* ---- library ---------------------------------------------------------------------------------- */
const char* dedloc_last_error(void);
long long dedloc_launch_count(void);          /* kernels launched by this library since load */
void dedloc_add_launch_count(long long n);    /* account for kernels replayed from a captured CUDA graph */
int dedloc_device_info(int* sm_count, int* cc_major, int* cc_minor, long long* total_mem);

/* ---- symmetric peer memory over CUDA IPC (SURVEY 5.8 data plane; replaces hivemind's gRPC tensors) */
int dedloc_symm_alloc(long long bytes, void** ptr, unsigned char handle[DEDLOC_IPC_HANDLE_BYTES]);
int dedloc_symm_open(const unsigned char handle[DEDLOC_IPC_HANDLE_BYTES], void** ptr);
int dedloc_symm_close(void* ptr);
int dedloc_symm_free(void* ptr);
int dedloc_enable_peer_access(int peer_device);
int dedloc_memset_async(void* ptr, int value, long long bytes, void* stream);
int dedloc_memcpy_async(void* dst, const void* src, long long bytes, void* stream);
int dedloc_memcpy_peer_async(void* dst, int dst_device, const void* src, int src_device, long long bytes,
                             void* stream);
int dedloc_memcpy2d_async(void* dst, long long dpitch, const void* src, long long spitch, long long width,
                          long long height, void* stream);

/* ---- flat gradient path (albert/run_trainer.py:279 HF Trainer clip + hivemind grad accumulation) */
int dedloc_sumsq(const void* x, int is_bf16, long long n, float* scratch, int scratch_floats, float* out,
                 void* stream);
int dedloc_accumulate_clip(float* acc, void* grad, int is_bf16, long long n, const float* normsq,
                           float max_norm, int zero_grad, void* stream);
/* wire codecs (hivemind CompressionType, albert/arguments.py:75-77); see csrc/codec.cuh */
int dedloc_encode(int codec, const float* src, long long n, float scale, void* payload, float* aux,
                  void* stream);
long long dedloc_codec_aux_floats(int codec, long long n);
int dedloc_codec_payload_bytes_per_elem(int codec);
int dedloc_chunk_elems(void);

/* ---- fused butterfly all-reduce + optimizer (hivemind averager.step + torch_optimizer.Lamb /
 *      apex LARC+SGD; albert/run_trainer.py:86-94,251-264; sgd_collaborative.py:135-157) ---------- */
enum { DEDLOC_AR_MEAN = 0, DEDLOC_AR_LAMB = 1, DEDLOC_AR_LARS = 2 };
enum { DEDLOC_TF_FROZEN = 1, DEDLOC_TF_NO_LARS = 2, DEDLOC_TF_ROWNORM128 = 4 };

typedef struct DedlocARParams {
  int P;                 /* group size */
  int me;                /* my index in the group */
  int mode;              /* DEDLOC_AR_* */
  int codec;             /* Codec enum of csrc/codec.cuh */
  int phases;            /* bit0 = phase A, bit1 = phase B (3 in production) */
  int xgpu_barriers;     /* 1 in production; 0 for single-GPU emulated ranks */
  uint32_t tag;          /* unique per round (low 2 bits ignored) */
  int T;                 /* number of tensors in the flat layout */
  unsigned long long timeout_ns;
  int replicate_state;   /* push optimizer moments to every member */
  int max_blocks;        /* 0 = as many as are co-resident */
  float lr, beta1, beta2, eps, bias_correction, clamp_value, momentum, trust;
  float w[DEDLOC_MAX_PEERS];                 /* normalised weights, sum = 1 */
  int slot[DEDLOC_MAX_PEERS];                /* flag slot of every member */
  long long cbeg[DEDLOC_MAX_PEERS];          /* owned chunk range of every member */
  long long cend[DEDLOC_MAX_PEERS];
  const uint8_t* send[DEDLOC_MAX_PEERS];     /* encoded payload of every member */
  const float* aux[DEDLOC_MAX_PEERS];
  float* out32[DEDLOC_MAX_PEERS];            /* MEAN: result; LAMB/LARS: fp32 master parameters */
  uint16_t* out16[DEDLOC_MAX_PEERS];         /* optional bf16 mirror */
  float* m[DEDLOC_MAX_PEERS];                /* exp_avg / momentum buffer */
  float* v[DEDLOC_MAX_PEERS];                /* exp_avg_sq */
  float* table[DEDLOC_MAX_PEERS];            /* [DEDLOC_MAX_PEERS][T+1][2] per-tensor partial sums */
  uint32_t* flags[DEDLOC_MAX_PEERS];         /* [4][DEDLOC_MAX_PEERS] signal pad */
  float* gavg;                               /* local scratch: averaged gradient of my shard */
  float* chunk_stats;                        /* local scratch: [my chunks][2] */
  const int* chunk_tid;                      /* tensor id of every chunk */
  const long long* tensor_cbeg;              /* [T+1] first chunk of every tensor */
  const float* t_wd;                         /* per-tensor weight decay */
  const float* t_lr_scale;                   /* per-tensor lr multiplier */
  const int* t_flags;                        /* DEDLOC_TF_* */
  uint32_t* status;                          /* local: [0] ARStatus, [2] non-finite chunks, [5..7] verdicts */
} DedlocARParams;

int dedloc_fused_allreduce(const DedlocARParams* params, void* stream);
int dedloc_ar_params_size(void);

/* ---- the same operation as a push-based, stage-pipelined dataflow kernel (csrc/fused_allreduce_v2.cu) -------------
 * Ownership is interleaved (chunk c belongs to owner c % K), the vector is cut into NS stages, and the CTAs of a
 * launch are split into four roles that run concurrently on different stages:
 *   PUSH    encode my gradient and store every chunk into its owner's staging slot (posted NVLink writes)
 *   REDUCE  owner: decode + weighted sum of all staged payloads; MEAN: store the average to every member;
 *           LAMB/LARS: per-tensor partial norms, published to every member as self-validating {value, tag} words
 *   UPDATE  owner: trust ratio from the summed norms, moments, and the update DIRECTION (fp16 u for LAMB, bf16
 *           momentum for LARS) stored to every member's delta buffer
 *   APPLY   every member: p32 -= scale_t * direction, bf16 mirror, optional row normalisation
 * Stages are handed from role to role and from GPU to GPU through round-tagged flags; there is no grid-wide or
 * cross-GPU barrier. fp32 master parameters stay replicated and bit-identical, moments live with the owner
 * (replicate_state pushes them too). `n_ranks` > 1 runs that many emulated members as one cooperative launch on one
 * GPU (tests). Same reference call sites as dedloc_fused_allreduce. */
#define DEDLOC_AR2_MAX_STAGES 64
typedef struct DedlocAR2Params {
  int P, me, mode, codec, T, K;               /* members, my index, DEDLOC_AR_*, codec, tensors, owners */
  uint32_t tag;                               /* non-zero, unique per round */
  int src_f32;                                /* 1: PUSH encodes acc * scale on the fly (NONE / FLOAT16 / BFLOAT16) */
  unsigned long long timeout_ns;
  int nP, nR, nU, nA;                         /* CTAs per role; their sum is the grid of one member */
  int NS;                                     /* stages */
  int replicate_state;
  long long Ls, C;                            /* chunks per stage (multiple of 32), chunks in the vector */
  long long slot_bytes, aux_slot_floats;      /* staging geometry of one sender slot */
  float scale;
  float lr, beta1, beta2, eps, bias_correction, clamp_value, momentum, trust;
  float w[DEDLOC_MAX_PEERS];
  int owner_of_k[DEDLOC_MAX_PEERS];           /* member index of owner k */
  int kidx[DEDLOC_MAX_PEERS];                 /* owner index of member q, -1 = owns nothing */
  /* local */
  const float* acc;                           /* fp32 gradient accumulators (src_f32) */
  const uint8_t* send;                        /* pre-encoded payload + aux (other codecs) */
  const float* aux;
  float* p32;
  uint16_t* p16;
  float* m;
  float* v;
  float* gavg;
  float* chunk_stats;
  const int* chunk_tid;
  const long long* tensor_cbeg;
  const float* t_wd;
  const float* t_lr_scale;
  const int* t_flags;
  uint32_t* status;                           /* [0] ARStatus [1] state was modified [2] non-finite chunks / skipped tensors */
  uint32_t* counters;                         /* [3][DEDLOC_AR2_MAX_STAGES], zeroed by the entry point */
  /* per member (peer memory) */
  uint8_t* stage_in[DEDLOC_MAX_PEERS];        /* [P slots][slot_bytes] */
  float* aux_in[DEDLOC_MAX_PEERS];            /* [P slots][aux_slot_floats] */
  uint16_t* delta[DEDLOC_MAX_PEERS];          /* [C * chunk] update directions */
  uint32_t* inflag[DEDLOC_MAX_PEERS];         /* [NS][MAX_PEERS] */
  uint32_t* outflag[DEDLOC_MAX_PEERS];        /* [NS][MAX_PEERS] */
  unsigned long long* table[DEDLOC_MAX_PEERS];/* [MAX_PEERS owners][T][2] {value, tag} words */
  uint32_t* abort_word[DEDLOC_MAX_PEERS];
  float* out32[DEDLOC_MAX_PEERS];             /* MEAN outputs */
  uint16_t* out16[DEDLOC_MAX_PEERS];
  float* m_all[DEDLOC_MAX_PEERS];             /* replicate_state targets */
  float* v_all[DEDLOC_MAX_PEERS];
  unsigned long long* trace;                  /* optional [4 roles][MAX_STAGES][4] %globaltimer stamps of CTA 0 of each role */
  int debug_flags;                            /* tests / diagnostics: 1 omit system fences, 2 omit per-warp fences, 4 keep PUSH local
                                                 (results undefined for 1, 2, 4); bits 8.. = start this member late by up to
                                                 (flags >> 8) * 16 us (bit 7: exactly) */
  int sequential;                             /* 1: every CTA runs PUSH, REDUCE, UPDATE, APPLY one after the other (all n* = grid) */
} DedlocAR2Params;

/* params: n_ranks structs in HOST memory; params_dev: device scratch of n_ranks * sizeof(DedlocAR2Params) bytes. */
int dedloc_fused_allreduce_v2(const DedlocAR2Params* params, int n_ranks, void* params_dev, void* stream);
int dedloc_ar2_params_size(void);
int dedloc_ar2_max_ctas(int mode, int codec);   /* co-resident CTAs of the kernel on the current device */

/* ---- NVLS variant of the weighted mean (csrc/nvls_allreduce.cu): in-switch reduction with multimem.ld_reduce,
 *      broadcast of the result with multimem.st, barriers with multimem.red. The windows are CUDA VMM allocations
 *      bound to a multicast object (dedloc_b200/parallel/nvls.py). Same reference call site as dedloc_fused_allreduce. */
typedef struct DedlocNvlsParams {
  void* mc_base;                 /* multicast address of the window (all members at once) */
  void* uc_base;                 /* ordinary address of MY window */
  long long n;                   /* elements, multiple of the chunk size */
  long long off_flags, off_payload, off_result;   /* byte offsets inside the window */
  int P, me;
  uint32_t round;                /* 1, 2, 3, ...: the barrier counters only grow */
  int ctas;                      /* grid size; must be the same on every member */
  int payload_is_bf16;           /* else fp16 */
  int reserved;
  unsigned long long timeout_ns;
  uint32_t* status;
  uint16_t* out16;               /* optional: bf16 copy of the result */
  float* out32;                  /* optional: fp32 copy of the result */
} DedlocNvlsParams;
int dedloc_nvls_allreduce_mean(const DedlocNvlsParams* params, void* stream);
int dedloc_nvls_params_size(void);

/* ---- tcgen05/TMEM/TMA GEMM with fused epilogues (cuBLAS calls under HF AlbertForPreTraining built at
 *      albert/run_trainer.py:64-68 and the VISSL swav_head, swav_prototypes_head.py:64-110) ----------- */
enum {
  DEDLOC_EPI_BIAS = 1,       /* + bias[n] */
  DEDLOC_EPI_BIAS_F32 = 2,   /* bias is fp32 (else same 16-bit type as A/B... bf16) */
  DEDLOC_EPI_GELU = 4,       /* gelu_new(x) */
  DEDLOC_EPI_DGELU = 8,      /* x * gelu_new'(aux_in[m,n]) */
  DEDLOC_EPI_RESIDUAL = 16,  /* + residual[m,n] (bf16) */
  DEDLOC_EPI_SAVE_PRE = 32,  /* also store the pre-activation (after bias) to pre_out as bf16 */
  DEDLOC_EPI_SAVE_DGELU = 64,/* with GELU + SAVE_PRE: pre_out receives gelu_new'(pre) instead (shares the tanh) */
  DEDLOC_EPI_MUL_AUX = 128   /* x * aux_in[m,n] (bf16): backward of an activation whose derivative was saved */
};

typedef struct DedlocGemmArgs {
  const void* A;      /* K-major: [M,K] row-major; MN-major: [K,M] row-major */
  const void* B;      /* K-major: [N,K] row-major; MN-major: [K,N] row-major */
  void* C;            /* [M,N] row-major */
  int M, N, K;
  long long lda, ldb, ldc;   /* row strides in elements */
  int a_mn_major, b_mn_major;
  int is_fp16;        /* inputs fp16 instead of bf16 */
  int epilogue;       /* DEDLOC_EPI_* */
  int out_kind;       /* 0 = bf16 store, 1 = fp32 store, 2 = fp32 atomic add (C += A B^T) */
  float alpha;
  const void* bias;
  const void* residual;
  long long ld_residual;
  const void* aux_in;
  long long ld_aux;
  void* pre_out;
  long long ld_pre;
  int split_k;        /* >1 only with out_kind == 2 */
  int block_n;        /* 0 = auto, 128 or 256 */
  int max_ctas;       /* 0 = one per SM */
  int cta_pair;       /* 0 = auto, 1 = one CTA per 128x256 tile, 2 = CTA pair (cta_group::2) per 256x256 tile */
} DedlocGemmArgs;

int dedloc_gemm(const DedlocGemmArgs* args, void* stream);

/* ---- row-wise kernels around the GEMMs (torch LayerNorm / bias-gradient reductions under HF ALBERT) ----- */
int dedloc_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd,
                         int M, int H, float eps, void* stream);
/* dgamma / dbeta are fp32 and ACCUMULATED (weight sharing: 24 uses per step) */
int dedloc_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                         void* dx, float* dgamma, float* dbeta, int M, int H, void* stream);
int dedloc_colsum_accumulate(const void* dy, float* out, int M, int N, long long ld, void* stream);

/* ---- ALBERT embedding front end and the 128-wide LayerNorm of the MLM head (HF AlbertEmbeddings / AlbertMLMHead under
 *      AlbertForPreTraining, albert/run_trainer.py:64-68) ------------------------------------------------------------ */
/* y[t] = LayerNorm(word[ids[t]] + type[token_types[t]] + pos[t % S]); tables, gamma, beta, y: bf16, E = 128 */
int dedloc_embed_ln_fwd(const long long* ids, const long long* token_types, const void* word, const void* pos,
                        const void* type, const void* gamma, const void* beta, void* y, float* mean, float* rstd, int M,
                        int S, int E, float eps, void* stream);
/* LayerNorm backward + the three embedding backwards; every gradient is ACCUMULATED into an fp32 buffer */
int dedloc_embed_ln_bwd(const void* dy, const long long* ids, const long long* token_types, const void* word,
                        const void* pos, const void* type, const void* gamma, const float* mean, const float* rstd,
                        float* gword, float* gpos, float* gtype, float* dgamma, float* dbeta, int M, int S, int E,
                        long long pad_id, int n_types, void* stream);
int dedloc_ln128_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd, int M,
                     float eps, void* stream);
int dedloc_ln128_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd, void* dx,
                     float* dgamma, float* dbeta, int M, void* stream);

/* ---- SwAV Sinkhorn-Knopp assignments in one persistent kernel (vissl/losses/swav_loss.py:177-244) -------- */
/* scores: [n, K] bf16 or fp32 (rows = samples: batch first, then the queue); out: [n_out, K] fp32 assignments of
 * the first n_out rows; scratch: dedloc_sinkhorn_scratch_floats(K, iters) floats, clobbered. */
long long dedloc_sinkhorn_scratch_floats(int K, int iters);
int dedloc_sinkhorn(const void* scores, int is_bf16, long long ld, int n, int K, int n_out, float epsilon, int iters,
                    float world_size, float* out, long long ld_out, float* scratch, void* stream);
/* Same computation for callers that cannot allocate device memory (no counterpart in the reference): the scratch is
 * a built-in per-GPU device array of dedloc_sinkhorn_noalloc_capacity_floats() floats (32 800 = 128.1 KiB), so
 * 32 + (iters + 1) * K must not exceed it -- larger requests fail with an error. The first call on a device does
 * one-time set-up and should precede any stream capture; calls whose GPU work overlaps race on the shared array. */
int dedloc_sinkhorn_noalloc_capacity_floats(void);
int dedloc_sinkhorn_noalloc(const void* scores, int is_bf16, long long ld, int n, int K, int n_out, float epsilon,
                            int iters, float world_size, float* out, long long ld_out, void* stream);

/* ---- SwAV head / loss kernels (vissl/models/heads/swav_prototypes_head.py:100-104, vissl/losses/swav_loss.py:275-302) */
/* y = x / max(||x||, eps), rows of 128 bf16 */
int dedloc_l2norm128_fwd(const void* x, void* y, int M, float eps, void* stream);
int dedloc_l2norm128_bwd(const void* x, const void* dy, void* dx, int M, float eps, void* stream);
/* swapped-prediction cross-entropy. scores: [num_crops * bs, K] bf16 or fp32; q: [n_assign, bs, K] fp32 assignments of the
 * crops assign_crop[0..n_assign) (device array); loss_row[r] = sum_{i: assign_crop[i] != crop(r)} sum_k q_i[b,k] *
 * log_softmax(scores[r] / T)[k]; lse[r] is kept for the backward, which writes
 * dscores[r,k] = grad_scale[0] / T * sum_i (q_i[b,k] - softmax[k] * sum_j q_i[b,j]). */
int dedloc_swav_ce_fwd(const void* scores, int is_bf16, long long ld, const float* q, int n_assign, const int* assign_crop,
                       int num_crops, int bs, int K, float temperature, float* loss_row, float* lse, void* stream);
int dedloc_swav_ce_bwd(const void* scores, int is_bf16, long long ld, const float* q, int n_assign, const int* assign_crop,
                       int num_crops, int bs, int K, float temperature, const float* lse, const float* grad_scale,
                       void* dscores, void* stream);

/* ---- shared-layer attention on the packed QKV layout (HF AlbertAttention: bmm + softmax + bmm) ----------- */
/* qkv: [B, S, 3, H, 64] bf16 (output of the fused QKV GEMM); mask: [B, S] bytes (1 = attend) or NULL;
 * out: [B, S, H*64] bf16; lse: [B, H, S] fp32 (log-sum-exp of the scaled scores, for backward) or NULL. */
int dedloc_attention_fwd(const void* qkv, const unsigned char* mask, void* out, float* lse, int B, int S, int H, int D,
                         float scale, void* stream);
/* backward: dqkv [B, S, 3, H, 64] bf16 from out / dout [B, S, H*64] and lse; scratch holds
 * dedloc_attention_bwd_scratch_floats(S) floats (per-CTA fp32 dQ accumulators), contents are clobbered. */
long long dedloc_attention_bwd_scratch_floats(int S);
int dedloc_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, const unsigned char* mask,
                         void* dqkv, float* scratch, int B, int S, int H, int D, float scale, void* stream);

/* ---- masked-LM cross-entropy on bf16 logits (torch cross_entropy(logits.float(), labels, reduction="none")) ---- */
/* logits: [rows, V] bf16, row pitch ld elements (ld % 8 == 0); labels: int64 per row (out of range -> loss 0);
 * loss, lse: fp32 per row. Backward: dlogits = (softmax - onehot(label)) * row_scale[row], bf16, row pitch ldd. */
int dedloc_ce_fwd(const void* logits, long long ld, const long long* labels, int rows, int V, float* loss, float* lse,
                  void* stream);
int dedloc_ce_bwd(const void* logits, long long ld, const long long* labels, const float* lse, const float* row_scale,
                  void* dlogits, long long ldd, int rows, int V, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DEDLOC_B200_H_ */
