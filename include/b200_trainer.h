/* b200_trainer.h -- C ABI of libb200trainer.so (hand-written sm_100a kernels + host runtime).
 *
 * Conventions
 *   - every function returns 0 on success, a negative value on failure; b200_last_error() then
 *     returns a thread-local message.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - pointers are device pointers unless stated otherwise; the library never allocates on behalf
 *     of a kernel: scratch buffers are always caller-provided (see each entry).
 *   - bf16 tensors are row-major with an explicit leading dimension in ELEMENTS.
 *   - "reference" citations are paths under the upstream project (zhc180/distributed-llm-trainer).
 */
#ifndef B200_TRAINER_H_
#define B200_TRAINER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------------------------------------------------------- runtime ------------------- */
const char* b200_last_error(void);
int b200_version(void);
int b200_sm_count(void);

/* Raw device allocations (zero-filled) that can be shared with peer processes through CUDA IPC.
 * Used for the NVLink communication buffers of the DDP / FSDP engines (reference: the NCCL-internal
 * buffers behind torch DDP/FSDP, ddp_trainer.py:167-172, fsdp_trainer.py:291-310). */
int b200_device_malloc(void** ptr, size_t bytes);
int b200_device_free(void* ptr);
int b200_ipc_export(void* ptr, void* handle64 /* host, 64 bytes out */);
int b200_ipc_open(const void* handle64 /* host */, void** ptr);
int b200_ipc_close(void* ptr);
int b200_enable_peer_access(int peer_device);
/* Stream-ordered copy-engine transfer between mapped device addresses (local or peer). */
int b200_memcpy_async(void* dst, const void* src, size_t bytes, void* stream);

/* ---------------------------------------------------------------- GEMM (tcgen05 / TMEM / TMA) */
enum {
  B200_EPI_BF16 = 0,          /* C(bf16) = A*B                                   */
  B200_EPI_F32 = 1,           /* C(fp32) = A*B                                   */
  B200_EPI_F32_ATOMIC = 2,    /* C(fp32) += A*B (red.global.add; split-K allowed) */
  B200_EPI_ROPE_BF16 = 3,     /* C(bf16) = rotary(A*B) on columns [0, rope_cols)  */
  B200_EPI_RESID_DROPOUT = 4, /* C(fp32) = residual(fp32) + dropout(A*B)          */
  B200_EPI_SWIGLU = 5,        /* C = [gate|up] (bf16), aux = silu(gate)*up (bf16) */
  B200_EPI_SWIGLU_BWD = 6     /* C = [dgate|dup] (bf16) from A*B = d_act, aux = [gate|up] */
};

typedef struct b200_gemm_args {
  const void* A;  /* bf16. K-major: [M, K] (ld = lda);  MN-major: [K, M] (ld = lda) */
  const void* B;  /* bf16. K-major: [N, K] (ld = ldb);  MN-major: [K, N] (ld = ldb) */
  void* C;        /* [M, N] (ld = ldc), dtype per epilogue */
  int M, N, K;
  int lda, ldb, ldc;
  int a_mn_major, b_mn_major;
  int epilogue;   /* B200_EPI_* */
  int split_k;    /* <=1: none.  >1 requires B200_EPI_F32_ATOMIC */
  int tile_n;     /* 128, 256, 384 (pair kernel, N % 384 == 0, plain / residual epilogues), or 0 = choose by wave efficiency */
  int max_ctas;   /* 0 = all SMs; otherwise caps the persistent grid (leaves SMs for comm kernels) */
  /* rotary epilogue: fp32 tables [seq_len, 32] (head_dim 64, half-split pairing; reference gpt.py:116-147) */
  const float* rope_cos;
  const float* rope_sin;
  int seq_len;
  int rope_cols;
  /* residual + dropout epilogue (reference gpt.py:240,282,308,314) */
  const float* residual;
  float dropout_p;
  const unsigned long long* rng_state; /* device {seed, offset}; may be NULL when dropout_p == 0 */
  unsigned int rng_stream;
  /* SwiGLU epilogues (reference gpt.py:278-281) */
  void* aux;
  int ld_aux;
  int half_n;
  /* pair kernel tile order: 0 = automatic, 1 = m fastest (neighbouring tiles share B), 2 = n fastest (the n-tiles
     of a row block run together: A streams from HBM once when it does not fit in L2) */
  int raster;
} b200_gemm_args;

/* Reference ops G3/G6/G8/G9 (gpt.py:185-187,239,278-281,447) and their backward GEMMs. */
int b200_gemm_bf16(const b200_gemm_args* args, void* stream);

/* CTA-pair variant: tcgen05.mma.cta_group::2 over 256x256 pair tiles (each CTA stages its A rows and half of
 * B), epilogue through swizzled smem + TMA store / TMA reduce-add.  Epilogues B200_EPI_BF16, _F32,
 * _F32_ATOMIC; C row pitch must be a multiple of 16 bytes.  Same reference ops as b200_gemm_bf16. */
int b200_gemm2_bf16(const b200_gemm_args* args, void* stream);

/* All-gather fused into the consuming GEMM (FSDP: the QKV projection is the first consumer of a block's
 * weights; reference: FSDP1 all_gather_into_tensor before the block forward, fsdp_trainer.py:379).  B is a
 * weight whose rows are sharded across ranks: inside ONE launch, `gather_pairs` CTA pairs stream the 128-row
 * blocks of B from their owners' shards (peer memory over NVLink) into args->B with bulk async copies and
 * publish a per-block flag, while the remaining pairs run the tcgen05 GEMM, their TMA producers waiting on
 * the flag of the rows they stage; tiles over locally owned rows are visited first, so the transfer overlaps
 * the math tile by tile.  Epilogues: B200_EPI_BF16, B200_EPI_ROPE_BF16.  K-major operands, ldb == K. */
typedef struct b200_ag_args {
  const void* const* shard_ptrs; /* host array [world]: every rank's bf16 shard of the unit (device pointers) */
  long long part_elems;          /* elements of the unit owned by each rank (rank r owns [r*part, (r+1)*part)) */
  long long b_offset_elems;      /* offset of B inside the unit */
  unsigned int* flags;           /* device uint32[ceil(N/128)]; must not already hold `epoch` */
  unsigned int epoch;            /* value published per block: use a new one for every call that reuses flags */
  int rank, world;
  int gather_pairs;              /* 0 = default (4 pairs = 8 SMs) */
  const unsigned int* epoch_dev; /* optional: device word holding the epoch (overrides `epoch`); bump it on the same
                                    stream before every launch -- what makes the launch replayable from a CUDA graph */
} b200_ag_args;
int b200_gemm2_ag_bf16(const b200_gemm_args* args, const b200_ag_args* ag, void* stream);

/* Reduce-scatter fused into the producing GEMM (the gradient path of DDP and FSDP; reference: DDP's bucketed
 * all-reduce fired from backward, ddp_trainer.py:329-342, and FSDP's reduce_scatter_tensor, fsdp_trainer.py:383).
 * C = A*B is a weight gradient (wgrad form: MN-major A and B, B200_EPI_F32_ATOMIC) whose rows are owned by
 * ranks in 128-row slabs: slab s of `slabs = ceil(M/128)` belongs to rank ((s*world)/slabs + rot) % world, so
 * every rank owns one contiguous run of slabs.  Each CTA's 128-row output slab is reduce-added by TMA
 * (cp.reduce.async.bulk.tensor .add) straight into its OWNER's fp32 buffer -- peer memory over NVLink unless
 * the owner is the calling rank -- from the same swizzled staging ring the local epilogue uses, so the
 * transfer rides under the main loop of the following tiles.  After every rank has run its GEMM (and a
 * barrier), owner o holds sum_r C_r for its rows: there is no separate reduce-scatter kernel.
 *   owner_base[o]: device pointer (mapped into this process) to the first row rank o owns, row pitch args->ldc;
 *                  NULL allowed for ranks that own no slab.  args->C is ignored (may be NULL).
 * Gradient accumulation: micro-steps before the last accumulate locally (b200_gemm2_bf16, reduce-add into a
 * private buffer); the last one passes that buffer as `prev`. */
typedef struct b200_rs_args {
  const void* const* owner_base; /* host array [world] */
  int world;                     /* 1..8 */
  int rot;                       /* rotation of the owner assignment (spreads short tensors over the ranks) */
  const void* prev;              /* optional: this rank's own fp32 accumulator of C ([M, N], pitch args->ldc) holding
                                    the gradient of the earlier micro-steps; when set, owner += A*B + prev, so a
                                    step's whole local gradient crosses NVLink once, inside the last wgrad GEMM */
} b200_rs_args;
int b200_gemm2_rs_bf16(const b200_gemm_args* args, const b200_rs_args* rs, void* stream);
/* The ownership rule above, for callers that lay out the owners' buffers: owner of a slab, and the slab run
 * [first, last) of an owner (empty when first >= last).  Host-only helpers, no GPU work. */
int b200_slab_owner(int slab, int slabs, int world, int rot);
int b200_owner_slab_range(int owner, int slabs, int world, int rot, int* first, int* last);

/* ---------------------------------------------------------------- normalisation / elementwise */
/* Token embedding gather into the fp32 residual stream (reference gpt.py:438). ids: int64 [rows]. */
int b200_embed_fwd(const int64_t* ids, const void* weight_bf16, float* out, int rows, int hidden, int vocab,
                   void* stream);
/* Scatter-add of the residual gradient into the (tied) embedding gradient (fp32 [vocab, hidden]). */
int b200_embed_bwd(const int64_t* ids, const float* dx, float* dweight, int rows, int hidden, int vocab,
                   void* stream);

/* RMSNorm, one pass: y(bf16) = x * rsqrt(mean(x^2)+eps) * w ; rstd saved (reference gpt.py:66-67). */
int b200_rmsnorm_fwd(const float* x, const void* w_bf16, void* y_bf16, float* rstd, int rows, int hidden, float eps,
                     void* stream);
/* dx (fp32) += d/dx ; dw (fp32 [hidden]) += d/dw (atomics).  dy is bf16. */
int b200_rmsnorm_bwd(const void* dy_bf16, const float* x, const void* w_bf16, const float* rstd, float* dx_accum,
                     float* dw_accum, int rows, int hidden, void* stream);
/* Same, and additionally dy_next(bf16) = dropout_mask(rng_stream) * dx_accum_updated / (1-p): the
   b200_dropout_bwd_cast that follows a norm backward in the block's backward pass, without its second
   pass over dx.  dy_next may be NULL (then identical to b200_rmsnorm_bwd). */
int b200_rmsnorm_bwd_dropout(const void* dy_bf16, const float* x, const void* w_bf16, const float* rstd,
                             float* dx_accum, float* dw_accum, int rows, int hidden, void* dy_next_bf16, float p,
                             const unsigned long long* rng_state, unsigned int rng_stream, void* stream);

/* out(fp32) = residual + dropout(y bf16)   -- unfused companion of B200_EPI_RESID_DROPOUT */
int b200_resid_dropout_fwd(const void* y_bf16, const float* residual, float* out, int rows, int cols, float p,
                           const unsigned long long* rng_state, unsigned int rng_stream, void* stream);
/* dy(bf16) = dropout_mask * dx / (1-p)     -- gradient fed to the o_proj / down_proj backward GEMMs */
int b200_dropout_bwd_cast(const float* dx, void* dy_bf16, int rows, int cols, float p,
                          const unsigned long long* rng_state, unsigned int rng_stream, void* stream);
/* keep-mask (uint8) of a [rows, cols] dropout op, for tests / oracles */
int b200_dropout_mask(uint8_t* mask, int rows, int cols, float p, const unsigned long long* rng_state,
                      unsigned int rng_stream, void* stream);

/* dst(bf16)[i] = src(fp32)[i]; used after split-K fp32 accumulation (reference: autocast's _to_copy casts, G11) */
int b200_cast_f32_bf16(const float* src, void* dst_bf16, long long n, void* stream);

/* In-place rotary embedding on the q and k column ranges of a packed [rows, 3h] bf16 tensor.
 * inverse != 0 applies the transposed rotation (backward).  Unfused companion of B200_EPI_ROPE_BF16. */
int b200_rope_inplace(void* qkv_bf16, const float* cos, const float* sin, int rows, int ld, int rope_cols,
                      int seq_len, int inverse, void* stream);

/* SwiGLU: act = silu(g)*u from gu=[g|u] ([rows, 2I]);  backward: dgu from d_act and gu. */
int b200_swiglu_fwd(const void* gu_bf16, void* act_bf16, int rows, int inter, void* stream);
int b200_swiglu_bwd(const void* dact_bf16, const void* gu_bf16, void* dgu_bf16, int rows, int inter, void* stream);

/* ---------------------------------------------------------------- attention ----------------- */
/* Causal flash attention over packed qkv [B*S, 3h] (bf16, q|k|v column blocks, head_dim 64) with
 * Philox dropout on the probabilities; never materialises [S,S] (reference gpt.py:199-234).
 * out: [B*S, h] bf16 in [B,S,H,D] order (what o_proj consumes, so the reference's
 * transpose+contiguous copy at gpt.py:238 disappears); lse: [B, H, S] fp32. */
int b200_attn_fwd(const void* qkv_bf16, void* out_bf16, float* lse, int batch, int seq, int heads, float dropout_p,
                  const unsigned long long* rng_state, unsigned int rng_stream, void* stream);
/* keep-mask (uint8, [rows = B*H*S, cols = S]) of the attention dropout for (p, rng_state, rng_stream), exactly
 * as b200_attn_fwd / b200_attn_bwd draw it: one Philox call per 16 consecutive kv positions of a query row, 8-bit
 * lanes, dropped when lane < round(p * 256) (p = 0.1 -> 26/256; the kernels rescale by the realised keep
 * probability).  For tests / oracles. */
int b200_attn_dropout_mask(uint8_t* mask, int rows, int cols, float p, const unsigned long long* rng_state,
                           unsigned int rng_stream, void* stream);
/* dqkv: [B*S, 3h] bf16.  delta: caller-provided scratch [B, H, S] fp32; dq_accum: caller-provided fp32
 * scratch [B*S, h], zeroed by the call. */
int b200_attn_bwd(const void* qkv_bf16, const void* out_bf16, const void* dout_bf16, const float* lse,
                  void* dqkv_bf16, float* delta, float* dq_accum, int batch, int seq, int heads, float dropout_p,
                  const unsigned long long* rng_state, unsigned int rng_stream, void* stream);
/* Same, for q / k that went through the rotary embedding (B200_EPI_ROPE_BF16 in the QKV GEMM): the dQ and dK
 * columns of dqkv come out already rotated back (b200_rope_inplace(..., inverse=1) folded into the final
 * pass, the dQ rotation done in fp32 before the single rounding), and the delta / zero-fill passes are one. */
int b200_attn_bwd_rope(const void* qkv_bf16, const void* out_bf16, const void* dout_bf16, const float* lse,
                       void* dqkv_bf16, float* delta, float* dq_accum, int batch, int seq, int heads, float dropout_p,
                       const unsigned long long* rng_state, unsigned int rng_stream, const float* rope_cos,
                       const float* rope_sin, void* stream);

/* Attention backward that allocates nothing and asks the caller for no scratch.  It has no counterpart in
 * the reference project; it exists for hosts that cannot hand this library device memory on demand (a
 * framework allocator that already owns all of HBM, a request path where cudaMalloc is forbidden because
 * it stalls the device, a bare C caller).  The row-sum and dQ-accumulator scratch that b200_attn_bwd takes
 * from its caller lives in a 32 MiB array compiled into the library, which every GPU that loads the
 * library keeps resident until the process exits.
 *   - same inputs, outputs and numerics as b200_attn_bwd;
 *   - refused with an error when the scratch it needs -- batch*seq*heads*64*4 + batch*heads*seq*4 bytes,
 *     each term rounded up to 256 -- exceeds b200_attn_bwd_noalloc_max_scratch_bytes() (32 MiB);
 *   - the first call on a device resolves the array's address and configures the kernel: make it at
 *     start-up, before any stream capture;
 *   - every call on a device reuses the same array, which still holds the previous call's data; what must
 *     be zero is cleared on `stream` at the start of each call.  Calls whose GPU work may overlap (other
 *     streams, other threads) race on it -- serialise them. */
unsigned long long b200_attn_bwd_noalloc_max_scratch_bytes(void);
int b200_attn_bwd_noalloc(const void* qkv_bf16, const void* out_bf16, const void* dout_bf16, const float* lse,
                          void* dqkv_bf16, int batch, int seq, int heads, float dropout_p,
                          const unsigned long long* rng_state, unsigned int rng_stream, void* stream);

/* ---------------------------------------------------------------- LM head loss -------------- */
/* Fused softmax cross-entropy forward+backward over a chunk of logits (reference gpt.py:449-453).
 * logits: bf16 [rows, ld] (ld >= vocab, padded columns ignored); labels: int64 [rows], <0 = ignore.
 * In place: logits <- (softmax - onehot) * grad_scale (0 for ignored rows / padded columns).
 * loss_sum (fp32 device scalar) += sum over rows of (lse - logit[label]) * loss_scale. */
int b200_ce_fwd_bwd(void* logits_bf16, const int64_t* labels, float* loss_sum, int rows, int vocab, int ld,
                    float loss_scale, float grad_scale, void* stream);

/* ---------------------------------------------------------------- optimizer ----------------- */
/* Fused AdamW over flat element range [lo, hi) (reference ddp_trainer.py:229-234, fsdp_trainer.py:338-343).
 * master/grad are indexed by the GLOBAL flat offset; exp_avg/exp_avg_sq by (offset - lo).
 * decay_chunks: uint8 per 128-element chunk (global chunk index).  grad_scale: device fp32 scalar
 * multiplied into the gradient (clip coefficient).  Writes the bf16 shadow copy when non-NULL. */
int b200_adamw(float* master, const float* grad, float* exp_avg, float* exp_avg_sq, void* shadow_bf16,
               const uint8_t* decay_chunks, long long lo, long long hi, int step, float lr, float beta1, float beta2,
               float eps, float weight_decay, const float* grad_scale, void* stream);
/* out[0] += sum(x[lo:hi]^2)  (grad-norm partial; reference ddp_trainer.py:350) */
int b200_sumsq(const float* x, long long lo, long long hi, float* out, void* stream);
/* scale[0] = pre_scale * min(1, max_norm / (sqrt(sumsq[0]) + 1e-6));  max_norm <= 0 disables clipping */
int b200_clip_coef(const float* sumsq, float max_norm, float pre_scale, float* scale, float* norm_out, void* stream);
/* rng_state[1] += delta  (device-side offset bump: keeps dropout fresh under CUDA-graph replay) */
int b200_rng_advance(unsigned long long* rng_state, unsigned long long delta, void* stream);

/* ---------------------------------------------------------------- NVLink communication -------- */
/* All of these take HOST arrays of `world` device pointers (one per rank, index = rank), each pointing
 * at that rank's copy of a symmetric buffer mapped into this process (b200_ipc_open).  `pad_ptrs` are
 * the signal pads (b200_comm_pad_bytes() bytes each, zero-initialised).  `epoch_counter` is a zero-initialised
 * uint32 in THIS rank's device memory that the pad kernels bump themselves (so launches carry no changing
 * host argument and can be captured in CUDA graphs); every rank must issue the same sequence of pad
 * operations and never have two of them in flight on different streams.  Kernels that wait on peers are
 * bounded by a 20 s device-side timeout and trap instead of hanging. */
int b200_comm_pad_bytes(void);
int b200_comm_barrier(const void* const* pad_ptrs, int rank, int world, unsigned int* epoch_counter, void* stream);
/* dst[lo:hi] = scale * sum_r grads_r[lo:hi]; sumsq[0] += |dst[lo:hi]|^2  (reference: DDP bucket all-reduce
 * ddp_trainer.py:342 + clip_grad_norm_ :350, FSDP reduce_scatter fsdp_trainer.py:383) */
int b200_reduce_scatter_f32(const void* const* grad_ptrs, float* dst, long long lo, long long hi, float scale,
                            float* sumsq, int rank, int world, int max_ctas, void* stream);
/* AdamW on the owned range, bf16 result stored into every rank's shadow buffer (reference: all-reduce
 * second half + optimizer.step, ddp_trainer.py:342,356) */
int b200_adamw_bcast(float* master, const float* grad, float* exp_avg, float* exp_avg_sq,
                     const void* const* shadow_ptrs, int rank, int world, const uint8_t* decay_chunks, long long lo,
                     long long hi, int step, float lr, float beta1, float beta2, float eps, float weight_decay,
                     const float* grad_scale, int max_ctas, void* stream);
/* out[i] = sum_r in_r[i], i < n <= 8 (reference: FSDP.clip_grad_norm_ scalar all-reduce, fsdp_trainer.py:388) */
int b200_allreduce_scalars(const void* const* pad_ptrs, int rank, int world, unsigned int* epoch_counter,
                           const float* in, float* out, int n, void* stream);
/* dst[r*shard + i] = shard_r[i]  (reference: FSDP all_gather_into_tensor, fsdp_trainer.py:379,383) */
int b200_all_gather_bf16(const void* const* shard_ptrs, void* dst, long long shard_elems, int rank, int world,
                         int max_ctas, void* stream);
/* Same, leaving elements [excl_lo, excl_hi) of dst untouched: the window a fused all-gather GEMM fills itself. */
int b200_all_gather_bf16_excl(const void* const* shard_ptrs, void* dst, long long shard_elems, int rank, int world,
                              long long excl_lo_elems, long long excl_hi_elems, int max_ctas, void* stream);
/* dst_shard[i] += scale * sum_r full_grad_r[src_offset + i]  (reference: FSDP reduce_scatter_tensor,
 * fsdp_trainer.py:383).  fp32 wire (exact) and bf16 wire (the reference's reduce_dtype) variants. */
int b200_reduce_scatter_f32_accum(const void* const* full_grad_ptrs, float* dst_shard, long long src_offset,
                                  long long n, float scale, int rank, int world, int max_ctas, void* stream);
int b200_reduce_scatter_bf16_accum(const void* const* full_grad_ptrs, float* dst_shard, long long src_offset,
                                   long long n, float scale, int rank, int world, int max_ctas, void* stream);


/* ---- optimizer side of the fused reduce-scatter (b200_gemm2_rs_bf16): chunk tables ---------------------------
 * A rank's owned parameters are many short runs (128-row slabs of weight matrices, whole norm vectors).  The
 * caller flattens them into chunks of <= 4096 elements (n % 4 == 0) in a device array so that ONE launch covers
 * them all.  Offsets are in elements:
 *   g   global flat index of the first element (weight-decay table / source offset for pulls)
 *   pg  offset into the local fp32 master / gradient arrays      mv  offset into the local moment arrays
 *   sh  offset into the destination bf16 parameter buffer(s)     dst receiving rank, or -1 = every rank */
typedef struct b200_chunk {
  long long g, pg, mv, sh;
  int n, dst;
} b200_chunk;
/* out[0] += sq_scale * sum(grad[chunks]^2)   (reference: clip_grad_norm_, ddp_trainer.py:350, fsdp_trainer.py:388) */
int b200_sumsq_chunks(const float* grad, const b200_chunk* chunks_dev, int n_chunks, float* out, float sq_scale,
                      void* stream);
/* AdamW over the chunks; updated bf16 parameters are stored into rank dst's buffer (param_ptrs[dst] + sh) or into
 * every rank's (dst = -1) -- with param_multicast != NULL (the NVLS multicast address of the same symmetric buffer)
 * by ONE multimem.st that the NVSwitch replicates, otherwise by world stores; zero_grad != 0 clears the gradient in
 * the same pass (reference: optimizer.step +
 * zero_grad + the parameter half of the all-reduce / next all-gather, ddp_trainer.py:356, fsdp_trainer.py:391). */
int b200_adamw_chunks(float* master, float* grad, float* exp_avg, float* exp_avg_sq, const void* const* param_ptrs,
                      void* param_multicast, int rank, int world, const uint8_t* decay_chunks,
                      const b200_chunk* chunks_dev, int n_chunks,
                      int step, float lr, float beta1, float beta2, float eps, float weight_decay,
                      const float* grad_scale, int zero_grad, int max_ctas, void* stream);
/* The parameter half of b200_adamw_chunks alone: bf16(master[pg ..]) stored to the chunks' holders, and
 * grad_to_zero[pg ..] (may be NULL) cleared -- used when AdamW itself ran on the host (reference: FSDP
 * CPUOffload(offload_params=True), fsdp_trainer.py:300-301). */
int b200_cast_scatter_chunks(const float* master, float* grad_to_zero, const void* const* param_ptrs, int rank, int world,
                             const b200_chunk* chunks_dev, int n_chunks, void* stream);
/* dst[pg ..] = sum_r src_r[g ..] over the chunks (small tensors that every rank accumulates locally);
 * zero_remote != 0 also clears every rank's source copy for the next step (dst may alias this rank's source).
 * src_multicast != NULL (NVLS multicast address of the source buffer): the sum is one multimem.ld_reduce done inside
 * the NVSwitch and the clearing one multimem.st, instead of world loads / stores per element. */
int b200_pull_reduce_chunks(const void* const* src_ptrs, void* src_multicast, float* dst, const b200_chunk* chunks_dev,
                            int n_chunks, int rank, int world, int zero_remote, void* stream);
/* b200_embed_bwd with owner-directed rows (same slab rule as b200_gemm2_rs_bf16 over `vocab` rows): owner_base[o]
 * points at the first embedding-gradient row rank o owns. */
int b200_embed_bwd_rs(const int64_t* ids, const float* dx, const void* const* owner_base, int rows, int hidden,
                      int vocab, int world, int rot, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_TRAINER_H_ */
