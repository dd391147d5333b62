/*
 * ftk.h -- C ABI of the B200 (sm_100a) training hot path of kubernetes-cloud's
 * finetuner-workflow/finetuner: window gather, fused lm_head + cross-entropy
 * forward/backward, data-parallel gradient all-reduce.
 *
 * The reference (a Python script) has no FFI of its own; each entry point
 * below cites the reference code whose work it replaces. Paths are relative
 * to the reference root; FT.py = finetuner-workflow/finetuner/finetuner.py.
 *
 * Conventions
 *   - plain C types only; device pointers unless a parameter says "host";
 *   - every function returns 0 on success or a negative FTK_E* code and never
 *     throws; ftk_last_error() returns a thread-local message for the last
 *     failure on the calling thread;
 *   - the caller owns every buffer; the library never allocates, frees or
 *     retains tensor memory (workspaces are caller-provided, sized by the
 *     *_workspace queries);
 *   - the CUDA stream is always explicit (a cudaStream_t passed as void*);
 *     the library never synchronises it;
 *   - the device is the one current on the calling thread;
 *   - safe to call forward from the Python thread and backward from the
 *     autograd engine thread (no thread-local CUDA state is kept).
 */
#ifndef FTK_H
#define FTK_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FTK_ABI_VERSION 1

#define FTK_OK 0
#define FTK_EINVAL (-1)    /* bad argument (null pointer, bad shape, misalignment) */
#define FTK_ECUDA (-2)     /* a CUDA runtime/driver call failed */
#define FTK_EWORKSPACE (-3) /* workspace too small */
#define FTK_ENCCL (-4)     /* NCCL missing or a NCCL call failed */
#define FTK_EUNSUPPORTED (-5) /* not an sm_100 device / feature unavailable */

/* dtype codes for ftk_lmhead_ce_bwd (dW) and ftk_dp_allreduce */
#define FTK_BF16 0
#define FTK_F32 1

/* ABI version of the loaded library; the Python loader checks it. */
int ftk_version(void);

/* Message for the last error raised on the calling thread ("" if none). */
const char* ftk_last_error(void);

/*
 * Process-wide tuning options. Each is one atomic integer read at launch time; nothing is read from the
 * environment. No counterpart in the reference.
 *   FTK_OPT_CTA_GROUP       2 (default): tcgen05 cta_group::2, a CTA pair shares one 256-row tile; 1: single CTAs
 *   FTK_OPT_GATHER_VARIANT  0 (default): pick by shape and size; 1: the direct (no shared memory) kernel where it
 *                           applies; 2: the shared-memory staged kernel where it applies
 *   FTK_OPT_GATHER_STORES   0 (default): streaming stores for launches of >= 4 M tokens; 1: always; 2: never
 *   FTK_OPT_SM_RESERVE      n in [0, 64]: SMs the persistent lm_head kernels launched from now on leave idle
 *                           (same as ftk_set_sm_reserve)
 */
#define FTK_OPT_CTA_GROUP 0
#define FTK_OPT_GATHER_VARIANT 1
#define FTK_OPT_GATHER_STORES 2
#define FTK_OPT_SM_RESERVE 3
#define FTK_OPT_COUNT_ 4
int ftk_set_option(int32_t key, int32_t value);

/*
 * Launch timeline: per-kernel device times inside a real step, without a profiler attached. No counterpart in
 * the reference (its only timing is the wall clock per optimizer step, FT.py:496-535).
 *   ftk_timeline_enable(1)  clears the timeline and starts recording: from now on every tagged launch is bracketed by
 *                           two CUDA events on its own stream (host cost: two cudaEventRecord per launch)
 *   ftk_timeline_enable(0)  stops recording and clears
 *   ftk_timeline_read       waits for the recorded events, writes (tag, milliseconds) pairs in launch order and
 *                           returns how many (<= max_n); negative on error. start_ms (optional) receives each
 *                           launch's start relative to the first recorded launch, across streams, so gaps and
 *                           overlap (compute stream vs communication stream) can be read off. Do not call it during
 *                           stream capture.
 * Tags name the kernels of this library.
 */
#define FTK_TL_GATHER 1
#define FTK_TL_FWD 2        /* fused lm_head GEMM + online softmax (ftk_lmhead_ce_fwd) */
#define FTK_TL_FWD_KEEP 3   /* the same, also storing bf16 numerators */
#define FTK_TL_FINALIZE 4
#define FTK_TL_DZ 5         /* logit recompute + dZ (ftk_lmhead_ce_bwd) */
#define FTK_TL_CONVERT 6    /* numerators -> dZ in place */
#define FTK_TL_DH 7
#define FTK_TL_DW 8
#define FTK_TL_DBIAS 9
#define FTK_TL_ADAMW 10
#define FTK_TL_SQNORM 11
#define FTK_TL_ALLREDUCE 12 /* the NCCL collective, on the communication stream */
#define FTK_TL_LINEAR_FWD 13
#define FTK_TL_LINEAR_DX 14
#define FTK_TL_ATTN_FWD 15
#define FTK_TL_ATTN_BWD 16
#define FTK_TL_LAYERNORM_FWD 17
#define FTK_TL_LAYERNORM_BWD 18
int ftk_timeline_enable(int32_t on);
int ftk_timeline_read(int32_t* tags, float* ms, float* start_ms, int32_t max_n);

/*
 * Window gather + label/mask construction, one launch per micro-batch.
 * Replaces TokenizedDataset.load (FT.py:662-692: slice, uint16->int64 widen,
 * mask rule), collector (FT.py:1030-1035: three stacks) and the label fill of
 * ModifiedTrainer.compute_loss (FT.py:476-477: labels[~mask] = -100), plus the
 * host->device copy of 17 B/token HF's Trainer does afterwards.
 *
 *   tokens      uint16 [n_ctx * T]   the token file, resident in HBM (FT.py:639-647)
 *   ctx_idx     int64  [B]           which contexts form this batch (FT.py:757-772 ordering)
 *   pad_id      tokenizer.pad_token_id; pad_is_eos = (pad_id == eos_id) (FT.py:650-651)
 *   shift_labels 0: labels[b,t] = mask ? ids[b,t] : -100   (what the reference hands to HF)
 *               1: labels[b,t] = (t+1<T && mask[b,t+1]) ? ids[b,t+1] : -100
 *                  (HF's causal shift, transformers loss_utils.ForCausalLMLoss, folded in:
 *                  the targets the fused loss consumes)
 *   input_ids   int64 [B,T]; labels int64 [B,T]; mask uint8 [B,T] (0/1, torch.bool layout)
 *   n_valid     int32 [B] or NULL: number of labels != -100 per row
 * Every ctx_idx[b] must lie in [0, n_ctx); the index list lives on the device
 * and is not validated here (the host-side dataset class validates it).
 * T % 16 == 0 and 16-byte aligned buffers take the vectorised path; other
 * shapes take a scalar path with the same results.
 */
int ftk_gather_windows(const uint16_t* tokens, int64_t n_ctx, int32_t T, const int64_t* ctx_idx, int32_t B,
                       int32_t pad_id, int32_t pad_is_eos, int32_t shift_labels, int64_t* input_ids, int64_t* labels,
                       uint8_t* mask, int32_t* n_valid, void* stream);

/*
 * Fused lm_head projection + causal-LM cross-entropy, forward.
 * Replaces logits = lm_head(hidden) (HF modeling_gpt_neo.py lm_head, reached
 * from FT.py:479 super().compute_loss -> model(**inputs)) followed by
 * logits.float() + cross_entropy(ignore_index=-100, reduction=mean)
 * (transformers loss_utils.py fixed_cross_entropy / ForCausalLMLoss).
 * The [M,V] logits never exist in HBM.
 *
 *   H        bf16 [M,d] row-major (hidden states at the lm_head input)
 *   W        bf16 [V,d] row-major (lm_head.weight; tied to wte.weight for GPT-2/GPT-Neo)
 *   bias     bf16 [V] or NULL (GPT-J has one)
 *   targets  int64 [M], already shifted; -100 = ignore
 *   lse      fp32 [M] out: log-sum-exp of each row's logits (saved for backward)
 *   loss_sum fp32 [1] out: sum over valid rows of (lse - logit[target])
 *   n_valid  int32 [1] out: number of rows with target != -100
 *            (mean loss = loss_sum / n_valid; NaN when n_valid == 0, as in the reference)
 * Requirements: d % 64 == 0, H/W 16-byte aligned. Every target must be -100 or lie in [0, V): a target >= V (the
 * uint16 token file can hold ids up to 65535) makes loss_sum NaN, where torch's cross_entropy raises a device assert.
 */
size_t ftk_lmhead_ce_fwd_workspace(int64_t M, int32_t d, int32_t V);
int ftk_lmhead_ce_fwd(const void* H, const void* W, const void* bias, const int64_t* targets, int64_t M, int32_t d,
                      int32_t V, float* lse, float* loss_sum, int32_t* n_valid, void* workspace,
                      size_t workspace_bytes, void* stream);

/*
 * Opt-in pair that trades memory for speed. ftk_lmhead_ce_fwd_keep is the forward above, and
 * additionally leaves the softmax numerators exp(z - c) in bf16 in `probs` ([M][round_up(V,64)],
 * ftk_lmhead_ce_probs_bytes) with the per-(256-column tile, row) reference maxima c in `cref`
 * (ftk_lmhead_ce_cref_bytes); c is the row's running maximum, so the exponent is never positive.
 * ftk_lmhead_ce_bwd_kept turns `probs` into dZ in place (one HBM-bound elementwise pass:
 * dZ = probs * exp(c - lse) * scale - onehot * scale) and runs the dH and dW GEMMs: the logit
 * recompute GEMM of ftk_lmhead_ce_bwd (a quarter of the tensor work of a training step) is gone.
 * Cost: the [M][V] bf16 array now lives from forward to backward instead of only inside the
 * backward, which is why this is not the default. `probs` is consumed by the backward.
 * Same results as the recompute path up to one extra bf16 rounding of the numerators.
 * No counterpart in the reference beyond the autograd it replaces (see ftk_lmhead_ce_bwd).
 */
size_t ftk_lmhead_ce_probs_bytes(int64_t M, int32_t V);
size_t ftk_lmhead_ce_cref_bytes(int64_t M, int32_t V);
int ftk_lmhead_ce_fwd_keep(const void* H, const void* W, const void* bias, const int64_t* targets, int64_t M, int32_t d,
                           int32_t V, float* lse, float* loss_sum, int32_t* n_valid, void* probs, float* cref,
                           void* workspace, size_t workspace_bytes, void* stream);
int ftk_lmhead_ce_bwd_kept(const void* H, const void* W, const int64_t* targets, const float* lse, void* probs,
                           const float* cref, int64_t M, int32_t d, int32_t V, float grad_scale,
                           const float* grad_scale_dev, const int32_t* n_valid_dev, void* dH, void* dW, int32_t dw_dtype,
                           int32_t accumulate, float* dbias, void* stream);

/*
 * The same forward for callers that cannot allocate device memory. It has no
 * counterpart in the reference.
 *
 * ftk_lmhead_ce_fwd wants a workspace, and getting one means cudaMalloc or a
 * framework allocator somewhere in the caller. Two kinds of host cannot do
 * that: a plugin living inside a runtime that reserved all of HBM for its own
 * pool at start-up and gives plugins no allocator, and a latency-critical loop
 * that must avoid the device-wide synchronisation cudaMalloc/cudaFree imply.
 * This variant keeps its scratch in a fixed array inside the library's device
 * image (about 8.5 MB per GPU that loads the library, held until exit), so the
 * call itself performs no allocation of any kind; inputs and outputs are still
 * the caller's buffers.
 *
 *   - M must be <= FTK_STATIC_MAX_ROWS; larger M fails with FTK_EWORKSPACE and
 *     a message naming the limit (use ftk_lmhead_ce_fwd for those).
 *   - The first call on each device looks up the array's device address
 *     (cudaGetSymbolAddress, with that device current) and caches it. Make that
 *     first call before starting any stream capture.
 *   - All calls share the one array and it keeps the previous call's contents.
 *     What has to start at zero is cleared on `stream` at the start of every
 *     call. Calls whose kernels can overlap on the GPU (two streams, two
 *     threads) would race on it: serialise them.
 * Results are identical to ftk_lmhead_ce_fwd.
 */
#define FTK_STATIC_MAX_ROWS 16384
int ftk_lmhead_ce_fwd_static(const void* H, const void* W, const void* bias, const int64_t* targets, int64_t M,
                             int32_t d, int32_t V, float* lse, float* loss_sum, int32_t* n_valid, void* stream);

/*
 * Fused backward of the above (autograd of lm_head + cross_entropy in the
 * reference, reached from loss.backward() inside HF Trainer.training_step).
 *   dZ = (softmax(z) - onehot(target)) * scale for valid rows, 0 for ignored,
 *   scale = grad_scale * (grad_scale_dev ? *grad_scale_dev : 1) / (n_valid_dev ? *n_valid_dev : 1)
 *   dH = dZ . W          bf16 [M,d]
 *   dW (+)= dZ^T . H     [V,d], dw_dtype FTK_BF16 or FTK_F32; accumulate != 0 adds into dW
 *                        (tied embeddings, gradient accumulation)
 *   dbias (+)= sum_m dZ  fp32 [V] or NULL
 * The logits are recomputed tile by tile from H, W and lse; dZ passes through
 * the workspace as bf16 in row slabs. slab_rows = 0 picks the default.
 */
size_t ftk_lmhead_ce_bwd_workspace(int64_t M, int32_t d, int32_t V, int64_t slab_rows);
int ftk_lmhead_ce_bwd(const void* H, const void* W, const void* bias, const int64_t* targets, const float* lse,
                      int64_t M, int32_t d, int32_t V, float grad_scale, const float* grad_scale_dev,
                      const int32_t* n_valid_dev, void* dH, void* dW, int32_t dw_dtype, int32_t accumulate,
                      float* dbias, void* workspace, size_t workspace_bytes, void* stream);

/*
 * Forward and backward of lm_head + cross-entropy in ONE call, for training steps (the default path of
 * ops.fused_lmhead_ce when gradients are wanted). Replaces the same reference code as ftk_lmhead_ce_fwd and
 * ftk_lmhead_ce_bwd together (FT.py:479 and the loss.backward() that follows it); the reference has no such
 * combined call.
 *
 * Doing both at once removes the logit recompute of ftk_lmhead_ce_bwd (a quarter of a step's tensor work) without
 * keeping anything of size [tokens x vocab] between forward and backward: the rows are processed in slabs, and for
 * each slab the forward sweep leaves bf16 softmax numerators in the workspace, a finalize pass gives lse, one
 * HBM-bound pass turns the numerators into dZ in place, and the dH and dW GEMMs consume them before the next slab
 * overwrites them. With two or more slabs there are two slab buffers, and the conversion of one slab runs on a
 * library-owned side stream (joined back into `stream` before the call's last kernels) while the next slab's sweep
 * fills the other. A slab is capped at 1 GiB whatever M is (at most 10,496 rows at V = 50,257; workspace <= 2.1 GiB);
 * ftk_lmhead_ce_fused_slab_rows reports the rows per slab the library would pick (slab_rows = 0) or clamps a
 * request.
 *
 *   dZ scale   grad_scale * (grad_scale_dev ? *grad_scale_dev : 1) / n_valid, with n_valid counted from `targets` at
 *              the start of the call. A caller that only learns the upstream gradient later (autograd) passes 1 and
 *              rescales dH / dW afterwards (ftk_scale_unless_one).
 *   dH         bf16 [M,d] out
 *   dW         [V,d] bf16 / fp32, accumulate != 0 adds into it (tied embeddings, gradient accumulation, fp32 main
 *              gradients); dbias fp32 [V] or NULL likewise
 *   dw_done_event   a cudaEvent_t, or NULL. When given, the last slab runs its dW GEMM before its dH GEMM, records
 *              the event on `stream` between the two, and launches that dH GEMM on (SMs - FTK_OPT_SM_RESERVE) SMs:
 *              dW is complete once the event fires, so a data-parallel caller hands the event to
 *              ftk_dp_allreduce_after and the all-reduce of dW runs on the idle SMs while dH is still being
 *              computed. Every other launch of the call uses all SMs.
 *   lse, loss_sum, n_valid   as ftk_lmhead_ce_fwd
 *
 * ftk_scale_unless_one multiplies n elements (FTK_BF16 / FTK_F32, 16-byte aligned, n a multiple of 8 / 4) by the
 * device scalar *factor_dev and does nothing when that scalar is exactly 1; with poison != 0 the elements become
 * NaN instead when the scalar is not 1 (for gradients that were already added into a caller's buffer and cannot be
 * rescaled: a wrong assumption then shows instead of passing silently).
 */
int64_t ftk_lmhead_ce_fused_slab_rows(int64_t M, int32_t d, int32_t V, int64_t slab_rows);
/* The slab planner on its own: rows per slab for a device with `sm_count` SMs (no device needed; host logic tests). */
int64_t ftk_lmhead_ce_fused_plan(int64_t M, int32_t d, int32_t V, int32_t sm_count, int32_t cta_group);
size_t ftk_lmhead_ce_fused_workspace(int64_t M, int32_t d, int32_t V, int64_t slab_rows);
int ftk_lmhead_ce_fused(const void* H, const void* W, const void* bias, const int64_t* targets, int64_t M, int32_t d,
                        int32_t V, float grad_scale, const float* grad_scale_dev, float* lse, float* loss_sum,
                        int32_t* n_valid, void* dH, void* dW, int32_t dw_dtype, int32_t accumulate, float* dbias,
                        void* dw_done_event, void* workspace, size_t workspace_bytes, int64_t slab_rows, void* stream);
int ftk_scale_unless_one(void* buf, size_t n, int32_t dtype, const float* factor_dev, int32_t poison, void* stream);

/*
 * Linear layers of the transformer trunk (the next row after the hot path): the same
 * TMA -> tcgen05.mma -> TMEM kernel template with a bias / GELU epilogue. Replaces torch.nn.Linear inside HF's
 * GPT-Neo / GPT-J blocks (the modules finetuner.py trains through model(**inputs), FT.py:479, with gradient
 * checkpointing on, FT.py:884) and the elementwise passes around their MLP activation.
 *   ftk_linear_fwd   Y[M,N] = act(X[M,K] . W[N,K]^T + bias[N]) + residual[M,N], all bf16; bias and residual may be
 *                    NULL; act FTK_ACT_GELU_NEW also stores gelu_new'(X . W^T + bias) in `dact` (may be NULL): the
 *                    factor the backward needs, computed from the same tanh as the activation itself.
 *                    The residual (HF GPTNeoBlock.forward's `attn_output + residual` and
 *                    `residual + feed_forward_hidden_states`) is added in fp32 before the one rounding to bf16 and
 *                    may alias Y.
 *   ftk_linear_bwd   dX[M,K] = dY[M,N] . W            (NULL to skip), multiplied elementwise by dx_mul (bf16 [M,K],
 *                    the `dact` of the GELU layer that produced X) when given: the gradient leaves the GEMM already
 *                    taken through the activation; plus dx_add[M,K] when given (may alias dX): layers that share an
 *                    input (q_proj, k_proj, v_proj) chain their dX GEMMs instead of leaving two elementwise sums to
 *                    autograd;
 *                    dW[N,K] (+)= dY^T . X           (NULL to skip; FTK_BF16 / FTK_F32, accumulate as elsewhere);
 *                    dbias[N] (+)= column sums of dY  (fp32, NULL to skip)
 * K and N must be multiples of 64; Y, dact, residual, dX, dW, dx_mul and dx_add 32-byte aligned (the epilogues use
 * 32-byte accesses: a thread owns a row, and 16-byte accesses would leave every sector half written per instruction).
 */
#define FTK_ACT_NONE 0
#define FTK_ACT_GELU_NEW 1
int ftk_linear_fwd(const void* X, const void* W, const void* bias, const void* residual, int64_t M, int32_t K, int32_t N,
                   int32_t act, void* Y, void* dact, void* stream);
int ftk_linear_bwd(const void* X, const void* W, const void* dY, const void* dx_mul, const void* dx_add, int64_t M,
                   int32_t K, int32_t N, void* dX, void* dW, int32_t dw_dtype, int32_t accumulate, float* dbias,
                   void* stream);

/*
 * LayerNorm of the transformer trunk. Replaces torch.nn.LayerNorm in HF's GPT-Neo blocks
 * (GPTNeoBlock.ln_1 / ln_2 and GPTNeoModel.ln_f, reached from FT.py:479 model(**inputs) and re-run by gradient
 * checkpointing, FT.py:884), its autograd, and the elementwise add that joins the residual branch's gradient to the
 * LayerNorm's where the two fork. Arithmetic of torch.nn.functional.layer_norm on bf16 tensors: fp32 statistics
 * (biased variance), rstd = 1 / sqrt(var + eps), one rounding at the end.
 *   X, Y, dY, dX, dx_add   bf16 [M, D];   gamma, beta   bf16 [D] (beta may be NULL);   mean, rstd   fp32 [M]
 *   ftk_layernorm_fwd   Y = (X - mean) * rstd * gamma + beta; also writes mean and rstd for the backward
 *   ftk_layernorm_bwd   dX = the LayerNorm's input gradient (+ dx_add when given: the gradient that reached X along
 *                       the residual branch; may alias dX); dgamma, dbeta fp32 [D] (dbeta may be NULL), summed over
 *                       rows in a fixed order (deterministic, no atomics); workspace from
 *                       ftk_layernorm_bwd_workspace (per-block partial column sums)
 * Requirements: D a multiple of 256, at most 8192; 16-byte aligned buffers.
 */
int ftk_layernorm_fwd(const void* X, const void* gamma, const void* beta, int64_t M, int32_t D, float eps, void* Y,
                      float* mean, float* rstd, void* stream);
size_t ftk_layernorm_bwd_workspace(int64_t M, int32_t D);
int ftk_layernorm_bwd(const void* dY, const void* X, const void* gamma, const float* mean, const float* rstd,
                      const void* dx_add, int64_t M, int32_t D, void* dX, float* dgamma, float* dbeta, void* workspace,
                      size_t workspace_bytes, void* stream);

/*
 * Causal / local-window self-attention of the transformer trunk. Replaces HF GPT-Neo's eager
 * attention (GPTNeoSelfAttention._attn: fp32 Q.K^T, where(causal or sliding-window bias), + attention_mask, softmax,
 * .V, reached from FT.py:479 model(**inputs) and re-run by gradient checkpointing, FT.py:884) and its autograd.
 * The [T, T] scores exist only in TMEM.
 *   Q, K, V, O, dO, dQ, dK, dV   bf16 [B * T, H * head_dim]: the projections' own outputs, heads side by side in a row
 *   key_mask   uint8 [B * T] (non-zero = this position may be attended) or NULL; 16-byte aligned
 *   window     0: causal (key j visible from query i iff j <= i); w > 0: also j > i - w (GPT-Neo's local layers, 256)
 *   scale      multiplies Q.K^T (1 for GPT-Neo, 1/sqrt(head_dim) for GPT-2 / GPT-J)
 *   lse        fp32 [B * H * T]: natural-log sum-exp of each row's visible scores (+inf for a row that sees no key,
 *              whose output is 0)
 *   workspace  ftk_attention_bwd_workspace bytes, 16-byte aligned (per query and head: lse in log2 units, row sum of dO * O)
 * Requirements: T % 128 == 0, head_dim 64 or 128, 16-byte aligned inputs, 32-byte aligned O / dQ / dK / dV.
 * Deterministic (no atomics).
 */
int ftk_attention_fwd(const void* Q, const void* K, const void* V, const uint8_t* key_mask, int32_t B, int32_t T, int32_t H,
                      int32_t head_dim, int32_t window, float scale, void* O, float* lse, void* stream);
size_t ftk_attention_bwd_workspace(int32_t B, int32_t T, int32_t H);
int ftk_attention_bwd(const void* Q, const void* K, const void* V, const void* O, const void* dO, const float* lse,
                      const uint8_t* key_mask, int32_t B, int32_t T, int32_t H, int32_t head_dim, int32_t window,
                      float scale, void* dQ, void* dK, void* dV, void* workspace, size_t workspace_bytes, void* stream);

/*
 * Data-parallel gradient all-reduce (replaces DeepSpeed ZeRO's bucketed
 * gradient reduction configured by finetuner/ds_config.json:27-45 and enabled
 * at FT.py:911-923). One communicator + one high-priority comm stream per rank.
 *   ftk_dp_unique_id  host: fills 128 bytes with an ncclUniqueId (rank 0 calls it, then
 *                     ships the bytes to the other ranks by any means)
 *   ftk_dp_init       host id bytes, rank, world -> opaque handle
 *   ftk_dp_allreduce  on the comm stream: wait for `ready_stream` work issued so far,
 *                     multiply by prescale (1/world for the reference's mean), ncclAllReduce(sum)
 *                     in place over `count` elements of dtype (FTK_BF16/FTK_F32); never blocks the host
 *   ftk_dp_allreduce_after  the same, but the bucket is ready when `ready_event` (a cudaEvent_t the caller had
 *                     recorded, e.g. ftk_lmhead_ce_fused's dw_done_event) fires: the collective can start while the
 *                     producing stream is still busy with later work
 *   ftk_dp_join       makes `stream` wait for everything issued on the comm stream so far
 *   ftk_set_sm_reserve  process-wide: the persistent lm_head kernels launched after this call leave n_sms SMs
 *                     idle (0 = use all). A NCCL kernel cannot share an SM with them, so without a reserve an
 *                     all-reduce issued on the comm stream only starts when the running GEMM ends; with one
 *                     (and NCCL_MAX_CTAS <= n_sms) it overlaps. Host-serial callers toggle it around the
 *                     kernels that should overlap.
 */
int ftk_set_sm_reserve(int32_t n_sms);
typedef struct ftk_dp ftk_dp_t;
int ftk_dp_unique_id(void* id128_host);
int ftk_dp_init(const void* id128_host, int32_t rank, int32_t world, ftk_dp_t** out);
int ftk_dp_allreduce(ftk_dp_t* dp, void* buf, size_t count, int32_t dtype, float prescale, void* ready_stream);
int ftk_dp_allreduce_after(ftk_dp_t* dp, void* buf, size_t count, int32_t dtype, float prescale, void* ready_event);
int ftk_dp_join(ftk_dp_t* dp, void* stream);
int ftk_dp_destroy(ftk_dp_t* dp);

/*
 * Optimizer step on the device, consuming the reduced gradient buckets in place
 * (next row of SURVEY.md 8f). Replaces the reference's DeepSpeed AdamW
 * (finetuner/ds_config.json:10-18, weight_decay 0.01 from FT.py:992, CPU-offloaded
 * by default at ds_config.json:35-40) and its gradient clipping
 * (ds_config.json:44). Arithmetic of torch.optim.AdamW: decoupled weight decay,
 * bias-corrected moments.
 *
 *   ftk_grad_sqnorm  out[0] (+)= sum g^2 over n elements (FTK_BF16 / FTK_F32), reduced in a
 *                    fixed order (deterministic); workspace from ftk_grad_sqnorm_workspace()
 *   ftk_clip_coef    coef[0] = min(1, max_norm / (sqrt(sqnorm[0]) + 1e-6))  (clip_grad_norm_)
 *   ftk_adamw_step   one AdamW update of n elements; step counts from 1.
 *                    param FTK_F32: updated in place (grad FTK_F32 or FTK_BF16, master unused);
 *                    param FTK_BF16: `master` (fp32) is updated and param is its bf16 rounding (grad FTK_BF16, or
 *                    FTK_F32 for fp32 main gradients: the reference accumulates and reduces gradients in fp32,
 *                    ds_config.json:45).
 *                    Every gradient is first multiplied by grad_scale * (grad_scale_dev ? *grad_scale_dev : 1).
 */
size_t ftk_grad_sqnorm_workspace(void);
int ftk_grad_sqnorm(const void* grad, size_t n, int32_t dtype, float* out, int32_t accumulate, void* workspace,
                    size_t workspace_bytes, void* stream);
int ftk_clip_coef(const float* sqnorm, float max_norm, float* coef, void* stream);
int ftk_adamw_step(void* param, const void* grad, float* master, float* exp_avg, float* exp_avg_sq, size_t n,
                   int32_t param_dtype, int32_t grad_dtype, float lr, float beta1, float beta2, float eps,
                   float weight_decay, int64_t step, float grad_scale, const float* grad_scale_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FTK_H */
