/* plaincv_b200.h — C ABI of libplaincv_b200.so: the sm_100a kernels behind the training-step hot path of
 * GeorgTirp/plainCV's models/LM causal Transformer (fwd + bwd of train_lm.py:compute_grads).
 *
 * The reference has no FFI of its own (it is pure JAX/Flax; every op below is "whatever XLA emits"), so each
 * entry point cites the reference *operation* it replaces (file:line under the reference checkout).  All
 * functions:
 *   - take raw device pointers, sizes and a cudaStream_t (as void*); they never allocate and never synchronise;
 *   - return 0 on success, a positive cudaError_t, or one of the negative PCV_ERR_* codes below;
 *   - are re-entrant: the only global state is a mutex-guarded cache of TMA descriptors, two launch counters and the
 *     tiling override of pcv_gemm_set_mode (relaxed atomics; every setting computes the same product).
 *     Measurement-only switches (epilogue cycle counters, leaving out operand loads / stores) exist only in
 *     -DPCV_MEASURE builds of the library and are not part of this ABI.
 * Activations are bf16 row-major [tokens, features]; master parameters and gradients are fp32.
 */
#ifndef PLAINCV_B200_H_
#define PLAINCV_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PCV_OK 0
#define PCV_ERR_ARG (-1)    /* bad shape / null pointer / unsupported combination */
#define PCV_ERR_ALIGN (-2)  /* pointer or leading dimension violates the 16-byte rule of TMA / vector access */
#define PCV_ERR_TMAP (-3)   /* cuTensorMapEncodeTiled failed */
#define PCV_ERR_DRIVER (-4) /* driver entry point unavailable */
#define PCV_ERR_SCRATCH (-5) /* problem larger than the built-in scratch of a *_static entry point */

/* GEMM epilogues */
#define PCV_EPI_BF16 0       /* C bf16  = alpha * acc              */
#define PCV_EPI_F32_ATOMIC 1 /* C fp32 += alpha * acc (red.add; split-K allowed) */
#define PCV_EPI_F32 2        /* C fp32  = alpha * acc              */

int pcv_version(void);
/* Number of kernels this library has launched in the calling process since load (bench.py's gpu_launches). */
long long pcv_launch_count(void);

/* C[M,N] (=|+=) alpha * sum_k A(m,k) B(n,k); bf16 operands, fp32 accumulate (tcgen05 / TMEM).
 * a_mn_major = 0: A stored [M][lda] (k contiguous); 1: A stored [K][lda] (m contiguous).  Same for B with N.
 * splits: split-K factor (PCV_EPI_F32_ATOMIC only; 0 = let the library pick one that fills whole waves).
 * Replaces flax.linen.Dense on the hot path — models/LM/transformer.py:110-134 (fc_gate/fc_up/fc2),
 * :194-201 (w_qkv), :246-253 (w_out), :398-405 (lm_head) — and their autodiff transposes (train_lm.py:267). */
int pcv_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                  int a_mn_major, int b_mn_major, int epilogue, int splits, float alpha, void* stream);

/* Tiling override for tests and benchmarks: 0 = automatic (CTA-pair 256x256 tiles, tcgen05 cta_group::2, whenever they
 * fill the machine; single-CTA 128xBN tiles otherwise), 1 = single-CTA tiles only, 2 = CTA-pair tiles always;
 * 3 / 4 = keep / halve the tiles of a sparse last wave; 5 / 6 = static / dynamic tile schedule of the CTA-pair kernel
 * (dynamic: one cluster launched per tile, running pairs take the tiles of the launches they cancel through cluster
 * launch control; env PCV_GEMM_DYNAMIC).  Every setting computes the same product. */
int pcv_gemm_set_mode(int mode);
/* Number of CTA-pair (tcgen05 cta_group::2) GEMM launches since load: lets a test assert which tiling ran. */
long long pcv_gemm_pair_launch_count(void);
/* Programmatic dependent launch for the GEMM / attention / RMSNorm kernels (their prologues may then run under the tail
 * of the preceding kernel in the stream; results are identical).  0 = off (default), 1 = on. */
int pcv_set_pdl(int on);

/* fp32 master -> bf16 compute copy, zero-filling the padded tail columns: dst[rows][dst_ld] <- src[rows][src_ld][:cols].
 * Replaces flax's per-call promote_dtype of `kernel`/`embedding` (transformer.py: every nn.Dense / nn.Embed). */
int pcv_cast_pad_bf16(const float* src, void* dst, int rows, int cols, int src_ld, int dst_ld, void* stream);

/* x[m,:] = bf16(E[ids[m],:])                         — nn.Embed, models/LM/transformer.py:361-369 */
int pcv_embed_fwd(const int32_t* ids, const void* table_bf16, void* x, int M, int d, void* stream);
/* dE[ids[m],:] += dx[m,:] (fp32 atomics)              — transpose of the gather above */
int pcv_embed_bwd(const int32_t* ids, const void* dx, float* dE, int M, int d, int ldE, void* stream);

/* Residual add fused with RMSNorm — Block: x = x + f(norm(x)), models/LM/transformer.py:293,334 with
 * nn.RMSNorm :41-47 (fp32 statistics, eps inside the rsqrt, bf16 output).
 *   res_out = res_in + delta   (delta may be NULL: plain norm of res_in)
 *   y       = bf16( fp32(res_out) * rsqrt(mean(res_out^2) + eps) * scale )
 *   rstd[m] saved for backward. */
int pcv_rmsnorm_residual_fwd(const void* res_in, const void* delta, const float* scale, void* res_out, void* y,
                             float* rstd, int M, int d, float eps, void* stream);
/* Backward of the above. dres_out: gradient arriving at res_out from later consumers (may be NULL);
 * dy: gradient of y.  Writes dres_in (= d delta) and accumulates dscale (fp32 atomics). */
int pcv_rmsnorm_residual_bwd(const void* dy, const void* dres_out, const void* res_out, const float* scale,
                             const float* rstd, void* dres_in, float* dscale, int M, int d, void* stream);

/* h = silu(g) * u with [g | u] the two halves of gu[M][2*hid] — GLU, models/LM/transformer.py:126 */
int pcv_swiglu_fwd(const void* gu, void* h, int M, int hid, void* stream);
int pcv_swiglu_bwd(const void* gu, const void* dh, void* dgu, int M, int hid, void* stream);

/* The GLU up-projection with its activation fused (models/LM/transformer.py:110-126): [g | u] = X[M,K] @ W_gu[K,2*hid]
 * (fc_gate | fc_up side by side, row stride ldw) written to gu[M][ldgu] for the backward pass, and
 * h[M][ldh] = bf16(silu(g) * u) computed from the bf16-rounded g and u in the same epilogue (each CTA pair stages
 * 128 gate columns next to the matching 128 up columns).  Same results as pcv_gemm_bf16 + pcv_swiglu_fwd. */
int pcv_gemm_bf16_swiglu(const void* X, const void* W_gu, void* gu, void* h, int M, int hid, int K, int ldx, int ldw,
                         int ldgu, int ldh, void* stream);
/* Its transpose through fc2 (transformer.py:127-134): dh = dY[M,K] @ W_down[hid,K]^T is rounded to bf16 in the
 * epilogue and turned into [dg | du] = dgu[M][ldgu] with the saved gu; dh never reaches memory.  Same results as
 * pcv_gemm_bf16 (dgrad) + pcv_swiglu_bwd. */
int pcv_gemm_bf16_dswiglu(const void* dY, const void* W_down, const void* gu, void* dgu, int M, int hid, int K,
                          int lddy, int ldw, int ldgu, void* stream);

/* Y[M,N] = bf16( rope( X[M,K] @ W[K,N] ) ): the w_qkv Dense (models/LM/transformer.py:194-201) with the interleaved-pair
 * RoPE of models/LM/embedding.py:28-67 applied in the epilogue to output columns [0, rope_cols) (the q | k thirds,
 * head_dim 64) at position (row % T), in fp32 before the bf16 rounding.  cos_sin = fp32 [T][32][2] = (cos, sin) of
 * precompute_freqs_cis (embedding.py:8-26), built once by the host. */
int pcv_gemm_bf16_rope(const void* X, const void* W, void* Y, int M, int N, int K, int ldx, int ldw, int ldy,
                       const float* cos_sin, int T, int rope_cols, void* stream);
/* The same rotation as a stand-alone in-place pass over the q | k columns of qkv[B*T][3*d]. */
int pcv_rope_qk(void* qkv, const float* cos_sin, int B, int T, int n_heads, void* stream);

/* Flash-style attention over qkv[B*T][3*d] (columns [q|k|v], each head-major [N][hd], hd = 64; q and k ALREADY
 * rotated by pcv_gemm_bf16_rope / pcv_rope_qk): softmax(q k^T / sqrt(hd) + mask) v
 * (jax.nn.dot_product_attention at models/LM/transformer.py:233-240).  Mask: causal, intersected with document
 * blocks when doc_start != NULL (doc_start[b*T + t] = first position of the document containing t; the
 * block-diagonal mask of data/datasets/data_prep_utils.py:14-43, never materialised).
 * out[B*T][d] bf16; lse[B][N][T] fp32 saved for backward.  T must be a multiple of 128. */
int pcv_attn_fwd(const void* qkv, const int32_t* doc_start, void* out, float* lse, int B, int T, int n_heads,
                 void* stream);
/* dqkv[B*T][3*d] (gradient w.r.t. the UN-rotated qkv: the transpose of the rotation is applied to dq and dk here)
 * from dout.  Scratch supplied by the caller: delta[B][N][T] fp32 (rowsum(dout*out)) and dq_acc[B*T][d] fp32
 * (zeroed here; dQ partials of the key tiles are reduce-added into it, then un-rotated). */
int pcv_attn_bwd(const void* qkv, const float* cos_sin, const int32_t* doc_start, const void* out, const void* dout,
                 const float* lse, float* delta, float* dq_acc, void* dqkv, int B, int T, int n_heads, void* stream);

/* pcv_attn_bwd without caller-supplied scratch and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  delta and the fp32 dQ accumulator live in a file-scope device array of the library:
 * limit B*T*n_heads*64 <= 4,194,304 (e.g. 4096 tokens at d = 1024; 16.25 MiB held per GPU until the process exits),
 * beyond it the call returns PCV_ERR_SCRATCH and launches nothing.  All calls share that array (it keeps the previous
 * call's contents; what must start at zero is cleared by each call): calls that overlap on the GPU race — serialise
 * them.  Same arguments, results and error codes as pcv_attn_bwd otherwise. */
int pcv_attn_bwd_static(const void* qkv, const float* cos_sin, const int32_t* doc_start, const void* out,
                        const void* dout, const float* lse, void* dqkv, int B, int T, int n_heads, void* stream);

/* Softmax cross-entropy over bf16 logits[M][ldl] (first V columns valid) with integer labels — _loss_and_acc,
 * train_lm.py:181-186 (fp32 logsumexp - logit[label], mean over all M tokens; acc = first-max argmax == label).
 * Writes per-row loss/correct into row_loss/row_correct and, if dlogits != NULL, overwrites
 * dlogits[M][ldl] = (softmax - onehot) * grad_scale (pad columns zero); dlogits may alias logits. */
int pcv_softmax_ce(const void* logits, const int32_t* labels, void* dlogits, float* row_loss, float* row_correct,
                   int M, int V, int ldl, float grad_scale, void* stream);
/* The LM head fused with softmax cross-entropy (models/LM/transformer.py:398-405 + train_lm.py:181-186), in two launches:
 *  (1) pcv_gemm_bf16_ce: logits[M][ldl] = bf16(X[M,K] @ W) exactly as pcv_gemm_bf16 (W stored [K][ldw] when w_mn_major = 1 —
 *      the lm_head kernel — or [N][ldw] when 0 — the tied embedding table), and, from the same epilogue registers, per
 *      (row, 128-column slot) the online-softmax partials (max, sum exp(logit - max)) of the bf16-rounded logits over the
 *      first n_valid columns: ce_ms = float2 [ceil(N/128)][M] (slot-major).  Always CTA-pair (cta_group::2) tiles.
 *  (2) pcv_ce_from_stats: merges the slots into each row's log-sum-exp (ce_ms is consumed: its first M entries are
 *      overwritten) and makes ONE streaming pass over the logits: reads the bf16
 *      logits, writes dlogits = (softmax - onehot) * grad_scale (may alias logits; pad columns zero), row_loss = lse -
 *      logit[label], row_correct = (first-max argmax == label).  One exponential per logit instead of the two of
 *      pcv_softmax_ce, no per-row staging in shared memory.  Same results as pcv_gemm_bf16 + pcv_softmax_ce up to the
 *      order of the fp32 sums. */
int pcv_gemm_bf16_ce(const void* X, const void* W, void* logits, int M, int N, int K, int ldx, int ldw, int ldl,
                     int w_mn_major, int n_valid, void* ce_ms, void* stream);
int pcv_ce_from_stats(const void* logits, const int32_t* labels, void* ce_ms, int n_slots, void* dlogits,
                      float* row_loss, float* row_correct, int M, int V, int ldl, float grad_scale, void* stream);
/* out[0] = sum(row_loss)/M, out[1] = sum(row_correct)/M  (deterministic single-block reduction) */
int pcv_mean2(const float* row_loss, const float* row_correct, float* out, int M, void* stream);

/* Un-gated MLP activations over n = M * hid_pad bf16 elements.  mode 0: silu — MLP,
 * models/LM/transformer.py:70-97; mode 1: relu(x)^2 — MLPReluSquared, transformer.py:138-165.
 * fwd: h = act(x); bwd: dx = dh * act'(x).  fp32 math, bf16 in / out. */
int pcv_act_fwd(const void* x, void* h, long long n, int mode, void* stream);
int pcv_act_bwd(const void* x, const void* dh, void* dx, long long n, int mode, void* stream);

/* out[i] += in[i] (bf16), helper for gradient joins of the residual stream */
int pcv_add_bf16(const void* a, const void* b, void* out, long long n, void* stream);

/* Gradient buckets on their way to / from a bf16 all-reduce: to_bf16 = 1: dst(bf16)[i] = rn(src(fp32)[i]);
 * to_bf16 = 0: dst(fp32)[i] = src(bf16)[i].  The data-parallel exchange replaces jax.lax.pmean(grads, "data")
 * (train_lm.py:207, 268); the reduction itself is NCCL's. */
int pcv_cast_f32_bf16(const void* src, void* dst, long long n, int to_bf16, void* stream);

/* ---- the step after the hot path: global-norm clip + AdamW over flat fp32 buffers ----
 * partials[b] = sum of g[i]^2 over block b's share (n_partials blocks, fixed order: reproducible norm).
 * Replaces optax.global_norm in _clip_grads, train_lm.py:173-178. */
int pcv_sumsq_partials(const float* g, long long n, float* partials, int n_partials, void* stream);
/* One pass over n parameters: g *= min(1, max_norm / (sqrt(sum partials) + 1e-6)) (skipped when max_norm <= 0);
 * m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; p -= lr (m/(1-b1^step) / (sqrt(v/(1-b2^step)) + eps) + weight_decay p);
 * p_bf16 (optional) = bf16(p), the compute copy.  step is 1-based.  Hyper-parameters are doubles so that 1 - b2 and
 * the bias corrections are formed before rounding to fp32 (the arithmetic on the data is fp32, as in the reference).  norm_out (optional, device) receives the norm.
 * Replaces _clip_grads + apply_grads (train_lm.py:173-178, 316-353) with optax.adamw (optim/factory.py:193-205)
 * and flax's per-call promote_dtype of the parameters. */
int pcv_adamw_step(float* p, const float* g, float* m, float* v, void* p_bf16, long long n, const float* partials,
                   int n_partials, float* norm_out, double max_norm, double lr, double b1, double b2, double eps,
                   double weight_decay, int step, void* stream);

/* ---- forward-mode (tangent) kernels for the GGN matvec of the LM ----
 * The reference forms GGN.v = J^T H (J v) with jax.jvp / jax.vjp around Transformer.__call__
 * (optim/ggn_utils.py:69-153, make_ggn_matvec_fn_lm).  J^T is the backward path above; J v needs the tangents below
 * (a Dense tangent is two pcv_gemm_bf16 launches, RoPE is linear: pcv_rope_qk on the tangent).
 *
 * Tangent of residual add + nn.RMSNorm (transformer.py:41-47, 293, 334).  x[M][d] = the primal residual stream after
 * the add, rstd[M] as saved by pcv_rmsnorm_residual_fwd; tres_out = bf16(tres_in + tdelta) (tdelta may be NULL: then
 * tres_out is unused); ty = rstd*scale*(tx - xh*mean(xh*tx)) + xh*tscale with xh = x*rstd. */
int pcv_rmsnorm_residual_jvp(const void* x, const float* rstd, const float* scale, const float* tscale,
                             const void* tres_in, const void* tdelta, void* tres_out, void* ty, int M, int d,
                             void* stream);
/* th[M][hid] = silu'(g)*tg*u + silu(g)*tu for gu, tgu [M][2*hid] = [gate | up] (GLU, transformer.py:126). */
int pcv_swiglu_jvp(const void* gu, const void* tgu, void* th, int M, int hid, void* stream);
/* th = act'(x)*tx over n bf16 elements; mode as pcv_act_fwd (MLP / MLPReluSquared, transformer.py:70-97, 138-165). */
int pcv_act_jvp(const void* x, const void* tx, void* th, long long n, int mode, void* stream);
/* Tangent of jax.nn.dot_product_attention (transformer.py:233-240) on tcgen05/TMEM: qkv as for pcv_attn_fwd (rotated
 * q, k), tqkv[B*T][3*d] its tangent (rotated likewise), out / lse the forward results; tout[B*T][d] =
 * P.tV + (P*tS).V - rowsum(P*tS)*out with tS = (tQ.K^T + Q.tK^T)/sqrt(64); same causal / document masking. */
int pcv_attn_jvp(const void* qkv, const void* tqkv, const int32_t* doc_start, const void* out, const float* lse,
                 void* tout, int B, int T, int n_heads, void* stream);
/* softmax_cross_entropy_hessian_vec (optim/ggn_utils.py:19-53) with the 1/(B*T) of :122-124 folded in:
 * hv[M][ldl] = (p*v - p*(p.v)) * out_scale per row, p = softmax(fp32(logits[:, :V])); pad columns zero; bf16 in / out;
 * hv may alias vec. */
int pcv_ce_hessian_vec(const void* logits, const void* vec, void* hv, int M, int V, int ldl, float out_scale,
                       void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PLAINCV_B200_H_ */
