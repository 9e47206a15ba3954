/* b200llm.h — C-ABI of the B200-native decoder training-step library (libb200llm.so).
 *
 * This is the lower face of the drop-in boundary: plain pointers and sizes,
 * no torch types. Every entry point enqueues hand-written sm_100a kernels on the CUDA stream
 * it is given and returns immediately (no hidden synchronisation, no per-call cudaMalloc;
 * the caller owns all device memory including workspaces).
 *
 * The reference (chemphenoms/Building_LLM_from_scratch) is pure PyTorch, so "the interface
 * each entry replaces" is an ATen call site in the reference's Python; each declaration cites
 * it as file:line relative to the reference root.
 *
 * Return value: 0 = OK; < 0 = argument / shape / alignment error (text in b200_last_error());
 * > 0 = a cudaError_t / CUresult passed through.
 *
 * Conventions: all matrices are row-major; "ld*" is the row pitch in ELEMENTS; bf16 data is
 * the raw 16-bit brain-float; device pointers must be 16-byte aligned unless stated otherwise;
 * `stream` is a cudaStream_t passed as void*.
 */
#ifndef B200LLM_H_
#define B200LLM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_ERR_ARG (-1)
#define B200_ERR_UNSUPPORTED (-2)
#define B200_ERR_INTERNAL (-3)

#define B200LLM_VERSION 100

/* ---- library plumbing ------------------------------------------------------------------ */
int b200_version(void);
/* Binds driver entry points and caches device properties; optional (done lazily otherwise). */
int b200_init(int device);
/* SMs the persistent kernels (GEMMs, element-wise grids) size their launches for: 0 = every SM (default), else an
 * even count >= 8 below the device's. A data-parallel step with a large gradient exchange (full fine-tuning: the
 * DDP all-reduce of build_components.py:176 moves every gradient) sets it a few SMs below 148 so that NCCL's CTAs
 * find room while backward runs — a grid of one persistent CTA per SM cannot overlap with another kernel. */
int b200_set_sm_limit(int sms);
/* Thread-local text of the last error returned on this thread. */
const char* b200_last_error(void);
/* Number of kernels this library has launched in this process (bench.py's gpu_launches). */
uint64_t b200_launch_count(void);

/* ---- K1: dense contraction on tcgen05/TMEM fed by TMA ------------------------------------
 * C[M,N] = A1[M,K1]·B1[N,K1]^T  (+ A2[M,K2]·B2[N,K2]^T)  (+ bias[N])  (gelu)  (+ R[M,N])
 * bf16 in, fp32 accumulate in tensor memory, bf16 out. The second product is the rank-r
 * LoRA update accumulated into the same TMEM tile (A2 = s·(x·A), B2 = B^T), pass K2 = 0 to
 * skip it. Replaces nn.Linear / LinearWithLoRA forward:  lora.py:26,46;
 * Models/Llama/Llama3.py:135-137,155,203; Models/Llama/common_components.py:121-124;
 * Models/GPT2/GPT2.py:33-35,49,58-62,121.
 * flags: B200_GEMM_BIAS | B200_GEMM_GELU | B200_GEMM_RESIDUAL.
 * K1, K2, lda*, ldb* must be multiples of 8; A/B bases 16-B aligned. */
#define B200_GEMM_BIAS 1
#define B200_GEMM_GELU 2
#define B200_GEMM_RESIDUAL 4
int b200_gemm_nt_bf16(const void* A1, int64_t lda1, const void* B1, int64_t ldb1, int64_t K1,
                      const void* A2, int64_t lda2, const void* B2, int64_t ldb2, int64_t K2,
                      const void* bias, const void* R, int64_t ldr, void* C, int64_t ldc,
                      int64_t M, int64_t N, int flags, void* stream);

/* Fully general form behind the two wrappers below and above: each operand of each product is
 * K-major (x_mn = 0: [rows, K]) or MN-major (x_mn = 1: stored [K, rows]); out_f32 writes fp32;
 * accumulate adds into C with fp32 atomics and allows split-K (splitk = 0: automatic);
 * trans_out stores element (i, j) at C[j*ldc + i]. Used for the LoRA backward
 * (dX = dY·W + s·u·A^T, dA = x^T·u, dB = t^T·dY; lora.py:26,46). */
int b200_gemm_ex(const void* A1, int64_t lda1, int a1_mn, const void* B1, int64_t ldb1, int b1_mn,
                 int64_t K1, const void* A2, int64_t lda2, int a2_mn, const void* B2, int64_t ldb2,
                 int b2_mn, int64_t K2, const void* bias, const void* R, int64_t ldr, void* C,
                 int64_t ldc, int64_t M, int64_t N, int flags, int out_f32, int accumulate,
                 int trans_out, float alpha, int splitk, void* stream);

/* Several LoRA-wrapped linears of the SAME input (q | k | v, fc1 | fc2) as one GEMM:
 *   C[M, N] = X[M, K]·W[N, K]^T + T[:, seg(n)*r : seg(n)*r + r]·Bcat[:, n]
 * W = row-concatenation of the base weights, T = [t_0 | t_1 | ...] with t_i = s·x·A_i [M, r] (ldt >= nseg*r),
 * Bcat = [B_0 | B_1 | ...] stored [r, N]; seg_start[i] = first output column of linear i (seg_start[0] = 0,
 * multiples of 256, nseg <= 4). Replaces the per-linear forward of lora.py:46 for
 * Models/Llama/Llama3.py:135-137 and common_components.py:121-122. */
int b200_gemm_lora_packed_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, int64_t K,
                               const void* T, int64_t ldt, const void* Bcat, int64_t ldb, int64_t r,
                               int nseg, const int64_t* seg_start, void* C, int64_t ldc, int64_t M,
                               int64_t N, void* stream);

/* Backward of the SwiGLU feed-forward's down projection input, SwiGLU backward fused into the epilogue:
 *   dh = dY[M, K]·W[K, F] (+ U[M, r]·A[F, r]^T),   (da | db)[M, 2F] = SwiGLU'(a, b)·dh,   (a | b) = AB[M, 2F]
 * W = fc3.weight [K = emb_dim, F = hidden_dim] read as an MN-major operand, U = s·dY·B^T, A = fc3's LoRA A
 * (r = 0: no LoRA). dh is never written. Replaces the autograd of common_components.py:121-124
 * (fc3 input gradient, silu(fc1 x) * fc2 x). M >= 256, F a multiple of 256. */
int b200_gemm_swiglu_bwd_bf16(const void* dY, int64_t lddy, const void* W, int64_t ldw, int64_t K,
                              const void* U, int64_t ldu, const void* A, int64_t lda, int64_t r,
                              const void* AB, int64_t ldab, void* dAB, int64_t lddab, int64_t M, int64_t F,
                              void* stream);

/* Forward of the SwiGLU feed-forward's up projections with SwiGLU in the epilogue:
 *   (a | b)[M, 2F] = x[M, K]·[W1; W2]^T (+ LoRA),   h[M, F] = silu(a) * b  (bf16 roundings of the reference)
 * Wil [2F, K]: rows of fc1.weight and fc2.weight interleaved in blocks of 128 (128 rows of W1, the same 128 of W2, ...);
 * T2 [M, 2r] = [s·x·A1 | s·x·A2]; Bil [2r, 2F]: rows 0..r-1 hold B1 in the a columns of that interleaving and zeros
 * in the b columns, rows r..2r-1 the other way round (r2 = 2r, 0: no LoRA). (a | b) is written in the plain packed
 * layout (a in columns [0, F), b in [F, 2F)). Replaces common_components.py:121-123 (fc1, fc2, silu, product).
 * M >= 256, F a multiple of 128. */
int b200_gemm_swiglu_fwd_bf16(const void* X, int64_t ldx, const void* Wil, int64_t ldw, int64_t K,
                              const void* T2, int64_t ldt, const void* Bil, int64_t ldb, int64_t r2,
                              void* AB, int64_t ldab, void* H, int64_t ldh, int64_t M, int64_t F, void* stream);

/* The packed q | k | v projection (b200_gemm_lora_packed_bf16; T / Bcat NULL and r = 0 for plain linears) with
 * compute_rope (Models/Llama/common_components.py:6-35 at Llama3.py:140-141) applied to the first rope_cols output
 * columns (the q and k heads) in the epilogue, with the reference's bf16 roundings: row m is position m % seq_len;
 * cos / sin are [>= seq_len, head_dim] bf16 tables whose two column halves are equal (precompute_rope_params builds
 * them as cat([angles, angles])); head_dim 64 or 128. Pair it with b200_attention_bwd_rope_bf16 in backward. */
int b200_gemm_qkv_rope_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, int64_t K, const void* T,
                            int64_t ldt, const void* Bcat, int64_t ldb, int64_t r, int nseg, const int64_t* seg_start,
                            const void* cos, const void* sin, int64_t seq_len, int64_t head_dim, int64_t rope_cols,
                            void* C, int64_t ldc, int64_t M, int64_t N, void* stream);

/* GPT-2 MLP and residual fusions (Models/GPT2/GPT2.py:58-62, 83-88); all need M, N >= 256 (CTA-pair kernel), bf16.
 * b200_gemm_gelu_fwd_bf16:  Z = X W^T + bias (+ T Bl), H = gelu(Z) (erf form, nn.GELU()), both written by the epilogue
 *                           (Z is what autograd saves for the GELU's backward).
 * b200_gemm_gelu_bwd_bf16:  dZ = gelu'(Z) * (dY W + U A^T): the input gradient of the MLP's second linear with the GELU
 *                           backward in the epilogue (W [K, N] row-major = layers.2.weight; N % 64 == 0).
 * b200_gemm_dropout_residual_bf16: Y = R + dropout(X W^T + bias (+ T Bl)), the `x + drop_shortcut(...)` of GPT2.py:85,87;
 *                           the mask of element i = row * N + col is the one b200_dropout_bf16 draws for (seed, seed_dev)
 *                           (N % 64 == 0).
 * T [M, r] = s X A and Bl [r, N] are the LoRA pair (lora.py:26) or NULL with r = 0; U [M, r] = s dY Bl^T, A [N, r]. */
int b200_gemm_gelu_fwd_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, int64_t K, const void* T,
                            int64_t ldt, const void* Bl, int64_t ldb, int64_t r, const void* bias, void* Z,
                            int64_t ldz, void* H, int64_t ldh, int64_t M, int64_t N, void* stream);
int b200_gemm_gelu_bwd_bf16(const void* dY, int64_t lddy, const void* W, int64_t ldw, int64_t K, const void* U,
                            int64_t ldu, const void* A, int64_t lda, int64_t r, const void* Z, int64_t ldz,
                            void* dZ, int64_t lddz, int64_t M, int64_t N, void* stream);
int b200_gemm_dropout_residual_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, int64_t K,
                                    const void* T, int64_t ldt, const void* Bl, int64_t ldb, int64_t r,
                                    const void* bias, const void* R, int64_t ldr, void* Y, int64_t ldy,
                                    int64_t M, int64_t N, float p, uint64_t seed, const uint64_t* seed_dev,
                                    void* stream);

/* Optional scratch for the large-GEMM kernel (one per device, caller-owned, 256-byte aligned,
 * b200_gemm_workspace_bytes() bytes, ZERO-FILLED when registered; pass NULL to unregister). With it,
 * b200_gemm_ex / b200_gemm_nt_bf16 cut the tiles of a last, at most half-filled wave along K so every
 * CTA pair stays busy (an 8192 x 2048 output is 256 tiles on 74 pairs); the partial sums meet in the
 * scratch, whose counters every launch leaves zero again. The device-wide scratch serves every stream
 * that has none of its own, so GEMMs using it must not overlap on different streams of the device; a
 * host that does run GEMMs concurrently on several streams registers one scratch per stream with
 * b200_gemm_set_stream_workspace (same size / alignment / zero-fill rules; NULL unregisters), which
 * makes every entry point re-entrant. No reference counterpart (scheduling detail of K1). */
/* Diagnostics (no reference counterpart): launches so far by scheduling path, out[0..n): [0] CTA-pair kernel,
 * [1] of those with a K-split tail wave, [2] single-CTA kernel, [3] grouped tile-scheduled, [4] grouped stream-K,
 * [5] packed LoRA segments, [6] SwiGLU-forward epilogue, [7] SwiGLU-backward epilogue, [8] grouped skinny GEMMs on
 * K-splitting CTA pairs. Returns the number of paths. */
int b200_gemm_path_counts(uint64_t* out, int n);
/* Scheduling switches for A/B measurements and tests (no reference counterpart). value: -1 = the library's own
 * heuristic (default), 0 = off, 1 = forced. Initial values come from B200_GEMM_TAIL (0/1), B200_GEMM_STREAMK_FORCE
 * (0/1), B200_GEMM_ORDER (m/n) and B200_GEMM_SKINNY_PAIR (0/1), read once per process. */
#define B200_GEMM_OPT_TAIL_SPLIT 0  /* K-split of a partly filled last wave (CTA-pair kernel) */
#define B200_GEMM_OPT_STREAMK 1     /* stream-K scheduling of the grouped skinny GEMM */
#define B200_GEMM_OPT_N_FASTEST 2   /* tile walk order of the CTA-pair kernel: 1 = N tiles of an M row back to back */
#define B200_GEMM_OPT_SKINNY_PAIR 3 /* grouped skinny GEMMs on CTA pairs that split K and reduce through distributed shared memory */
int b200_gemm_set_option(int option, int value);
size_t b200_gemm_workspace_bytes(void);
int b200_gemm_set_workspace(void* workspace, size_t bytes);
int b200_gemm_set_stream_workspace(void* stream, void* workspace, size_t bytes);

/* Grouped launch of up to B200_GEMM_GROUP_MAX independent skinny GEMMs (every N <= 64, one
 * operand pair each; same operand conventions as b200_gemm_ex) in ONE persistent kernel. Used
 * for the LoRA weight gradients dA = x^T u, dB = (dY^T t)^T of several linears at once
 * (lora.py:26 under autograd) and for the rank-r projections t = x A of linears that share x. */
#define B200_GEMM_GROUP_MAX 16
typedef struct {
  const void* A; int64_t lda; int a_mn;
  const void* B; int64_t ldb; int b_mn;
  void* C; int64_t ldc;
  int64_t M, N, K;
  int out_f32, accumulate, trans_out;
  float alpha;
} b200_gemm_problem;
int b200_gemm_grouped_bf16(int nprob, const b200_gemm_problem* probs, void* stream);

/* Same group, stream-K scheduled: the (problem, 128-row tile, 64-wide K block) units of all problems
 * are cut into equal ranges, one per SM, so every SM streams the same number of bytes whatever the
 * shapes (u = dY·B^T of the vocab projection is 32 tiles x 2004 K blocks; t = x·A is 64 x 32).
 * Tiles whose K range is shared by several CTAs are summed in `workspace` (fp32) and written by the
 * CTA that completes them. The caller owns `workspace`: b200_gemm_grouped_workspace_bytes() bytes,
 * 256-byte aligned, ZERO-FILLED before the first use; every launch leaves it zero-filled again, so one
 * buffer serves all launches on a stream (launches on different streams need their own).
 * Same call sites as b200_gemm_grouped_bf16 (lora.py:26 forward and under autograd). */
size_t b200_gemm_grouped_workspace_bytes(int nprob, const b200_gemm_problem* probs);
int b200_gemm_grouped_streamk_bf16(int nprob, const b200_gemm_problem* probs, void* workspace,
                                   size_t workspace_bytes, void* stream);

/* General bf16 GEMM with operand layouts, C[M,N] (+)= op(A)·op(B):
 *   transa = 0: A is [M,K] (K contiguous);  1: A is stored [K,M] (M contiguous)
 *   transb = 0: B is [N,K] (K contiguous);  1: B is stored [K,N] (N contiguous)
 * out_f32 != 0 writes fp32 C (used for gradient accumulation), `accumulate` adds to C.
 * Replaces the autograd MmBackward of the call sites above (dX = dY·W, dW = dY^T·X,
 * LoRA dA / dB: lora.py:26). */
int b200_gemm_bf16(const void* A, int64_t lda, int transa, const void* B, int64_t ldb, int transb,
                   void* C, int64_t ldc, int out_f32, int accumulate, float alpha, int64_t M,
                   int64_t N, int64_t K, void* stream);

/* ---- K4: RMSNorm (Models/Llama/common_components.py:54-70) ------------------------------
 * y = (x * rsqrt(mean(x^2) + eps)) * w, x,y bf16 [M,d], w fp32 [d]. With bf16_steps != 0 the
 * intermediate roundings of the reference's bf16 eager evaluation (x^2, mean, +eps, rsqrt,
 * x*r rounded to bf16 before the fp32 weight multiply) are reproduced. If `res` is non-NULL
 * the kernel first forms s = x + res (bf16), writes it to `sum_out` and normalises s
 * (fuses the residual add of Llama3.py:178,180). d % 8 == 0, d <= 16384. */
int b200_rmsnorm_fwd_bf16(const void* x, const void* res, const float* w, void* sum_out, void* y,
                          int64_t M, int64_t d, float eps, int bf16_steps, void* stream);
/* dx = r*(g - xn*mean(g*xn)), g = dy*w; dw[d] (fp32, may be NULL) += sum_m dy*xn. If dres_in
 * is non-NULL it is added to dx (gradient arriving through the residual branch). */
int b200_rmsnorm_bwd_bf16(const void* x, const float* w, const void* dy, const void* dres_in,
                          void* dx, float* dw, int64_t M, int64_t d, float eps, void* stream);
/* fp32 instantiation for the 1e-4 parity mode (no intermediate roundings). */
int b200_rmsnorm_fwd_f32(const float* x, const float* w, float* y, int64_t M, int64_t d, float eps,
                         void* stream);
int b200_rmsnorm_bwd_f32(const float* x, const float* w, const float* dy, float* dx, float* dw,
                         int64_t M, int64_t d, float eps, void* stream);

/* ---- K4': LayerNorm (nn.LayerNorm, Models/GPT2/GPT2.py:79-80,85,87,105,120), eps 1e-5 ---- */
int b200_layernorm_fwd_bf16(const void* x, const void* w, const void* b, void* y, float* mean,
                            float* rstd, int64_t M, int64_t d, float eps, void* stream);
int b200_layernorm_bwd_bf16(const void* x, const void* w, const void* dy, const float* mean,
                            const float* rstd, void* dx, float* dw, float* db, int64_t M,
                            int64_t d, void* stream);
/* The same with the gradient of the residual stream added to dx as it is written (x feeds both the norm and the
 * block's residual add, GPT2.py:83-87: autograd would sum the two contributions with a separate add kernel); dres may be
 * NULL. */
int b200_layernorm_bwd_res_bf16(const void* x, const void* w, const void* dy, const void* dres, const float* mean,
                                const float* rstd, void* dx, float* dw, float* db, int64_t M, int64_t d,
                                void* stream);

/* ---- K5: SwiGLU gate (Models/Llama/common_components.py:78-88,121-123) --------------------
 * h = silu(a) * b; a, b [M,F] with row pitches lda, ldb (they may be the two halves of one
 * fused fc1|fc2 GEMM output), h [M,F] contiguous. bf16_steps reproduces the bf16 roundings of
 * sigmoid, a*sig and the product. F % 8 == 0. */
int b200_swiglu_fwd_bf16(const void* a, int64_t lda, const void* b, int64_t ldb, void* h, int64_t M,
                         int64_t F, int bf16_steps, void* stream);
/* da = dh*b*silu'(a), db = dh*silu(a); dh [M,F] contiguous. */
int b200_swiglu_bwd_bf16(const void* a, int64_t lda, const void* b, int64_t ldb, const void* dh,
                         void* da, int64_t ldda, void* db, int64_t lddb, int64_t M, int64_t F,
                         void* stream);

/* ---- K5': GELU (erf form, nn.GELU(), Models/GPT2/GPT2.py:60) ------------------------------ */
int b200_gelu_fwd_bf16(const void* x, void* y, int64_t n, void* stream);
int b200_gelu_bwd_bf16(const void* x, const void* dy, void* dx, int64_t n, void* stream);

/* ---- K6: rotary embedding (Models/Llama/common_components.py:6-35) -------------------------
 * Rotate-half RoPE on `nheads` heads of width hd laid out consecutively from column 0 of a
 * [B*T, ldx] activation (token row = b*T + t, position = t); result written to y (pitch ldy;
 * y may alias x for in-place use). cos/sin are the module's [ctx, hd] tables (bf16).
 * inverse != 0 applies the transpose (the backward of the rotation). bf16_steps reproduces
 * the x*cos, rot*sin roundings of the reference's bf16 evaluation. */
int b200_rope_bf16(const void* x, int64_t ldx, void* y, int64_t ldy, const void* cos, const void* sin,
                   int64_t B, int64_t T, int64_t nheads, int64_t hd, int inverse, int bf16_steps,
                   void* stream);

/* ---- K7: embedding gather (Models/Llama/Llama3.py:197; GPT2.py:110-112) --------------------
 * out[m,:] = tok_w[idx[m],:] (+ pos_w[m % T,:] when pos_w != NULL). idx int64. */
int b200_embedding_fwd_bf16(const int64_t* idx, const void* tok_w, const void* pos_w, void* out,
                            int64_t M, int64_t T, int64_t d, int64_t vocab, void* stream);
/* dW[idx[m],:] += dout[m,:] in fp32 (full pre-training only). */
int b200_embedding_bwd_f32acc(const int64_t* idx, const void* dout, float* dtok_w, float* dpos_w,
                              int64_t M, int64_t T, int64_t d, int64_t vocab, void* stream);

/* ---- K3: cross-entropy over materialised logits (train.py:91) ------------------------------
 * F.cross_entropy(logits, targets) with ignore_index = -100, reduction = mean over the
 * non-ignored rows; all rows ignored -> NaN like the reference. A target outside [0, V) that is
 * not ignore_index makes the loss NaN (the reference's kernel asserts; a device-side assert would
 * poison the context, a NaN loss is the loud, recoverable form of the same refusal). Overwrites
 * `logits` in place with dloss/dlogits (already divided by the valid count) when write_grad != 0.
 * loss_out: 1 fp32. workspace: b200_cross_entropy_workspace_bytes(M) bytes. */
size_t b200_cross_entropy_workspace_bytes(int64_t M);
int b200_cross_entropy_bf16(void* logits, int64_t ld, const int64_t* targets, float* loss_out,
                            int64_t M, int64_t V, int write_grad, void* workspace,
                            size_t workspace_bytes, void* stream);

/* x[0..n) *= *scale_dev (one fp32 in device memory); returns at once when it is exactly 1. The fused
 * vocab projection writes dlogits for d(loss) = 1 during forward; this applies the incoming gradient of
 * a scaled loss (loss / accumulation_steps, loss scaling) in backward without a host read. */
int b200_scale_bf16_dev(void* x, int64_t n, const float* scale_dev, void* stream);

/* Row compaction in front of the vocab projection: rows whose target is ignore_index (-100)
 * contribute nothing to the mean loss of train.py:91 nor to any gradient, so the projection and
 * the cross-entropy only have to see the others. b200_compact_rows lists the non-ignored rows
 * (idx[capacity], -1 padded; meta[0] = their count, meta[1] = 1 if count > capacity),
 * b200_gather_rows_bf16 builds the compact activation / target arrays (an overflowing compaction
 * is turned into a NaN loss downstream rather than silently dropping rows), and
 * b200_scatter_rows_bf16 puts the compact input gradient back (caller zeroes dst first). */
int b200_compact_rows(const int64_t* targets, int64_t M, int64_t capacity, int32_t* idx, int32_t* meta,
                      void* stream);
int b200_gather_rows_bf16(const void* x, int64_t ldx, const int64_t* targets, const int32_t* idx,
                          const int32_t* meta, void* out, int64_t* targets_out, int64_t capacity,
                          int64_t d, void* stream);
int b200_scatter_rows_bf16(const void* src, const int32_t* idx, void* dst, int64_t ldd, int64_t capacity,
                           int64_t d, void* stream);

/* Cross-entropy for callers that cannot give the library a workspace. (No counterpart in the
 * reference: it exists for hosts whose device memory is fully planned ahead — every byte of HBM
 * already assigned, allocator calls forbidden while training runs because cudaMalloc stalls all
 * streams and may fail — and for embedders that pass only inputs and outputs.) Same arithmetic
 * and results as b200_cross_entropy_bf16; the (M + 2)-float workspace is a fixed array inside
 * the library, 256 KiB on every GPU that loads it, so M is limited to
 * B200_CE_NOALLOC_MAX_ROWS rows (larger M is rejected with B200_ERR_ARG and a message).
 * The call allocates nothing. All calls on a device share that one array: two calls whose
 * kernels can overlap on the GPU (different streams without an ordering between them) race and
 * give wrong losses — serialise them. The first call on each device looks the array's address
 * up once; make that first call outside any stream capture. */
#define B200_CE_NOALLOC_MAX_ROWS 65536
int b200_cross_entropy_bf16_noalloc(void* logits, int64_t ld, const int64_t* targets,
                                    float* loss_out, int64_t M, int64_t V, int write_grad,
                                    void* stream);

/* ---- K8a: global gradient norm + clip coefficient (train.py:115-120) ----------------------
 * out[0] = ||g||_2 over n elements; out[1] = min(1, max_norm / (out[0] + 1e-6)), the factor
 * clip_grad_norm_ multiplies every gradient by. workspace: b200_grad_norm_workspace_bytes(n)
 * bytes of scratch for the per-block partial sums. _f32 reads an fp32 gradient buffer. */
size_t b200_grad_norm_workspace_bytes(int64_t n);
int b200_grad_norm_bf16(const void* g, int64_t n, float max_norm, float* out, void* workspace,
                        size_t workspace_bytes, void* stream);
int b200_grad_norm_f32(const float* g, int64_t n, float max_norm, float* out, void* workspace,
                       size_t workspace_bytes, void* stream);

/* ---- K8b: AdamW over a flat buffer (torch.optim.AdamW, build_components.py:258) ------------
 * decoupled weight decay, bias correction from `step` (1-based); grad_scale (device pointer,
 * may be NULL) is the clip coefficient from b200_grad_norm_* (out + 1). p, m, v bf16; g bf16
 * or (g_is_f32) fp32. dev_hyper (may be NULL): 7 floats in device memory that override the
 * scalar arguments — fill a host copy with b200_adam_hyper and upload it before replaying a
 * captured CUDA graph. */
int b200_adam_hyper(float* host7, float lr, float beta1, float beta2, float eps, float weight_decay,
                    int64_t step);
int b200_adamw_bf16(void* p, const void* g, int g_is_f32, void* m, void* v, int64_t n, float lr,
                    float beta1, float beta2, float eps, float weight_decay, int64_t step,
                    const float* dev_hyper, const float* grad_scale, void* stream);
int b200_adamw_f32(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1,
                   float beta2, float eps, float weight_decay, int64_t step, const float* dev_hyper,
                   const float* grad_scale, void* stream);
/* fp32 master weights — the bf16_hybrid policy of datautils/mixed_precision.py:24-28 (parameters and
 * optimizer state fp32, compute / communication copy bf16): the same update as b200_adamw_f32 on the
 * master p (g, m, v fp32) that also rewrites p_bf16 = bf16(p), the copy the GEMMs read. */
int b200_adamw_master_f32(float* p, const float* g, float* m, float* v, void* p_bf16, int64_t n, float lr,
                          float beta1, float beta2, float eps, float weight_decay, int64_t step,
                          const float* dev_hyper, const float* grad_scale, void* stream);

/* ---- K9: dtype casts (datautils/mixed_precision.py:10-35) ---------------------------------- */
int b200_cast_f32_to_bf16(const float* in, void* out, int64_t n, void* stream);
int b200_cast_bf16_to_f32(const void* in, float* out, int64_t n, void* stream);

/* ---- helpers: transposed copy and residual add ------------------------------------------- */
/* out[c, r] = alpha * in[r, c]; in [rows, cols] pitch ldi, out [cols, rows] pitch ldo. */
int b200_transpose_bf16(const void* in, int64_t ldi, void* out, int64_t ldo, int64_t rows,
                        int64_t cols, float alpha, void* stream);
int b200_add_bf16(const void* a, const void* b, void* out, int64_t n, void* stream);
/* out[n] += sum_m x[m, n] (fp32 atomics; zero `out` first): the bias gradient of
 * nn.Linear(bias=True) in Models/GPT2/GPT2.py:23,59-61. */
int b200_colsum_bf16(const void* x, int64_t ld, float* out, int64_t M, int64_t N, void* stream);

/* ---- K2: fused causal attention (Models/Llama/Llama3.py:131-153; GPT2.py:29-47) ------------
 * q [B*T, ldq] with h heads of width hd starting at q, k/v likewise with hkv heads; head j
 * of q attends kv head j / (h / hkv). o [B*T, ldo]. lse [B, h, T] fp32 (natural-log
 * logsumexp of the scaled scores) is saved for backward. scale = 1/sqrt(hd). hd in {64,128}. */
int b200_attention_fwd_bf16(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                            int64_t ldv, void* o, int64_t ldo, float* lse, int64_t B, int64_t T,
                            int64_t h, int64_t hkv, int64_t hd, float scale, float dropout_p,
                            uint64_t seed, const uint64_t* seed_dev, void* stream);
/* Backward. dq/dk/dv have the layouts of q/k/v. `workspace` is scratch of
 * b200_attention_bwd_workspace_bytes(B, T, h, hd) bytes: -lse*log2(e) and -rowsum(dO*O) per
 * (b, h, t), laid out per 64-query step. */
size_t b200_attention_bwd_workspace_bytes(int64_t B, int64_t T, int64_t h, int64_t hd);
int b200_attention_bwd_bf16(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                            int64_t ldv, const void* o, int64_t ldo, const void* dout,
                            int64_t lddo, const float* lse, void* workspace, size_t workspace_bytes,
                            void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv,
                            int64_t B, int64_t T, int64_t h, int64_t hkv, int64_t hd, float scale,
                            float dropout_p, uint64_t seed, const uint64_t* seed_dev, void* stream);
/* Backward for q / k that were rotated by the projection's epilogue (b200_gemm_qkv_rope_bf16): dq / dk leave through
 * the transposed rotation, i.e. as gradients of the UN-rotated projections (compute_rope under autograd), so no
 * separate inverse-RoPE pass over dq | dk runs. No dropout on this path (Llama has none). */
int b200_attention_bwd_rope_bf16(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v,
                                 int64_t ldv, const void* o, int64_t ldo, const void* dout, int64_t lddo,
                                 const float* lse, void* workspace, size_t workspace_bytes, void* dq, int64_t lddq,
                                 void* dk, int64_t lddk, void* dv, int64_t lddv, int64_t B, int64_t T, int64_t h,
                                 int64_t hkv, int64_t hd, float scale, const void* cos, const void* sin, void* stream);
/* dropout_p > 0 drops attention weights after the softmax like GPT2.py:43; the mask is a counter-based
 * function of (seed + *seed_dev, b, h, q, k), so backward must be given the forward's seed and seed_dev.
 * seed_dev (may be NULL = 0) is one uint64 in device memory read by the kernels when they run: a step
 * captured in a CUDA graph draws fresh masks on every replay if the host bumps it between replays.
 * b200_attention_dropout_mask_u8 writes that keep mask ([B, H, T, T] bytes) for tests. */
int b200_attention_dropout_mask_u8(void* out, int64_t B, int64_t H, int64_t T, float dropout_p,
                                   uint64_t seed, void* stream);

/* ---- K10: dropout (nn.Dropout, Models/GPT2/GPT2.py:24,81,102) ------------------------------
 * y = (res +) keep(i) * x / (1 - p), keep a counter-based function of (seed + *seed_dev, i) (seed_dev:
 * optional device uint64, see the attention note above): applying the same call to the upstream
 * gradient is the backward. b200_dropout_mask_u8 dumps the mask of a given effective seed. */
int b200_dropout_bf16(const void* x, const void* res, void* y, int64_t n, float p, uint64_t seed,
                      const uint64_t* seed_dev, void* stream);
int b200_dropout_mask_u8(void* out, int64_t n, float p, uint64_t seed, void* stream);

/* ---- fp32 parity mode (csrc/fp32.cu) ------------------------------------------------------------------------
 * The same operators in plain fp32 (FFMA accumulation, no tensor cores) so that the fp32 tolerance class of the
 * north star — 1e-4 relative per op — can be checked on the GPU: the reference's GPT-2 modules are fp32 as written
 * (Models/GPT2/GPT2.py:20-23,59-61) and its Llama sub-modules accept dtype=float32. Correctness mode, not tuned.
 * b200_gemm_f32: C[b][m,n] = alpha * sum_k A[b][m,k] B[b][n,k] (+ bias[n]) (+ beta C[b][m,n]); every matrix is given by
 * (pointer, batch stride, row stride, column stride) in elements, so transposes and head slices are views.
 * b200_softmax_causal_f32: rows of [.., T, T] scores: softmax over k <= q of S * scale, zeros beyond (GPT2.py:38-42,
 * Llama3.py:148-152); b200_softmax_bwd_f32: dS = P (dP - sum_k dP P) scale in place over dP.
 * b200_elementwise_f32 op: 0 y = gelu(a) (erf), 1 y = b gelu'(a), 2 y = a + b, 3 y = silu(a) b, 4 y = a * b[0]. */
int b200_gemm_f32(const float* A, int64_t a_bs, int64_t a_rs, int64_t a_cs, const float* B, int64_t b_bs, int64_t b_rs,
                  int64_t b_cs, const float* bias, float* C, int64_t c_bs, int64_t c_rs, int64_t c_cs, int64_t batch,
                  int64_t M, int64_t N, int64_t K, float alpha, float beta, void* stream);
int b200_softmax_causal_f32(float* S, int64_t rows, int64_t T, float scale, void* stream);
int b200_softmax_bwd_f32(const float* P, float* dP, int64_t rows, int64_t T, float scale, void* stream);
int b200_layernorm_fwd_f32(const float* x, const float* w, const float* b, float* y, float* mean, float* rstd, int64_t M,
                           int64_t d, float eps, void* stream);
int b200_layernorm_bwd_f32(const float* x, const float* w, const float* dy, const float* mean, const float* rstd, float* dx,
                           float* dw, float* db, int64_t M, int64_t d, void* stream);
int b200_elementwise_f32(const float* a, const float* b, float* y, int64_t n, int op, void* stream);
int b200_swiglu_bwd_f32(const float* a, const float* b, const float* dh, float* da, float* db, int64_t n, void* stream);
int b200_rope_f32(const float* x, float* y, const float* cos, const float* sin, int64_t rows, int64_t T, int64_t nheads,
                  int64_t hd, int inverse, void* stream);
size_t b200_cross_entropy_f32_workspace_bytes(int64_t M);
int b200_cross_entropy_f32(float* logits, int64_t ld, const int64_t* targets, float* loss_out, int64_t M, int64_t V,
                           int write_grad, void* workspace, size_t workspace_bytes, void* stream);
int b200_embedding_fwd_f32(const int64_t* idx, const float* tok_w, const float* pos_w, float* out, int64_t M, int64_t T,
                           int64_t d, void* stream);
int b200_embedding_bwd_f32(const int64_t* idx, const float* dout, float* dtok_w, float* dpos_w, int64_t M, int64_t T,
                           int64_t d, void* stream);

/* ---- K2 decode: one new query token per sequence against a KV cache ---------------------------
 * o[b, head, :] = softmax(q[b, head, :]·K[b, 0..T, kv(head), :]^T * scale)·V[b, 0..T, kv(head), :]
 * q, o: [B, h*hd] (row pitches ldq, ldo); k_cache, v_cache: [B, t_max, hkv*hd] bf16 with row pitch ldc
 * (elements) and batch stride t_max*ldc; keys are stored already rotated (RoPE at their own position).
 * hd 64 or 128, h / hkv <= 8. The reference's generate.py:4-75 keeps no cache (every token re-runs
 * the window through Models/Llama/Llama3.py:131-155 / Models/GPT2/GPT2.py:29-49); this is the same
 * attention evaluated for the last position only. */
int b200_attention_decode_bf16(const void* q, int64_t ldq, const void* k_cache, const void* v_cache,
                               int64_t ldc, int64_t t_max, void* o, int64_t ldo, int64_t B, int64_t T,
                               int64_t h, int64_t hkv, int64_t hd, float scale, void* stream);

/* ---- data side on the device: integer work, bit-exact ----------------------
 * b200_collate_i64 = custom_collate_fn (datautils/dataloader_instruction_finetune.py:10-50) over a packed
 * token buffer: item r is tokens[offsets[r] : offsets[r+1]] with instruction length instr_len[r];
 * width = longest item of the batch (the row length after appending the end-of-text token and shifting),
 * out_cols = min(width, allowed_max_length). inputs / targets: [batch, out_cols] int64.
 * Every target equal to pad_id except the first one of its row becomes ignore_index (this also hits pad
 * ids occurring naturally in the text, like the reference), then targets[:instr_len - 1] = ignore_index
 * with Python slice semantics (instr_len 0 blanks all but the last column). */
int b200_collate_i64(const int64_t* tokens, const int64_t* offsets, const int64_t* instr_len, int64_t batch,
                     int64_t width, int64_t out_cols, int64_t pad_id, int64_t ignore_index,
                     int64_t* inputs, int64_t* targets, void* stream);
/* DatasetPT windows (datautils/dataset.py:21-34) gathered for one batch from a device-resident token
 * stream: inputs[b] = tokens[starts[b] : starts[b]+L], targets[b] = tokens[starts[b]+1 : starts[b]+L+1]. */
int b200_window_batch_i64(const int64_t* tokens, int64_t n_tokens, const int64_t* starts, int64_t batch,
                          int64_t L, int64_t* inputs, int64_t* targets, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200LLM_H_ */
