/*
 * psrc_b200.h -- C ABI of the B200-native prompted-CLIP training step.
 *
 * Drop-in boundary for the ONE hot path named in BASELINE.json: the prompted
 * CLIP train step of the reference's PromptSRC / CoOp / CoCoOp / MaPLe trainers.
 * Each entry point cites the reference code it replaces (paths relative to the
 * reference repo root).  The reference is pure Python on top of PyTorch, so the
 * "FFI" a maintainer binds is ctypes; nothing in these
 * signatures is a torch type -- only device pointers, sizes and a CUDA stream.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host
 *   - every function is asynchronous on `stream`, stateless, re-entrant and
 *     safe to capture into a CUDA graph (no allocation, no host sync)
 *     (this sentence does not cover psrc_loss_fwd_bwd_noalloc, which keeps a workspace inside the
 *     library and therefore is neither stateless nor re-entrant -- read its own comment)
 *   - return value: 0 = ok, otherwise a PSRC_ERR_* code; the message for the
 *     calling thread's last error is psrc_last_error()
 *   - activations are batch-first: row r = sequence * L + position
 *   - the residual stream is always fp32; "act" tensors (LN outputs, q/k/v,
 *     MLP hidden, and their gradients) are bf16 in the fast mode and fp32 in
 *     the validation mode (dtype arguments below)
 */
#ifndef PSRC_B200_H_
#define PSRC_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* psrc_stream_t; /* cudaStream_t */

enum { PSRC_F32 = 0, PSRC_BF16 = 1, PSRC_F16 = 2 };

enum {
  PSRC_OK = 0,
  PSRC_ERR_ARG = 1,      /* bad argument (shape / alignment / dtype)        */
  PSRC_ERR_CUDA = 2,     /* a CUDA runtime / driver call failed             */
  PSRC_ERR_UNSUPPORTED = 3
};

enum { PSRC_ACT_NONE = 0, PSRC_ACT_QUICKGELU = 1, PSRC_ACT_QUICKGELU_GRAD = 2 };
enum { PSRC_GEMM_AUTO = 0, PSRC_GEMM_SIMT = 1, PSRC_GEMM_TCGEN05 = 2 /* CTA-pair kernel, 1-CTA fallback */,
       PSRC_GEMM_TCGEN05_1CTA = 3 /* force the single-CTA kernel */ };

const char* psrc_last_error(void);
int psrc_version(void);
/* number of kernels this library has launched since load (all threads) --
 * bench.py reports the per-step delta as "gpu_launches" */
long long psrc_launch_count(void);

/* ------------------------------------------------------------------ GEMM
 * out[M,N] = act(alpha * A[M,K] . W[N,K]^T + bias[N]) + residual[M,N]
 * Replaces nn.Linear / nn.MultiheadAttention in/out projections / conv1-as-GEMM
 * / `x @ proj`:  PromptSRC/clip/model.py:174-176 (c_fc, c_proj), :171,:183
 * (MHA in_proj/out_proj), :376,:402 (conv1), :428-429 (visual.proj),
 * PromptSRC/trainers/promptsrc.py:68 (text_projection); and their autograd
 * input-gradients (dX = dY . W, by passing the pre-transposed weight).
 *   act = QUICKGELU       : model.py:162-164; `out2` (optional) receives the
 *                           pre-activation (alpha*acc + bias) for the backward
 *   act = QUICKGELU_GRAD  : out = (alpha*acc + bias) * QuickGELU'(aux[m,n])
 * The tcgen05 backend needs ab_dtype = BF16, K % 8 == 0, 16-byte aligned A/W. */
typedef struct {
  int32_t M, N, K;
  int32_t ab_dtype;            /* dtype of A and W                               */
  const void* A; int64_t lda;  /* row stride in elements                         */
  const void* W; int64_t ldw;
  const float* bias;           /* [N] fp32 or NULL                               */
  int32_t act;
  int32_t aux_dtype;
  const void* aux; int64_t ldaux;
  const float* residual; int64_t ldres; /* fp32 [M,N] or NULL; may alias out: with   */
                                        /* residual == out the sum is formed by the  */
                                        /* TMA store (add at the L2), in place       */
  void* out; int64_t ldo; int32_t out_dtype;
  int32_t out2_dtype;
  void* out2; int64_t ldo2;    /* NULL or pre-activation (QUICKGELU only)        */
  float alpha;
  int32_t backend;             /* PSRC_GEMM_*                                    */
} psrc_gemm_args;
int psrc_gemm(const psrc_gemm_args* args, psrc_stream_t stream);

/* C[M,N] = alpha * sum_k A(m,k) * B(n,k) + beta * C, all fp32, arbitrary element
 * strides (CUDA-core kernel; used for the [B,C] cosine logits and their two
 * gradients): promptsrc.py:193, :201 and autograd of those lines. */
int psrc_gemm_f32_strided(int M, int N, int K,
                          const float* A, int64_t a_sm, int64_t a_sk,
                          const float* B, int64_t b_sn, int64_t b_sk,
                          float* C, int64_t ldc, float alpha, float beta,
                          psrc_stream_t stream);

/* ------------------------------------------------------------- LayerNorm
 * model.py:153-159 (LayerNorm computed in fp32, eps 1e-5, biased variance).
 * x fp32 [rows, D] -> y (out_dtype) ; mean/rstd fp32 [rows] saved for backward. */
int psrc_layernorm_fwd(const float* x, const float* gamma, const float* beta,
                       void* y, int out_dtype, float* mean, float* rstd,
                       int rows, int D, float eps, psrc_stream_t stream);
/* dgrad only (gamma/beta are frozen: promptsrc.py:238-247):
 *   dx_out[r,:] = dx_in[r,:] (residual branch, may be NULL) + LN'(dy)[r,:]
 * dx_out may alias dx_in.  dx_act (optional) receives dx_out cast to act dtype
 * (the A operand of the next dgrad GEMM). */
int psrc_layernorm_bwd(const void* dy, int dy_dtype, const float* x,
                       const float* gamma, const float* mean, const float* rstd,
                       const float* dx_in, float* dx_out, void* dx_act, int act_dtype,
                       int rows, int D, psrc_stream_t stream);

/* ------------------------------------------------------------- Attention
 * nn.MultiheadAttention core (model.py:183 / :227 / :285): per (sequence, head)
 * softmax(q k^T / sqrt(64) [+ causal -inf mask, model.py:592-598]) v, head_dim 64.
 * qkv [n_seq*L, 3*D] with column blocks [q | k | v], head h at h*64 inside each.
 * out [n_seq*L, D]; lse fp32 [n_seq, H, L] (log-sum-exp of the scaled scores).
 * backend: 0 auto, 1 CUDA-core (any dtype), 2 mma.sync tensor-core (bf16), 3 tcgen05/TMEM (bf16, 32 < L <= 256),
 * 5 mma.sync tile kernel (bf16 or fp16, L <= 32: the truncated text tower, one problem per warp; one 16x16 score
 *   tile up to L = 16, 2 x 2 tiles up to 32).  (4 was a CUDA-core short-sequence kernel, superseded by 5.)
 * dtype F16 ("fp16 forward operands"): qkv / out are IEEE half, in the backward dout / dqkv stay bf16; backend 5 only. */
int psrc_attn_fwd(const void* qkv, void* out, float* lse, int dtype,
                  int n_seq, int L, int H, int causal, int backend,
                  psrc_stream_t stream);
int psrc_attn_bwd(const void* qkv, const void* out, const void* dout,
                  const float* lse, void* dqkv, int dtype,
                  int n_seq, int L, int H, int causal, int backend,
                  psrc_stream_t stream);

/* One-query form for the LAST block of a tower: only one row per sequence leaves it
 * (class token, model.py:426; EOT row, promptsrc.py:64-68), so the attention core
 * is needed for the query at row idx[s] (idx == NULL: row 0) of each sequence only.
 * fwd writes that row of out [n_seq*L, D] and that entry of lse [n_seq,H,L]; the
 * other rows / entries are left untouched.  bwd reads that row of dout and writes
 * ALL of dqkv (dK, dV of every key <- the one query; dQ zero for every other row).
 * dtype: F32 / BF16 / F16 (qkv, out); grad_dtype (dout, dqkv): F32 with F32, else BF16. */
int psrc_attn_row_fwd(const void* qkv, void* out, float* lse, int dtype,
                      const int32_t* idx, int n_seq, int L, int H, int causal,
                      psrc_stream_t stream);
int psrc_attn_row_bwd(const void* qkv, const void* dout, const float* lse, void* dqkv,
                      int dtype, int grad_dtype, const int32_t* idx,
                      int n_seq, int L, int H, int causal, psrc_stream_t stream);

/* ------------------------------------------------- token / prompt assembly
 * model.py:401-420 (VisionTransformer.forward up to ln_pre):
 *   im2col: image fp32 [B,3,R,R] -> patches [B*G*G, 3*P*P] (column = c*P*P+ky*P+kx,
 *           i.e. conv1.weight.reshape(width,-1) is the GEMM weight)
 *   assemble: x[b,0]=cls+pos[0]; x[b,1+p]=patch_embed[b,p]+pos[1+p];
 *             x[b,1+G*G+j]=fp16_round(VPT[j]) (model.py:414);
 *             then ln_pre over every row -> fp32 residual stream [B*L, D].
 *   (The backward into VPT is composed from psrc_prompt_grad + psrc_layernorm_bwd on
 *   the n_ctx rounded prompt rows, which are identical for every image.) */
int psrc_im2col_patches(const float* image, void* patches, int out_dtype,
                        int B, int R, int P, psrc_stream_t stream);
int psrc_assemble_vision_tokens(const void* patch_embed, int pe_dtype,
                                const float* cls, const float* pos, const float* vpt,
                                const float* ln_g, const float* ln_b,
                                float* x, int B, int n_patches, int n_ctx, int D,
                                float eps, psrc_stream_t stream);
/* promptsrc.py:144-168 (construct_prompts) + :60 (+ positional_embedding),
 * coop.py:259-296 (class_token_position == "end"), truncated to L positions
 * (exact under the causal mask).  ctx is [n_ctx,D] (ctx_per_class=0)
 * or [C,n_ctx,D] (class-specific contexts, coop.py:222-230).  CoCoOp's
 * instance-conditioned prompts (cocoop.py:173-198) are n_groups > 1 with
 * ctx_bias [n_groups, D]: sequence s = g*C + c uses ctx[j] + ctx_bias[g].
 *   x[s,0]=prefix[c]+pos[0]; x[s,1+j]=ctx[j](+bias[g])+pos[1+j];
 *   x[s,1+n_ctx+k]=suffix[c,k]+pos[1+n_ctx+k] */
int psrc_assemble_text_prompts(const float* prefix, const float* suffix, int suffix_len,
                               const float* ctx, int ctx_per_class,
                               const float* ctx_bias, int n_groups, const float* pos,
                               float* x, int C, int L, int n_ctx, int D,
                               psrc_stream_t stream);
/* model.py:232-252 deep-prompt replacement: overwrite rows [start, start+n_ctx) of
 * every sequence with fp16_round(prompt[j]) (or the raw value if round_fp16 == 0). */
int psrc_insert_prompt(float* x, const float* prompt, int n_seq, int L, int start,
                       int n_ctx, int D, int round_fp16, psrc_stream_t stream);
/* backward of the replacement: grad[j,:] (+)= sum_seq r(dx[seq,start+j,:]); then the
 * rows are zeroed when zero_rows != 0 (no gradient reaches the layer below there),
 * in dx and in its optional act-dtype copy dx_act (the next dgrad GEMM's A operand).
 * round_fp16 != 0: r = fp16 round trip of every per-sequence element before the sum --
 * autograd of the reference's `.expand(N).half()` casts the incoming gradient to fp16
 * (model.py:238,249,414); r = identity for `ctx`, which has no `.half()`. */
int psrc_prompt_grad(float* dx, void* dx_act, int act_dtype, float* grad, int n_seq,
                     int L, int start, int n_ctx, int D, int zero_rows, int accumulate,
                     int round_fp16, psrc_stream_t stream);

/* ------------------------------------------------------------------ heads
 * Gather one row per sequence (vision: row 0, model.py:426; text: EOT row
 * argmax(tokenized_prompts), promptsrc.py:68), LayerNorm it (ln_post / ln_final)
 * and emit it in act dtype for the projection GEMM.  idx == NULL -> row 0. */
int psrc_gather_ln_fwd(const float* x, const int32_t* idx, const float* gamma,
                       const float* beta, void* y, int out_dtype, float* mean,
                       float* rstd, int n_seq, int L, int D, float eps,
                       psrc_stream_t stream);
/* backward: dx (fp32 [n_seq*L, D]) is fully written: zeros everywhere except the
 * gathered rows, which receive LN'(dy) (+ dx_in at those rows when dx_in != NULL --
 * the residual-branch gradient; dx_in may alias dx); dx_act (optional) is the same
 * in act dtype.  With dx_in this is also the LN2 backward of a block whose MLP ran
 * on the gathered rows only (the last block: only ln_post(x[:,0]) / the EOT row is
 * ever read, model.py:426, promptsrc.py:64-68). */
int psrc_gather_ln_bwd(const void* dy, int dy_dtype, const float* x, const int32_t* idx,
                       const float* gamma, const float* mean, const float* rstd,
                       const float* dx_in, float* dx, void* dx_act, int act_dtype,
                       int n_seq, int L, int D, psrc_stream_t stream);
/* promptsrc.py:190-191 `f / f.norm(dim=-1, keepdim=True)`; inv_norm saved. */
int psrc_l2norm_fwd(const float* x, float* y, float* inv_norm, int rows, int D,
                    psrc_stream_t stream);
/* dx = (dy - y * <y,dy>) * inv_norm */
int psrc_l2norm_bwd(const float* dy, const float* y, const float* inv_norm, float* dx,
                    int rows, int D, psrc_stream_t stream);

/* ------------------------------------------ trainable small layers (MaPLe, CoCoOp)
 * The only trainable dense layers on the path act on a handful of rows; fp32, batched over G
 * independent (X, W, b) groups, all tensors contiguous:
 *   X [G,R,K]  W [G,N,K]  b [G,N] (or NULL)  Y [G,R,N]
 * MaPLe coupling functions (PromptSRC/trainers/maple.py:145 `proj`, :155-156
 * `compound_prompt_projections`, applied at :199-204): G = prompt depth, R = n_ctx.
 * CoCoOp meta-net (PromptSRC/trainers/cocoop.py:139-143, applied at :187): G = 1, R = images,
 * linear1 with relu != 0, then linear2. */
int psrc_small_linear_fwd(const float* X, const float* W, const float* b, float* Y,
                          int G, int R, int N, int K, int relu, psrc_stream_t stream);
/* autograd of the above: dW [G,N,K] = dY^T . X, db [G,N] = sum_r dY (both overwritten, skipped
 * when dW == NULL), dX [G,R,K] (+)= dY . W (skipped when dX == NULL).  Y_relu != NULL: the layer
 * ended in a ReLU whose OUTPUT is Y_relu; dY is masked by (Y_relu > 0) first. */
int psrc_small_linear_bwd(const float* X, const float* W, const float* dY, const float* Y_relu,
                          float* dX, int accumulate_dX, float* dW, float* db,
                          int G, int R, int N, int K, psrc_stream_t stream);
/* CoCoOp logits with one text-feature set per image (cocoop.py:246-255):
 *   logits[b,c] = scale * <f_hat[b,:], t_hat[b,c,:]>,  t_hat [B,C,E]
 * backward: d_that[b,c,:] = scale * d_logits[b,c] * f_hat[b,:] (the image tower is frozen, so no
 * gradient w.r.t. f_hat is produced). */
int psrc_group_logits_fwd(const float* f_hat, const float* t_hat, float* logits, int B, int C,
                          int E, float scale, psrc_stream_t stream);
int psrc_group_logits_bwd(const float* f_hat, const float* d_logits, float* d_that, int B, int C,
                          int E, float scale, psrc_stream_t stream);
/* autograd of `ctx_shifted = ctx[None] + bias[:,None]` (cocoop.py:185-188): gg [G,n_ctx,D] is the
 * context-row gradient summed over the classes of image g (psrc_prompt_grad per image);
 * dctx[j,:] = sum_g gg[g,j,:], dbias[g,:] = sum_j gg[g,j,:]. */
int psrc_ctx_bias_grad(const float* gg, float* dctx, float* dbias, int G, int n_ctx, int D,
                       psrc_stream_t stream);

/* CoOp CLASS_TOKEN_POSITION "middle" / "front" (PromptSRC/trainers/coop.py:278-293): the prompt of class s is the
 * "end" layout [prefix | ctx | class tokens | rest] with its rows permuted (per-class, because the class-name length
 * differs).  perm int32 [n_seq, L], a bijection of [0, L) per sequence.
 *   inverse == 0: dst[s,p,:] = src[s,perm[s,p],:] (+ add[p,:], the positional embedding, may be NULL)
 *   inverse != 0: dst[s,perm[s,p],:] = src[s,p,:]   (autograd of the gather: routes d(x0) back to the "end" layout) */
int psrc_permute_rows(const float* src, const int32_t* perm, const float* add, float* dst,
                      int n_seq, int L, int D, int inverse, psrc_stream_t stream);

/* ------------------------------------------------------------------- loss
 * PromptSRC objective and its gradient in one pass (promptsrc.py:204, :304-324;
 * focal CE: coop.py:146-163 / promptsrc_pets_focal.py:93-117):
 *   ce    = mean_b CE(logits_b, y_b)                 (alpha == NULL)
 *         = mean_b alpha[y_b] (1-p_b)^gamma CE_b      (focal, alpha != NULL)
 *   kl    = w_kl * sum_{b,c} exp(q)(q - p) / (Bg*C),  p = log_softmax(logits),
 *           q = log_softmax(zs_logits)                 (zs_logits NULL -> skipped)
 *   l1_t  = w_text  * mean|t_hat - t_zs|   over [Ct, E]   (NULL -> skipped)
 *   l1_i  = w_image * mean|f_hat - f_zs|   over [B, E]
 * B_global / C_l1_global are the divisors (global batch, global class count) so a
 * rank holding a shard produces its additive share of the single-device loss.
 * losses[4] = {ce, kl, l1_text, l1_image} (already weighted; partial sums of this
 * rank); d_logits [B,C]; d_that [Ct,E] and d_fhat [B,E] receive ONLY the L1 part
 * (the logits part is added by the caller's strided GEMMs with beta = 1). */
typedef struct {
  int32_t B, C, E, Ct;          /* Ct = rows of t_hat handled here (class shard)   */
  int32_t B_global, Ct_global;
  const float* logits; const float* zs_logits; const int32_t* labels;
  const float* focal_alpha; float focal_gamma;
  const float* t_hat; const float* t_zs; const float* f_hat; const float* f_zs;
  float w_kl, w_text, w_image;
  float* d_logits; float* d_that; float* d_fhat;
  float* losses;                /* [4] fp32, overwritten                           */
  float* row_scratch;           /* [>= 4*B + 2*296] fp32 workspace                 */
} psrc_loss_args;
int psrc_loss_fwd_bwd(const psrc_loss_args* args, psrc_stream_t stream);

/* The same objective for hosts that have no workspace to give.  (No counterpart in the reference:
 * PyTorch allocates freely; this exists only for embedders of this library.)
 *
 * Who needs it: a caller that receives nothing but its input and output buffers and is not allowed to
 * obtain device memory itself -- a plugin inside a runtime that has already reserved the whole GPU,
 * code that runs while some thread is capturing a CUDA graph (cudaMalloc is illegal then), or a
 * latency-critical loop that must stay clear of the allocator's locks.  Such a caller cannot supply
 * `row_scratch`, so here the workspace is a fixed array inside the library's device image.
 *
 *   - args->row_scratch is ignored (NULL is fine); everything else means what it means above, and
 *     the results are those of the entry point above on the same inputs.
 *   - limit: args->B <= 4096 logits rows per call (C, Ct and E are not limited -- the workspace holds
 *     per-row and per-CTA partial sums only).  A larger B returns PSRC_ERR_ARG with a message that
 *     names the limit; split the batch or use the entry point above.
 *   - cost: 34 KB of device memory on every GPU that loads the library, held until the process exits.
 *   - the call performs no cudaMalloc / cudaFree and calls no allocating library routine.
 *   - the FIRST call on each device looks up that device's address of the array
 *     (cudaGetSymbolAddress, with the device current) and remembers it.  Make that first call during
 *     start-up, with the target device current, before any stream capture begins; `stream` must belong
 *     to the current device, as for every entry point here.
 *   - all calls on one GPU share the one array.  It still holds the previous call's partial sums when
 *     the next call starts; that is harmless because each call overwrites every slot before reading it
 *     (nothing needs to start at zero).  But two calls that are in flight on the same GPU at the same
 *     time -- different streams without an ordering between them -- overwrite each other's partial
 *     sums and return wrong losses: run them on one stream or order them with events. */
int psrc_loss_fwd_bwd_noalloc(const psrc_loss_args* args, psrc_stream_t stream);

/* torch.optim.SGD semantics (Dassl optim/optimizer.py:105-113) on a flat buffer:
 * g += wd*p ; buf = first ? g : mom*buf + (1-damp)*g ; p -= lr*(nesterov? g+mom*buf : buf)
 * lr is read from device memory so a captured graph follows the scheduler. */
int psrc_sgd_step(float* p, const float* g, float* buf, const float* lr_dev,
                  float momentum, float dampening, float weight_decay, int nesterov,
                  int first_step, int64_t n, psrc_stream_t stream);

/* y[i] += a * x[i] (fp32): joins the reduce-scattered logits part of dL/dt_hat with the local
 * L1 part; also the GPA running average over the prompt parameters (promptsrc.py:341-347). */
int psrc_axpy(float* y, const float* x, float a, int64_t n, psrc_stream_t stream);

/* utility casts (fp32 <-> bf16) */
int psrc_cast(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n,
              psrc_stream_t stream);

/* ---------------------------------------------------- input pipeline
 * uint8 HWC images -> CLIP-normalised fp32 [B,3,R,R] in one launch.  Replaces the per-image PIL /
 * torchvision chain Dassl composes (Dassl.pytorch/dassl/data/transforms/transforms.py:176-326 train:
 * RandomResizedCrop -> RandomHorizontalFlip -> ToTensor -> Normalize; :329-357 test: Resize ->
 * CenterCrop -> ToTensor -> Normalize); the random parameters are drawn on the host with the
 * reference's own RNG calls (psrc_b200/input_pipeline.py) and arrive here as geometry:
 *   src      uint8, image b at src + b*image_stride, pixel (y,x,c) at y*row_stride + 3*x + c
 *   geom_f   [B,4] fp32  {sx, sy, ox, oy}: source centre of output pixel x = (x + ox + 0.5) * sx
 *   geom_i   [B,5] int32 {left, top, w, h, flip}: the sampled window (coordinates are relative to
 *            its origin and the filter taps are clamped to it), flip = mirror the output columns
 *   interpolation 0 = bilinear, 1 = bicubic (a = -0.5), both antialiased like PIL (support scaled by
 *            max(scale, 1)); quantize 0 = float, 1 = round the result to uint8 (the PIL image
 *            handed to ToTensor), 2 = also round after the horizontal pass (PIL's two-pass resize)
 *   max_taps >= the largest filter window of the batch (host-computed: 2*ceil(support)+2)
 *   mean_host / std_host: 3 HOST floats (cfg.INPUT.PIXEL_MEAN / PIXEL_STD)                        */
int psrc_crop_resize_normalize(const uint8_t* src, int64_t image_stride, int32_t row_stride,
                               const float* geom_f, const int32_t* geom_i, float* out, int32_t B,
                               int32_t R, int32_t interpolation, int32_t quantize, int32_t max_taps,
                               const float* mean_host, const float* std_host, psrc_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* PSRC_B200_H_ */
