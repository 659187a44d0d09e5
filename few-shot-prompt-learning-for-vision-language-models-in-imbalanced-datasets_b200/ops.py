"""Thin tensor-level wrappers over the C ABI (one function per kernel family).

PyTorch only provides device memory and the stream here; all arithmetic runs in the
hand-written sm_100a kernels of ``csrc/``.  Shapes follow ``include/psrc_b200.h``.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import cabi
from .cabi import (ACT_NONE, ACT_QUICKGELU, ACT_QUICKGELU_GRAD, BF16, F32, GEMM_AUTO, GEMM_SIMT,
                   GEMM_TCGEN05, GEMM_TCGEN05_1CTA, dt, ptr)

LN_EPS = 1e-5  # nn.LayerNorm default used by PromptSRC/clip/model.py:153

# debugging knobs (numerics bisection on the GPU box): force a kernel family, 0 = automatic
import os as _os
_FORCE_ATTN = int(_os.environ.get("PSRC_ATTN_BACKEND", "0"))
_FORCE_GEMM = int(_os.environ.get("PSRC_GEMM_BACKEND", "0"))


def _ld(t: torch.Tensor) -> int:
    assert t.dim() == 2 and t.stride(1) == 1, "row-major 2-D tensor expected"
    return t.stride(0)


def gemm(a, w, out, bias=None, act=ACT_NONE, aux=None, residual=None, out2=None, alpha=1.0,
         backend=GEMM_AUTO):
    """out[M,N] = act(alpha * a[M,K] @ w[N,K]^T + bias) + residual   (see psrc_gemm)."""
    M, K = a.shape
    N = w.shape[0]
    assert w.shape[1] == K and out.shape == (M, N) and a.dtype == w.dtype
    g = cabi.GemmArgs()
    g.M, g.N, g.K, g.ab_dtype = M, N, K, dt(a)
    g.A, g.lda, g.W, g.ldw = ptr(a), _ld(a), ptr(w), _ld(w)
    g.bias = ptr(bias)
    g.act = act
    if aux is not None:
        g.aux, g.ldaux, g.aux_dtype = ptr(aux), _ld(aux), dt(aux)
    if residual is not None:
        assert residual.dtype == torch.float32
        g.residual, g.ldres = ptr(residual), _ld(residual)
    g.out, g.ldo, g.out_dtype = ptr(out), _ld(out), dt(out)
    if out2 is not None:
        g.out2, g.ldo2, g.out2_dtype = ptr(out2), _ld(out2), dt(out2)
    g.alpha, g.backend = float(alpha), (backend or _FORCE_GEMM)
    cabi.check(cabi.load().psrc_gemm(C.byref(g), cabi.stream()), "psrc_gemm")
    return out


def gemm_f32_strided(M, N, K, a, a_sm, a_sk, b, b_sn, b_sk, c, alpha=1.0, beta=0.0):
    """c[M,N] = alpha * sum_k a(m,k) b(n,k) + beta * c  (fp32, element strides)."""
    assert a.dtype == b.dtype == c.dtype == torch.float32
    cabi.check(cabi.load().psrc_gemm_f32_strided(M, N, K, ptr(a), a_sm, a_sk, ptr(b), b_sn, b_sk, ptr(c),
                                                 _ld(c), float(alpha), float(beta), cabi.stream()),
               "psrc_gemm_f32_strided")
    return c


def layernorm_fwd(x, gamma, beta, y, mean=None, rstd=None, eps=LN_EPS):
    rows, D = x.shape
    cabi.check(cabi.load().psrc_layernorm_fwd(ptr(x), ptr(gamma), ptr(beta), ptr(y), dt(y), ptr(mean), ptr(rstd),
                                              rows, D, eps, cabi.stream()), "psrc_layernorm_fwd")
    return y


def layernorm_bwd(dy, x, gamma, mean, rstd, dx_out, dx_in=None, dx_act=None):
    rows, D = x.shape
    cabi.check(cabi.load().psrc_layernorm_bwd(ptr(dy), dt(dy), ptr(x), ptr(gamma), ptr(mean), ptr(rstd), ptr(dx_in),
                                              ptr(dx_out), ptr(dx_act), dt(dx_act) if dx_act is not None else F32,
                                              rows, D, cabi.stream()), "psrc_layernorm_bwd")
    return dx_out


def attn_fwd(qkv, out, lse, n_seq, L, H, causal, backend=0):
    backend = backend or _FORCE_ATTN
    cabi.check(cabi.load().psrc_attn_fwd(ptr(qkv), ptr(out), ptr(lse), dt(qkv), n_seq, L, H, int(causal), backend,
                                         cabi.stream()), "psrc_attn_fwd")
    return out


def attn_bwd(qkv, out, dout, lse, dqkv, n_seq, L, H, causal, backend=0):
    backend = backend or _FORCE_ATTN
    cabi.check(cabi.load().psrc_attn_bwd(ptr(qkv), ptr(out), ptr(dout), ptr(lse), ptr(dqkv), dt(qkv), n_seq, L, H,
                                         int(causal), backend, cabi.stream()), "psrc_attn_bwd")
    return dqkv


def attn_row_fwd(qkv, out, lse, n_seq, L, H, causal, idx=None):
    """One-query attention of a tower's last block: the query is row idx[s] (None: row 0) of every sequence; writes
    that row of `out` and that entry of `lse` only (psrc_attn_row_fwd)."""
    cabi.check(cabi.load().psrc_attn_row_fwd(ptr(qkv), ptr(out), ptr(lse), dt(qkv), ptr(idx), n_seq, L, H, int(causal),
                                             cabi.stream()), "psrc_attn_row_fwd")
    return out


def attn_row_bwd(qkv, dout, lse, dqkv, n_seq, L, H, causal, idx=None):
    """Backward of attn_row_fwd: reads the query row of `dout`, writes ALL of dqkv (psrc_attn_row_bwd)."""
    cabi.check(cabi.load().psrc_attn_row_bwd(ptr(qkv), ptr(dout), ptr(lse), ptr(dqkv), dt(qkv), dt(dqkv), ptr(idx),
                                             n_seq, L, H, int(causal), cabi.stream()), "psrc_attn_row_bwd")
    return dqkv


def im2col_patches(image, patches, R, P):
    B = image.shape[0]
    cabi.check(cabi.load().psrc_im2col_patches(ptr(image), ptr(patches), dt(patches), B, R, P, cabi.stream()),
               "psrc_im2col_patches")
    return patches


def assemble_vision_tokens(pe, cls, pos, vpt, ln_g, ln_b, x, B, n_patches, n_ctx, D, eps=LN_EPS):
    cabi.check(cabi.load().psrc_assemble_vision_tokens(ptr(pe), dt(pe), ptr(cls), ptr(pos), ptr(vpt), ptr(ln_g),
                                                       ptr(ln_b), ptr(x), B, n_patches, n_ctx, D, eps,
                                                       cabi.stream()), "psrc_assemble_vision_tokens")
    return x


def assemble_text_prompts(prefix, suffix, ctx, pos, x, C_, L, n_ctx, D, ctx_per_class=False, ctx_bias=None,
                          n_groups=1):
    suffix_len = suffix.shape[1]
    cabi.check(cabi.load().psrc_assemble_text_prompts(ptr(prefix), ptr(suffix), suffix_len, ptr(ctx),
                                                      int(ctx_per_class), ptr(ctx_bias), n_groups, ptr(pos), ptr(x),
                                                      C_, L, n_ctx, D, cabi.stream()), "psrc_assemble_text_prompts")
    return x


def insert_prompt(x, prompt, n_seq, L, start, n_ctx, D, round_fp16=True):
    cabi.check(cabi.load().psrc_insert_prompt(ptr(x), ptr(prompt), n_seq, L, start, n_ctx, D, int(round_fp16),
                                              cabi.stream()), "psrc_insert_prompt")
    return x


def prompt_grad(dx, grad, n_seq, L, start, n_ctx, D, zero_rows=True, accumulate=False, dx_act=None,
                round_fp16=False):
    cabi.check(cabi.load().psrc_prompt_grad(ptr(dx), ptr(dx_act), dt(dx_act) if dx_act is not None else F32,
                                            ptr(grad), n_seq, L, start, n_ctx, D, int(zero_rows),
                                            int(accumulate), int(round_fp16), cabi.stream()), "psrc_prompt_grad")
    return grad


def gather_ln_fwd(x, idx, gamma, beta, y, mean, rstd, n_seq, L, D, eps=LN_EPS):
    cabi.check(cabi.load().psrc_gather_ln_fwd(ptr(x), ptr(idx), ptr(gamma), ptr(beta), ptr(y), dt(y), ptr(mean),
                                              ptr(rstd), n_seq, L, D, eps, cabi.stream()), "psrc_gather_ln_fwd")
    return y


def gather_ln_bwd(dy, x, idx, gamma, mean, rstd, dx, n_seq, L, D, dx_act=None, dx_in=None):
    cabi.check(cabi.load().psrc_gather_ln_bwd(ptr(dy), dt(dy), ptr(x), ptr(idx), ptr(gamma), ptr(mean), ptr(rstd),
                                              ptr(dx_in), ptr(dx), ptr(dx_act), dt(dx_act) if dx_act is not None else F32,
                                              n_seq, L, D, cabi.stream()), "psrc_gather_ln_bwd")
    return dx


def l2norm_fwd(x, y, inv_norm):
    rows, D = x.shape
    cabi.check(cabi.load().psrc_l2norm_fwd(ptr(x), ptr(y), ptr(inv_norm), rows, D, cabi.stream()), "psrc_l2norm_fwd")
    return y


def l2norm_bwd(dy, y, inv_norm, dx):
    rows, D = y.shape
    cabi.check(cabi.load().psrc_l2norm_bwd(ptr(dy), ptr(y), ptr(inv_norm), ptr(dx), rows, D, cabi.stream()),
               "psrc_l2norm_bwd")
    return dx


def loss_fwd_bwd(logits, labels, d_logits, losses, scratch, zs_logits=None, focal_alpha=None, focal_gamma=2.0,
                 t_hat=None, t_zs=None, d_that=None, f_hat=None, f_zs=None, d_fhat=None, w_kl=1.0, w_text=25.0,
                 w_image=10.0, B_global=None, Ct_global=None):
    B, Cn = logits.shape
    a = cabi.LossArgs()
    a.B, a.C = B, Cn
    a.E = (t_hat.shape[1] if t_hat is not None else (f_hat.shape[1] if f_hat is not None else 0))
    a.Ct = t_hat.shape[0] if t_hat is not None else 0
    a.B_global = B_global or B
    a.Ct_global = Ct_global or a.Ct
    assert labels.dtype == torch.int32
    assert scratch.numel() >= 4 * B + 2 * 296
    a.logits, a.zs_logits, a.labels = ptr(logits), ptr(zs_logits), ptr(labels)
    a.focal_alpha, a.focal_gamma = ptr(focal_alpha), float(focal_gamma)
    a.t_hat, a.t_zs, a.f_hat, a.f_zs = ptr(t_hat), ptr(t_zs), ptr(f_hat), ptr(f_zs)
    a.w_kl, a.w_text, a.w_image = float(w_kl), float(w_text), float(w_image)
    a.d_logits, a.d_that, a.d_fhat = ptr(d_logits), ptr(d_that), ptr(d_fhat)
    a.losses, a.row_scratch = ptr(losses), ptr(scratch)
    cabi.check(cabi.load().psrc_loss_fwd_bwd(C.byref(a), cabi.stream()), "psrc_loss_fwd_bwd")
    return losses


def sgd_step(p, g, buf, lr_dev, momentum, dampening, weight_decay, nesterov, first_step):
    cabi.check(cabi.load().psrc_sgd_step(ptr(p), ptr(g), ptr(buf), ptr(lr_dev), momentum, dampening, weight_decay,
                                         int(nesterov), int(first_step), p.numel(), cabi.stream()), "psrc_sgd_step")
    return p


def cast(src, dst):
    assert src.numel() == dst.numel() and src.is_contiguous() and dst.is_contiguous()
    cabi.check(cabi.load().psrc_cast(ptr(src), dt(src), ptr(dst), dt(dst), src.numel(), cabi.stream()), "psrc_cast")
    return dst


def axpy(y, x, a=1.0):
    assert y.dtype == x.dtype == torch.float32 and y.numel() == x.numel() and y.is_contiguous() and x.is_contiguous()
    cabi.check(cabi.load().psrc_axpy(ptr(y), ptr(x), float(a), y.numel(), cabi.stream()), "psrc_axpy")
    return y


def _f32c(*ts):
    for t in ts:
        assert t is None or (t.dtype == torch.float32 and t.is_contiguous()), "contiguous fp32 tensor expected"


def small_linear_fwd(x, w, b, y, relu=False):
    """y[G,R,N] = x[G,R,K] @ w[G,N,K]^T + b[G,N] (optionally ReLU) -- see psrc_small_linear_fwd."""
    G, R, K = x.shape
    N = w.shape[1]
    assert w.shape == (G, N, K) and y.shape == (G, R, N) and (b is None or b.shape == (G, N))
    _f32c(x, w, b, y)
    cabi.check(cabi.load().psrc_small_linear_fwd(ptr(x), ptr(w), ptr(b), ptr(y), G, R, N, K, int(relu), cabi.stream()),
               "psrc_small_linear_fwd")
    return y


def small_linear_bwd(x, w, dy, dx=None, dw=None, db=None, y_relu=None, accumulate_dx=False):
    G, R, K = x.shape
    N = w.shape[1]
    assert w.shape == (G, N, K) and dy.shape == (G, R, N)
    assert dx is None or dx.shape == (G, R, K)
    assert dw is None or (dw.shape == (G, N, K) and db is not None and db.shape == (G, N))
    _f32c(x, w, dy, dx, dw, db, y_relu)
    cabi.check(cabi.load().psrc_small_linear_bwd(ptr(x), ptr(w), ptr(dy), ptr(y_relu), ptr(dx), int(accumulate_dx),
                                                 ptr(dw), ptr(db), G, R, N, K, cabi.stream()), "psrc_small_linear_bwd")


def group_logits_fwd(f_hat, t_hat, logits, scale):
    B, E = f_hat.shape
    Cn = logits.shape[1]
    assert t_hat.shape == (B * Cn, E) and logits.shape == (B, Cn)
    _f32c(f_hat, t_hat, logits)
    cabi.check(cabi.load().psrc_group_logits_fwd(ptr(f_hat), ptr(t_hat), ptr(logits), B, Cn, E, float(scale),
                                                 cabi.stream()), "psrc_group_logits_fwd")
    return logits


def group_logits_bwd(f_hat, d_logits, d_that, scale):
    B, E = f_hat.shape
    Cn = d_logits.shape[1]
    assert d_that.shape == (B * Cn, E) and d_logits.shape == (B, Cn)
    _f32c(f_hat, d_logits, d_that)
    cabi.check(cabi.load().psrc_group_logits_bwd(ptr(f_hat), ptr(d_logits), ptr(d_that), B, Cn, E, float(scale),
                                                 cabi.stream()), "psrc_group_logits_bwd")
    return d_that


def ctx_bias_grad(gg, dctx, dbias):
    G, n_ctx, D = gg.shape
    assert dctx.shape == (n_ctx, D) and dbias.shape == (G, D)
    _f32c(gg, dctx, dbias)
    cabi.check(cabi.load().psrc_ctx_bias_grad(ptr(gg), ptr(dctx), ptr(dbias), G, n_ctx, D, cabi.stream()),
               "psrc_ctx_bias_grad")


def permute_rows(src, perm, dst, n_seq, L, D, add=None, inverse=False):
    assert src.shape == dst.shape == (n_seq * L, D) and perm.shape == (n_seq, L) and perm.dtype == torch.int32
    _f32c(src, dst, add)
    cabi.check(cabi.load().psrc_permute_rows(ptr(src), ptr(perm), ptr(add), ptr(dst), n_seq, L, D, int(inverse),
                                             cabi.stream()), "psrc_permute_rows")
    return dst


def crop_resize_normalize(src_u8, geom_f, geom_i, out, interpolation, quantize, max_taps, mean, std):
    """uint8 [B, Hc, Wc, 3] canvases + per-image geometry -> normalised fp32 [B, 3, R, R] (psrc_crop_resize_normalize)."""
    B, Hc, Wc, ch = src_u8.shape
    R = out.shape[-1]
    assert src_u8.dtype == torch.uint8 and ch == 3 and src_u8.is_contiguous()
    assert out.shape == (B, 3, R, R) and out.dtype == torch.float32 and out.is_contiguous()
    assert geom_f.shape == (B, 4) and geom_f.dtype == torch.float32 and geom_f.is_contiguous()
    assert geom_i.shape == (B, 5) and geom_i.dtype == torch.int32 and geom_i.is_contiguous()
    m3, s3 = (C.c_float * 3)(*mean), (C.c_float * 3)(*std)
    cabi.check(cabi.load().psrc_crop_resize_normalize(ptr(src_u8), Hc * Wc * 3, Wc * 3, ptr(geom_f), ptr(geom_i),
                                                      ptr(out), B, R, int(interpolation), int(quantize),
                                                      int(max_taps), m3, s3, cabi.stream()),
               "psrc_crop_resize_normalize")
    return out
