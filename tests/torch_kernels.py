"""TEST INFRASTRUCTURE -- plain-PyTorch (CPU-capable) restatement of every entry point of include/psrc_b200.h.

`TorchKernels.install(monkeypatch)` swaps `psrc_b200.ops.*` for these, so the PRODUCT's host orchestration
(engine.py / step.py / trainers.py: buffer plumbing, launch order, class sharding, collectives, loss scaling) runs
unchanged on CPU tensors -- single process against the golden fixtures, and world_size-2 `gloo` against the
single-process result.  Each function follows the contract text of the header, not the CUDA source.  Never
imported by the package; never used on a GPU box (the GPU tests call the real kernels through the C ABI).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

ACT_NONE, ACT_QUICKGELU, ACT_QUICKGELU_GRAD = 0, 1, 2
EPS = 1e-5


def _gelu(x):
    return x * torch.sigmoid(1.702 * x)


def _gelu_grad(x):
    s = torch.sigmoid(1.702 * x)
    return s * (1 + 1.702 * x * (1 - s))


def _attn(qkv, n_seq, L, H, causal):
    D = H * 64
    q, k, v = qkv.float().view(n_seq, L, 3, H, 64).permute(2, 0, 3, 1, 4)
    s = (q @ k.transpose(-1, -2)) / 8.0
    if causal:
        s = s + torch.full((L, L), float("-inf")).triu(1)
    o = (s.softmax(-1) @ v).permute(0, 2, 1, 3).reshape(n_seq * L, D)
    return o, torch.logsumexp(s, -1)


def _ln(x, g, b, eps):
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    rstd = (var + eps).rsqrt()
    return (x - mu) * rstd * g + b, mu.squeeze(-1), rstd.squeeze(-1)


def _ln_bwd(dy, x, g, mean, rstd):
    xh = (x - mean[:, None]) * rstd[:, None]
    gy = dy.float() * g
    return rstd[:, None] * (gy - gy.mean(-1, keepdim=True) - xh * (gy * xh).mean(-1, keepdim=True))


def _fp16(x):
    return x.to(torch.float16).to(torch.float32)


class TorchKernels:
    NAMES = ("gemm", "gemm_f32_strided", "layernorm_fwd", "layernorm_bwd", "attn_fwd", "attn_bwd", "attn_row_fwd",
             "attn_row_bwd", "im2col_patches",
             "assemble_vision_tokens", "assemble_text_prompts", "insert_prompt", "prompt_grad", "gather_ln_fwd",
             "gather_ln_bwd", "l2norm_fwd", "l2norm_bwd", "loss_fwd_bwd", "sgd_step", "cast", "axpy",
             "small_linear_fwd", "small_linear_bwd", "group_logits_fwd", "group_logits_bwd", "ctx_bias_grad", "permute_rows")

    def install(self, monkeypatch):
        from psrc_b200 import ops
        monkeypatch.setattr(ops, "_FAKE_KERNELS", True, raising=False)
        for n in self.NAMES:
            monkeypatch.setattr(ops, n, getattr(self, n))

    # ---------------------------------------------------------------- GEMM
    @staticmethod
    def gemm(a, w, out, bias=None, act=ACT_NONE, aux=None, residual=None, out2=None, alpha=1.0, backend=0):
        y = alpha * (a.float() @ w.float().t())
        if bias is not None:
            y = y + bias
        if act == ACT_QUICKGELU:
            if out2 is not None:
                out2.copy_(y)
            y = _gelu(y)
        elif act == ACT_QUICKGELU_GRAD:
            y = y * _gelu_grad(aux.float())
        if residual is not None:
            y = y + residual
        out.copy_(y)
        return out

    @staticmethod
    def gemm_f32_strided(M, N, K, a, a_sm, a_sk, b, b_sn, b_sk, c, alpha=1.0, beta=0.0):
        A = torch.as_strided(a, (M, K), (a_sm, a_sk), a.storage_offset())
        B = torch.as_strided(b, (N, K), (b_sn, b_sk), b.storage_offset())
        r = alpha * (A @ B.t())
        c.copy_(r + beta * c if beta != 0.0 else r)
        return c

    # ---------------------------------------------------------------- LayerNorm
    @staticmethod
    def layernorm_fwd(x, gamma, beta, y, mean=None, rstd=None, eps=EPS):
        v, mu, rs = _ln(x, gamma, beta, eps)
        y.copy_(v)
        if mean is not None:
            mean.copy_(mu)
        if rstd is not None:
            rstd.copy_(rs)
        return y

    @staticmethod
    def layernorm_bwd(dy, x, gamma, mean, rstd, dx_out, dx_in=None, dx_act=None):
        d = _ln_bwd(dy, x, gamma, mean, rstd)
        if dx_in is not None:
            d = d + dx_in
        dx_out.copy_(d)
        if dx_act is not None:
            dx_act.copy_(d)
        return dx_out

    # ---------------------------------------------------------------- attention
    @staticmethod
    def attn_fwd(qkv, out, lse, n_seq, L, H, causal, backend=0):
        o, l = _attn(qkv, n_seq, L, H, causal)
        out.copy_(o)
        lse.copy_(l)
        return out

    @staticmethod
    def attn_bwd(qkv, out, dout, lse, dqkv, n_seq, L, H, causal, backend=0):
        with torch.enable_grad():
            q = qkv.detach().float().clone().requires_grad_(True)
            o, _ = _attn(q, n_seq, L, H, causal)
            o.backward(dout.float())
        dqkv.copy_(q.grad)
        return dqkv

    @classmethod
    def attn_row_fwd(cls, qkv, out, lse, n_seq, L, H, causal, idx=None):
        """Contract of psrc_attn_row_fwd: only the query row of `out` / entry of `lse` is written."""
        o, l = _attn(qkv, n_seq, L, H, causal)
        r = cls._rows(idx, n_seq, L)
        out[r] = o[r].to(out.dtype)
        pos = (r - torch.arange(n_seq) * L)
        lse[torch.arange(n_seq), :, pos] = l[torch.arange(n_seq), :, pos]
        return out

    @classmethod
    def attn_row_bwd(cls, qkv, dout, lse, dqkv, n_seq, L, H, causal, idx=None):
        r = cls._rows(idx, n_seq, L)
        d = torch.zeros(n_seq * L, H * 64)
        d[r] = dout[r].float()                   # only the query row of dout is read
        with torch.enable_grad():
            q = qkv.detach().float().clone().requires_grad_(True)
            o, _ = _attn(q, n_seq, L, H, causal)
            o.backward(d)
        dqkv.copy_(q.grad)
        return dqkv

    # ---------------------------------------------------------------- tokens
    @staticmethod
    def im2col_patches(image, patches, R, P):
        B = image.shape[0]
        cols = F.unfold(image, kernel_size=P, stride=P)            # [B, 3*P*P, G*G], column = c*P*P + ky*P + kx
        patches.copy_(cols.transpose(1, 2).reshape(B * (R // P) ** 2, -1))
        return patches

    @staticmethod
    def assemble_vision_tokens(pe, cls, pos, vpt, ln_g, ln_b, x, B, n_patches, n_ctx, D, eps=EPS):
        t = torch.cat([cls.expand(B, 1, D), pe.float().view(B, n_patches, D)], 1) + pos
        if n_ctx > 0:
            t = torch.cat([t, _fp16(vpt).expand(B, n_ctx, D)], 1)
        x.copy_(_ln(t, ln_g, ln_b, eps)[0].reshape(-1, D))
        return x

    @staticmethod
    def assemble_text_prompts(prefix, suffix, ctx, pos, x, C_, L, n_ctx, D, ctx_per_class=False, ctx_bias=None,
                              n_groups=1):
        outs = []
        for g in range(n_groups):
            c = ctx if ctx_per_class else ctx.expand(C_, n_ctx, D)
            if ctx_bias is not None:
                c = c + ctx_bias[g]
            outs.append((torch.cat([prefix, c, suffix], 1)[:, :L] + pos[:L]))
        x.copy_(torch.cat(outs, 0).reshape(-1, D))
        return x

    @staticmethod
    def insert_prompt(x, prompt, n_seq, L, start, n_ctx, D, round_fp16=True):
        x.view(n_seq, L, D)[:, start:start + n_ctx] = _fp16(prompt) if round_fp16 else prompt
        return x

    @staticmethod
    def prompt_grad(dx, grad, n_seq, L, start, n_ctx, D, zero_rows=True, accumulate=False, dx_act=None,
                    round_fp16=False):
        rows = dx.view(n_seq, L, D)[:, start:start + n_ctx]
        s = (_fp16(rows) if round_fp16 else rows).sum(0)
        grad.copy_(grad + s if accumulate else s)
        if zero_rows:
            rows.zero_()
            if dx_act is not None:
                dx_act.view(n_seq, L, D)[:, start:start + n_ctx] = 0
        return grad

    # ---------------------------------------------------------------- heads
    @staticmethod
    def _rows(idx, n_seq, L):
        base = torch.arange(n_seq) * L
        return base if idx is None else base + idx.long()

    @classmethod
    def gather_ln_fwd(cls, x, idx, gamma, beta, y, mean, rstd, n_seq, L, D, eps=EPS):
        v, mu, rs = _ln(x[cls._rows(idx, n_seq, L)], gamma, beta, eps)
        y.copy_(v)
        if mean is not None:
            mean.copy_(mu)
        if rstd is not None:
            rstd.copy_(rs)
        return y

    @classmethod
    def gather_ln_bwd(cls, dy, x, idx, gamma, mean, rstd, dx, n_seq, L, D, dx_act=None, dx_in=None):
        r = cls._rows(idx, n_seq, L)
        d = _ln_bwd(dy, x[r], gamma, mean, rstd)
        if dx_in is not None:
            d = d + dx_in[r]
        dx.zero_()
        dx[r] = d
        if dx_act is not None:
            dx_act.copy_(dx)
        return dx

    @staticmethod
    def l2norm_fwd(x, y, inv_norm):
        inv = 1.0 / x.norm(dim=-1)
        y.copy_(x * inv[:, None])
        if inv_norm is not None:
            inv_norm.copy_(inv)
        return y

    @staticmethod
    def l2norm_bwd(dy, y, inv_norm, dx):
        dx.copy_((dy - y * (dy * y).sum(-1, keepdim=True)) * inv_norm[:, None])
        return dx

    # ---------------------------------------------------------------- loss
    @staticmethod
    def loss_fwd_bwd(logits, labels, d_logits, losses, scratch, zs_logits=None, focal_alpha=None, focal_gamma=2.0,
                     t_hat=None, t_zs=None, d_that=None, f_hat=None, f_zs=None, d_fhat=None, w_kl=1.0, w_text=25.0,
                     w_image=10.0, B_global=None, Ct_global=None):
        B, C = logits.shape
        Bg = B_global or B
        with torch.enable_grad():
            lg = logits.detach().clone().requires_grad_(True)
            ce = F.cross_entropy(lg, labels.long(), reduction="none")
            if focal_alpha is not None:
                pt = torch.exp(-ce)
                ce = focal_alpha[labels.long()] * (1 - pt) ** focal_gamma * ce
            l_ce = ce.sum() / Bg
            l_kl = torch.zeros(())
            if zs_logits is not None:
                l_kl = w_kl * F.kl_div(F.log_softmax(lg, 1), F.log_softmax(zs_logits, 1), reduction="sum",
                                       log_target=True) / (Bg * C)
            (l_ce + l_kl).backward()
        d_logits.copy_(lg.grad)
        l_t = l_i = torch.zeros(())
        if t_hat is not None and d_that is not None:
            sc = w_text / ((Ct_global or t_hat.shape[0]) * t_hat.shape[1])
            l_t = (t_hat - t_zs).abs().sum() * sc
            d_that.copy_(torch.sign(t_hat - t_zs) * sc)
        if f_hat is not None and d_fhat is not None:
            sc = w_image / (Bg * f_hat.shape[1])
            l_i = (f_hat - f_zs).abs().sum() * sc
            d_fhat.copy_(torch.sign(f_hat - f_zs) * sc)
        losses.copy_(torch.stack([l_ce.detach(), l_kl.detach(), l_t, l_i]))
        return losses

    # ---------------------------------------------------------------- optimizer / utilities
    @staticmethod
    def sgd_step(p, g, buf, lr_dev, momentum, dampening, weight_decay, nesterov, first_step):
        d = g + weight_decay * p
        if momentum != 0:
            b = d.clone() if first_step else momentum * buf + (1 - dampening) * d
            buf.copy_(b)
            d = d + momentum * b if nesterov else b
        p.sub_(lr_dev[0] * d)
        return p

    @staticmethod
    def cast(src, dst):
        dst.copy_(src)
        return dst

    @staticmethod
    def axpy(y, x, a=1.0):
        y.add_(x, alpha=a)
        return y

    # ---------------------------------------------------------------- trainable small layers (MaPLe, CoCoOp)
    @staticmethod
    def small_linear_fwd(x, w, b, y, relu=False):
        o = torch.einsum("grk,gnk->grn", x, w)
        if b is not None:
            o = o + b[:, None]
        y.copy_(o.clamp_min(0) if relu else o)
        return y

    @staticmethod
    def small_linear_bwd(x, w, dy, dx=None, dw=None, db=None, y_relu=None, accumulate_dx=False):
        d = dy if y_relu is None else dy * (y_relu > 0)
        if dw is not None:
            dw.copy_(torch.einsum("grn,grk->gnk", d, x))
            db.copy_(d.sum(1))
        if dx is not None:
            g = torch.einsum("grn,gnk->grk", d, w)
            dx.copy_(dx + g if accumulate_dx else g)

    @staticmethod
    def group_logits_fwd(f_hat, t_hat, logits, scale):
        B, E = f_hat.shape
        logits.copy_(scale * torch.einsum("be,bce->bc", f_hat, t_hat.view(B, -1, E)))
        return logits

    @staticmethod
    def group_logits_bwd(f_hat, d_logits, d_that, scale):
        B, E = f_hat.shape
        d_that.copy_((scale * d_logits[:, :, None] * f_hat[:, None, :]).reshape(-1, E))
        return d_that

    @staticmethod
    def ctx_bias_grad(gg, dctx, dbias):
        dctx.copy_(gg.sum(0))
        dbias.copy_(gg.sum(1))

    @staticmethod
    def permute_rows(src, perm, dst, n_seq, L, D, add=None, inverse=False):
        idx = (torch.arange(n_seq)[:, None] * L + perm.long()).reshape(-1)
        if inverse:
            dst[idx] = src
        else:
            dst.copy_(src[idx] + (add[:L].repeat(n_seq, 1) if add is not None else 0))
        return dst
