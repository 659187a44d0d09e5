"""Host-side orchestration of the prompted-CLIP towers over the C-ABI kernels.

Pure plumbing: this module owns device buffers (torch tensors) and the order of kernel
launches; every FLOP runs in ``csrc/``.  Buffers are allocated once per (shape, mode) so a
whole step can be captured into a CUDA graph.

Data layout in HBM: activations are batch-first ``[n_seq * L, D]``; the residual
stream is fp32; the 16-bit tensors of the fast mode follow ``TowerDtypes`` (fp16 forward operands,
bf16 gradients; everything fp32 in the validation mode); frozen weights are kept as ``W [out, in]``
(forward GEMM operand, K-major) plus ``W^T [in, out]`` (dgrad GEMM operand, gradient format) --
they are never given a weight-gradient pass
(PromptSRC/trainers/promptsrc.py:238-247).
"""
from __future__ import annotations

from typing import NamedTuple

import torch

from . import ops
from .ops import ACT_QUICKGELU, ACT_QUICKGELU_GRAD

ACT_DTYPES = {"bf16": torch.bfloat16, "fp32": torch.float32}


class TowerDtypes(NamedTuple):
    """Storage formats of one tower's 16-bit tensors in the fast mode (all fp32 in the validation mode).

    fwd:  forward GEMM operands -- LN outputs, MLP hidden / pre-activation, W_qkv, W_fc1, W_fc2
    attn: q/k/v, the attention output and W_o (the attention kernel's format)
    grad: every gradient tensor and the transposed weights of the dgrad GEMMs

    "fp16 forward" (DESIGN 2): forward operands in IEEE half -- CLIP's native inference format -- carry 3 more
    mantissa bits than bf16 at the same tensor-core rate; gradients keep bf16's range.  The tcgen05 attention kernels
    are bf16, so a tower whose sequences are longer than 32 tokens keeps bf16 q/k/v/O."""
    fwd: torch.dtype
    attn: torch.dtype
    grad: torch.dtype


def tower_dtypes(prec: str, L: int, fwd16: str = "fp16") -> TowerDtypes:
    if prec == "fp32":
        return TowerDtypes(torch.float32, torch.float32, torch.float32)
    assert prec == "bf16" and fwd16 in ("fp16", "bf16"), (prec, fwd16)
    if fwd16 == "bf16":
        return TowerDtypes(torch.bfloat16, torch.bfloat16, torch.bfloat16)
    return TowerDtypes(torch.float16, torch.float16 if L <= 32 else torch.bfloat16, torch.bfloat16)


def _dev(t, device, dtype=torch.float32):
    return t.detach().to(device=device, dtype=dtype).contiguous()


class BlockWeights:
    """One ResidualAttentionBlock (clip/model.py:167-188) in kernel layout."""

    def __init__(self, sd, p, device, dt: TowerDtypes, need_backward):
        f = lambda k: _dev(sd[p + k], device)            # noqa: E731  fp32 vectors
        a = lambda k: _dev(sd[p + k], device, dt.fwd)    # noqa: E731  forward GEMM operands
        self.ln1_g, self.ln1_b = f("ln_1.weight"), f("ln_1.bias")
        self.ln2_g, self.ln2_b = f("ln_2.weight"), f("ln_2.bias")
        self.wqkv, self.bqkv = a("attn.in_proj_weight"), f("attn.in_proj_bias")
        self.wo, self.bo = _dev(sd[p + "attn.out_proj.weight"], device, dt.attn), f("attn.out_proj.bias")
        self.w1, self.b1 = a("mlp.c_fc.weight"), f("mlp.c_fc.bias")
        self.w2, self.b2 = a("mlp.c_proj.weight"), f("mlp.c_proj.bias")
        if need_backward:
            t = lambda k: _dev(sd[p + k].t(), device, dt.grad)  # noqa: E731
            self.wqkv_t, self.wo_t = t("attn.in_proj_weight"), t("attn.out_proj.weight")
            self.w1_t, self.w2_t = t("mlp.c_fc.weight"), t("mlp.c_proj.weight")


class TowerWeights:
    def __init__(self, sd, prefix, n_layers, device, dt, need_backward):
        if not isinstance(dt, TowerDtypes):              # one dtype for everything (fp32 validation mode, tests)
            dt = TowerDtypes(dt, dt, dt)
        self.dt = dt
        self.blocks = [BlockWeights(sd, f"{prefix}resblocks.{i}.", device, dt, need_backward)
                       for i in range(n_layers)]
        self.need_backward = need_backward


ROW_ATTN_MAX_L = 288    # psrc_attn_row_fwd / _bwd: 32 lanes x 9 keys per lane (csrc/attention_row.cu:22)


class Tower:
    """A stack of residual attention blocks over ``n_seq`` sequences of length ``L``.

    prompt_layers: {layer i: (start, n_ctx)} -- rows [start, start+n_ctx) of every sequence are
    overwritten with that layer's deep prompt before the block runs (clip/model.py:232-252).
    """

    def __init__(self, weights: TowerWeights, n_seq, L, D, H, causal, save, device, prompt_layers=None):
        self.w, self.n_seq, self.L, self.D, self.H, self.causal = weights, n_seq, L, D, H, causal
        self.dt = dt = weights.dt            # storage formats follow the weights this tower multiplies with
        self.act, self.save, self.device = dt.grad, save, device
        self.prompt_layers = dict(prompt_layers or {})
        # towers whose consumer reads only row 0 of every sequence (ln_post on the class token, model.py:426) and whose
        # gradient therefore enters on row 0 only: the last block's out-projection, LN2 and MLP -- and their
        # input-gradients -- run on n_seq rows instead of n_seq * L (set by the owner)
        self.cls_only_last_mlp = False
        self.nl = len(weights.blocks)
        M = self.M = n_seq * L
        e = lambda *s, dt=torch.float32: torch.empty(*s, device=device, dtype=dt)  # noqa: E731
        n_saved = self.nl if save else 1
        # a forward-only tower updates ONE residual-stream buffer in place (psrc_gemm then adds through the TMA store and
        # no residual byte crosses the SM); a tower that is differentiated keeps every layer's LayerNorm inputs
        self.X = [e(M, D) for _ in range(self.nl + 1 if save else 1)]
        self.X1 = [e(M, D) for _ in range(n_saved)] if save else [self.X[0]]
        self.qkv = [e(M, 3 * D, dt=dt.attn) for _ in range(n_saved)]
        self.O = [e(M, D, dt=dt.attn) for _ in range(n_saved)]
        self.lse = [e(n_seq, H, L) for _ in range(n_saved)]
        self.u = [e(M, 4 * D, dt=dt.fwd) for _ in range(n_saved)]
        self.stats = [[e(M) for _ in range(4)] for _ in range(n_saved)]  # mean1, rstd1, mean2, rstd2
        self.h = e(M, D, dt=dt.fwd)
        self.g = e(M, 4 * D, dt=dt.fwd)
        self.hc = e(n_seq, D, dt=dt.fwd)         # class-token rows of the last block (cls_only_last_mlp)
        if save:
            assert weights.need_backward
            self.du = e(M, 4 * D, dt=dt.grad)
            self.dh = e(M, D, dt=dt.grad)
            self.dhc = e(n_seq, D, dt=dt.grad)       # class-token rows of the last block (cls_only_last_mlp)
            self.dqkv = e(M, 3 * D, dt=dt.grad)
            self.dx = e(M, D)
            self.dxa = e(M, D, dt=dt.grad) if dt.grad != torch.float32 else None

    def _row_only_last(self, i):
        """Block i is the last one and runs in the class-token-only form.  Sequences longer than the one-query attention
        kernels take (ViT-L/14@336px: 577 tokens) keep the ordinary full-row block -- same result on row 0, any L."""
        return self.cls_only_last_mlp and i == self.nl - 1 and self.L <= ROW_ATTN_MAX_L

    # -- buffers of layer i
    def _xin(self, i):
        return self.X[i] if self.save else self.X[0]

    def _xout(self, i):
        return self.X[i + 1] if self.save else self.X[0]

    @property
    def x0(self):
        return self.X[0]

    @property
    def x_final(self):
        return self.X[self.nl] if self.save else self.X[0]

    def forward(self, prompts=None):
        """prompts: {layer i: fp32 [n_ctx, D] tensor}.  Input must already be in ``self.x0``."""
        n_seq, L, D, H = self.n_seq, self.L, self.D, self.H
        for i, b in enumerate(self.w.blocks):
            s = i if self.save else 0
            x0, x1, x2 = self._xin(i), self.X1[s], self._xout(i)
            if i in self.prompt_layers:
                start, n_ctx = self.prompt_layers[i]
                ops.insert_prompt(x0, prompts[i], n_seq, L, start, n_ctx, D, round_fp16=True)
            mean1, rstd1, mean2, rstd2 = self.stats[s]
            ops.layernorm_fwd(x0, b.ln1_g, b.ln1_b, self.h, mean1, rstd1)
            ops.gemm(self.h, b.wqkv, self.qkv[s], bias=b.bqkv)
            if self._row_only_last(i):
                # Last block, consumer reads row 0 only: every token's K / V still feed the class-token query, so LN1
                # and the q/k/v projection run on all rows; the attention core runs for ONE query per (sequence,
                # head), and the out-projection, LN2 and the MLP on the n_seq class-token rows, addressed as strided
                # views (row stride L * D) of the full-size buffers.  Row-wise identical arithmetic: exact.  The other
                # rows of O / x1 / x2 are never written nor read.
                ops.attn_row_fwd(self.qkv[s], self.O[s], self.lse[s], n_seq, L, H, self.causal)
                oc, x0c, x1c, x2c = (t.view(n_seq, L, D)[:, 0] for t in (self.O[s], x0, x1, x2))
                gc = self.g[:n_seq]
                ops.gemm(oc, b.wo, x1c, bias=b.bo, residual=x0c)
                ops.gather_ln_fwd(x1, None, b.ln2_g, b.ln2_b, self.hc, mean2[:n_seq] if self.save else None,
                                  rstd2[:n_seq] if self.save else None, n_seq, L, D)
                ops.gemm(self.hc, b.w1, gc, bias=b.b1, act=ACT_QUICKGELU, out2=self.u[s][:n_seq] if self.save else None)
                ops.gemm(gc, b.w2, x2c, bias=b.b2, residual=x1c)
                continue
            ops.attn_fwd(self.qkv[s], self.O[s], self.lse[s], n_seq, L, H, self.causal)
            ops.gemm(self.O[s], b.wo, x1, bias=b.bo, residual=x0)
            ops.layernorm_fwd(x1, b.ln2_g, b.ln2_b, self.h, mean2, rstd2)
            ops.gemm(self.h, b.w1, self.g, bias=b.b1, act=ACT_QUICKGELU, out2=self.u[s] if self.save else None)
            ops.gemm(self.g, b.w2, x2, bias=b.b2, residual=x1)
        return self.x_final

    def backward(self, prompt_grads=None, stop_at_layer=0):
        """Input-gradient chain.  ``self.dx`` (fp32) and ``self.dxa`` (act copy; fp32 mode: dx itself)
        must hold d(loss)/d(x_final).  On return ``self.dx`` holds d(loss)/d(x0) (rows of replaced
        prompt tokens zeroed) and prompt_grads[i] ([n_ctx, D] fp32) are overwritten."""
        assert self.save
        n_seq, L, D, H = self.n_seq, self.L, self.D, self.H
        dx = self.dx
        dxa = self.dxa if self.dxa is not None else self.dx
        for i in range(self.nl - 1, stop_at_layer - 1, -1):
            b = self.w.blocks[i]
            mean1, rstd1, mean2, rstd2 = self.stats[i]
            if self._row_only_last(i):
                # the incoming gradient is non-zero on the class-token rows only (it came through gather_ln_bwd with
                # idx = None, which zeroes every other row of dx / dxa): the MLP branch and the out-projection are
                # differentiated on those n_seq rows; d(attention output) of every other row is zero
                dxac = dxa.view(n_seq, L, D)[:, 0]
                duc = self.du[:n_seq]
                ops.gemm(dxac, b.w2_t, duc, act=ACT_QUICKGELU_GRAD, aux=self.u[i][:n_seq])
                ops.gemm(duc, b.w1_t, self.dhc)
                ops.gather_ln_bwd(self.dhc, self.X1[i], None, b.ln2_g, mean2[:n_seq], rstd2[:n_seq], dx, n_seq, L, D,
                                  dx_act=self.dxa, dx_in=dx)
                # d(attention output) of the class-token rows; the one-query backward reads no other row of dh
                ops.gemm(dxac, b.wo_t, self.dh.view(n_seq, L, D)[:, 0])
                ops.attn_row_bwd(self.qkv[i], self.dh, self.lse[i], self.dqkv, n_seq, L, H, self.causal)
            else:
                # MLP branch: x2 = x1 + W2 . gelu(W1 . LN2(x1))
                ops.gemm(dxa, b.w2_t, self.du, act=ACT_QUICKGELU_GRAD, aux=self.u[i])
                ops.gemm(self.du, b.w1_t, self.dh)
                ops.layernorm_bwd(self.dh, self.X1[i], b.ln2_g, mean2, rstd2, dx, dx_in=dx, dx_act=self.dxa)
                # attention branch: x1 = x0 + Wo . attn(Wqkv . LN1(x0))
                ops.gemm(dxa, b.wo_t, self.dh)
                ops.attn_bwd(self.qkv[i], self.O[i], self.dh, self.lse[i], self.dqkv, n_seq, L, H, self.causal)
            ops.gemm(self.dqkv, b.wqkv_t, self.dh)
            ops.layernorm_bwd(self.dh, self.X[i], b.ln1_g, mean1, rstd1, dx, dx_in=dx, dx_act=self.dxa)
            if i in self.prompt_layers:
                start, n_ctx = self.prompt_layers[i]
                ops.prompt_grad(dx, prompt_grads[i], n_seq, L, start, n_ctx, D, zero_rows=True,
                                dx_act=self.dxa, round_fp16=True)
        return dx
