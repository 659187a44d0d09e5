"""Per-kernel numerics checks against plain PyTorch fp32 on the same device.

Used two ways: `tests/test_kernels_gpu.py` (pytest -m gpu) asserts every check, and
`python tests/kernel_checks.py [--only NAME]` prints a table (each group in its own
subprocess so a CUDA fault in one kernel cannot hide the others).
"""
from __future__ import annotations

import math
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


# the PyTorch side is the fp32 reference: no TF32 in its conv / matmul
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.detach().float(), b.detach().float()
    return float((a - b).norm() / (b.norm() + 1e-30))


def max_err(a, b) -> float:
    return float((a.detach().float() - b.detach().float()).abs().max())


def quickgelu(x):
    return x * torch.sigmoid(1.702 * x)


def quickgelu_grad(x):
    s = torch.sigmoid(1.702 * x)
    return s * (1 + 1.702 * x * (1 - s))


class Results:
    def __init__(self):
        self.rows = []

    def add(self, name, err, tol):
        ok = (err == err) and err <= tol
        self.rows.append((name, err, tol, ok))
        print(f"{'PASS' if ok else 'FAIL'}  {name:58s} err={err:.3e} tol={tol:.1e}", flush=True)

    @property
    def ok(self):
        return all(r[3] for r in self.rows)


# ----------------------------------------------------------------------------- GEMM
def check_gemm(res: Results, backend_name: str):
    from psrc_b200 import ops
    dev = "cuda"
    backend = {"simt": ops.GEMM_SIMT, "tc": ops.GEMM_TCGEN05}[backend_name]
    g = torch.Generator(device=dev).manual_seed(0)
    shapes = [(256, 256, 64), (128, 128, 128), (300, 384, 768), (1000, 200, 72), (4, 512, 768),
              (2048, 768, 3072), (16384, 2304, 768)]
    if backend_name == "simt":
        shapes = shapes[:5] + [(1024, 768, 512)]
    for (M, N, K) in shapes:
        dts = [torch.bfloat16] if backend_name == "tc" else [torch.bfloat16, torch.float32]
        for dtp in dts:
            a = (torch.randn(M, K, device=dev, generator=g) * 0.5).to(dtp)
            w = (torch.randn(N, K, device=dev, generator=g) * 0.5).to(dtp)
            bias = torch.randn(N, device=dev, generator=g)
            resid = torch.randn(M, N, device=dev, generator=g)
            ref0 = a.float() @ w.float().t()
            tag = f"gemm[{backend_name},{str(dtp)[6:]}] {M}x{N}x{K}"
            # plain, fp32 out
            out = torch.full((M, N), float("nan"), device=dev)
            ops.gemm(a, w, out, backend=backend)
            res.add(tag + " plain/f32", rel_err(out, ref0), 2e-3)
            # bias + bf16 out
            outb = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
            ops.gemm(a, w, outb, bias=bias, backend=backend)
            res.add(tag + " bias/bf16", rel_err(outb, ref0 + bias), 6e-3)
            # bias + residual, in place on the fp32 residual stream, alpha
            rs = resid.clone()
            ops.gemm(a, w, rs, bias=bias, residual=rs, alpha=0.5, backend=backend)
            res.add(tag + " bias+res inplace", rel_err(rs, 0.5 * ref0 + bias + resid), 2e-3)
            # the same, out of place (the differentiated towers keep every layer's residual stream)
            ro = torch.full((M, N), float("nan"), device=dev)
            ops.gemm(a, w, ro, bias=bias, residual=resid, alpha=0.5, backend=backend)
            res.add(tag + " bias+res", rel_err(ro, 0.5 * ref0 + bias + resid), 2e-3)
            # quickgelu with pre-activation copy
            og = torch.empty(M, N, device=dev, dtype=dtp)
            pre = torch.empty(M, N, device=dev, dtype=dtp)
            ops.gemm(a, w, og, bias=bias, act=ops.ACT_QUICKGELU, out2=pre, backend=backend)
            res.add(tag + " quickgelu", rel_err(og, quickgelu(ref0 + bias)), 6e-3)
            res.add(tag + " quickgelu pre-act", rel_err(pre, ref0 + bias), 6e-3)
            # quickgelu grad
            u = torch.randn(M, N, device=dev, generator=g).to(dtp)
            od = torch.empty(M, N, device=dev, dtype=dtp)
            ops.gemm(a, w, od, act=ops.ACT_QUICKGELU_GRAD, aux=u, backend=backend)
            res.add(tag + " quickgelu_grad", rel_err(od, ref0 * quickgelu_grad(u.float())), 6e-3)
    # strided (non-contiguous rows) A: a view into a wider buffer
    M, N, K = 512, 256, 128
    big = (torch.randn(M, 3 * K, device=dev, generator=g)).to(torch.bfloat16)
    a = big[:, K:2 * K]
    w = torch.randn(N, K, device=dev, generator=g).to(torch.bfloat16)
    out = torch.empty(M, N, device=dev)
    ops.gemm(a, w, out, backend=backend)
    res.add(f"gemm[{backend_name}] strided A view", rel_err(out, a.float() @ w.float().t()), 2e-3)


def check_gemm_strided(res: Results):
    from psrc_b200 import ops
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(1)
    B, Cn, E = 37, 1000, 512
    f = torch.randn(B, E, device=dev, generator=g)
    t = torch.randn(Cn, E, device=dev, generator=g)
    logits = torch.empty(B, Cn, device=dev)
    ops.gemm_f32_strided(B, Cn, E, f, E, 1, t, E, 1, logits, alpha=14.3)
    res.add("gemm_f32_strided logits", rel_err(logits, 14.3 * f @ t.t()), 1e-5)
    dl = torch.randn(B, Cn, device=dev, generator=g)
    df = torch.randn(B, E, device=dev, generator=g)
    ref = df + 2.0 * dl @ t
    ops.gemm_f32_strided(B, E, Cn, dl, Cn, 1, t, 1, E, df, alpha=2.0, beta=1.0)
    res.add("gemm_f32_strided dF (beta=1)", rel_err(df, ref), 1e-5)
    dtt = torch.empty(Cn, E, device=dev)
    ops.gemm_f32_strided(Cn, E, B, dl, 1, Cn, f, 1, E, dtt)
    res.add("gemm_f32_strided dT (A transposed)", rel_err(dtt, dl.t() @ f), 1e-5)


# ----------------------------------------------------------------------------- LayerNorm & co
def check_layernorm(res: Results):
    from psrc_b200 import ops
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(2)
    for D in (128, 512, 768, 1024):
        rows = 1003
        x = torch.randn(rows, D, device=dev, generator=g) * 2 + 0.3
        gam = torch.randn(D, device=dev, generator=g)
        bet = torch.randn(D, device=dev, generator=g)
        ref = F.layer_norm(x, (D,), gam, bet, 1e-5)
        for dtp in (torch.float32, torch.bfloat16):
            y = torch.empty(rows, D, device=dev, dtype=dtp)
            mean = torch.empty(rows, device=dev)
            rstd = torch.empty(rows, device=dev)
            ops.layernorm_fwd(x, gam, bet, y, mean, rstd)
            res.add(f"layernorm_fwd D={D} {str(dtp)[6:]}", rel_err(y, ref), 1e-5 if dtp == torch.float32 else 4e-3)
        res.add(f"layernorm_fwd D={D} mean", max_err(mean, x.mean(1)), 1e-5)
        res.add(f"layernorm_fwd D={D} rstd", rel_err(rstd, (x.var(1, unbiased=False) + 1e-5).rsqrt()), 1e-5)
        # backward (dgrad), with residual accumulate and act copy
        dy = torch.randn(rows, D, device=dev, generator=g)
        xr = x.clone().requires_grad_(True)
        F.layer_norm(xr, (D,), gam, bet, 1e-5).backward(dy)
        dx_in = torch.randn(rows, D, device=dev, generator=g)
        dx = dx_in.clone()
        dxa = torch.empty(rows, D, device=dev, dtype=torch.bfloat16)
        ops.layernorm_bwd(dy, x, gam, mean, rstd, dx, dx_in=dx, dx_act=dxa)
        res.add(f"layernorm_bwd D={D} f32 (+residual, in place)", rel_err(dx, xr.grad + dx_in), 1e-5)
        res.add(f"layernorm_bwd D={D} act copy", rel_err(dxa, xr.grad + dx_in), 4e-3)
        dyb = dy.to(torch.bfloat16)
        dx2 = torch.empty(rows, D, device=dev)
        ops.layernorm_bwd(dyb, x, gam, mean, rstd, dx2)
        xr.grad = None
        F.layer_norm(xr, (D,), gam, bet, 1e-5).backward(dyb.float())
        res.add(f"layernorm_bwd D={D} bf16 dy", rel_err(dx2, xr.grad), 1e-5)
    # gathered LN
    n_seq, L, D = 57, 13, 512
    x = torch.randn(n_seq * L, D, device=dev, generator=g)
    idx = torch.randint(0, L, (n_seq,), device=dev, generator=g, dtype=torch.int32)
    gam = torch.randn(D, device=dev, generator=g)
    bet = torch.randn(D, device=dev, generator=g)
    rows_sel = torch.arange(n_seq, device=dev) * L + idx.long()
    xs = x[rows_sel].clone().requires_grad_(True)
    ref = F.layer_norm(xs, (D,), gam, bet, 1e-5)
    y = torch.empty(n_seq, D, device=dev)
    mean = torch.empty(n_seq, device=dev)
    rstd = torch.empty(n_seq, device=dev)
    ops.gather_ln_fwd(x, idx, gam, bet, y, mean, rstd, n_seq, L, D)
    res.add("gather_ln_fwd (EOT rows)", rel_err(y, ref), 1e-5)
    dy = torch.randn(n_seq, D, device=dev, generator=g)
    ref.backward(dy)
    dx = torch.full((n_seq * L, D), float("nan"), device=dev)
    ops.gather_ln_bwd(dy, x, idx, gam, mean, rstd, dx, n_seq, L, D)
    ref_dx = torch.zeros(n_seq * L, D, device=dev)
    ref_dx[rows_sel] = xs.grad
    res.add("gather_ln_bwd (zeros elsewhere)", rel_err(dx, ref_dx), 1e-5)
    y0 = torch.empty(n_seq, D, device=dev)
    ops.gather_ln_fwd(x, None, gam, bet, y0, mean, rstd, n_seq, L, D)
    res.add("gather_ln_fwd (row 0)", rel_err(y0, F.layer_norm(x[0::L], (D,), gam, bet, 1e-5)), 1e-5)
    # l2norm
    x = torch.randn(101, 512, device=dev, generator=g)
    y = torch.empty_like(x)
    inv = torch.empty(101, device=dev)
    ops.l2norm_fwd(x, y, inv)
    xr = x.clone().requires_grad_(True)
    ref = xr / xr.norm(dim=-1, keepdim=True)
    res.add("l2norm_fwd", rel_err(y, ref), 1e-6)
    dy = torch.randn_like(x)
    ref.backward(dy)
    dx = torch.empty_like(x)
    ops.l2norm_bwd(dy, y, inv, dx)
    res.add("l2norm_bwd", rel_err(dx, xr.grad), 1e-5)


# ----------------------------------------------------------------------------- attention
def _attn_ref(qkv, n_seq, L, H, causal):
    D = H * 64
    q, k, v = qkv.float().view(n_seq, L, 3, H, 64).permute(2, 0, 3, 1, 4)  # [3][n,H,L,64]
    s = (q @ k.transpose(-1, -2)) / 8.0
    if causal:
        s = s + torch.full((L, L), float("-inf"), device=qkv.device).triu(1)
    p = s.softmax(-1)
    o = (p @ v).permute(0, 2, 1, 3).reshape(n_seq * L, D)
    return o, torch.logsumexp(s, -1)


def check_attention(res: Results, backend_name: str):
    from psrc_b200 import ops
    dev = "cuda"
    backend = {"simt": 1, "mma": 2, "tc": 3, "small_mma": 5}[backend_name]
    g = torch.Generator(device=dev).manual_seed(3)
    cases = [(3, 201, 12, False), (5, 77, 8, True), (7, 16, 8, True), (2, 197, 12, False), (4, 9, 2, True),
             (2, 64, 2, False), (1, 259, 4, False)]
    dts = [torch.bfloat16] if backend_name in ("mma", "tc", "small_mma") else [torch.float32, torch.bfloat16]
    if backend_name == "small_mma":  # one 16 x 16 score tile per warp: L <= 16
        cases = [(7, 16, 8, True), (4, 9, 2, True), (5, 11, 8, True), (250, 11, 8, True), (2, 1, 2, True),
                 (3, 16, 3, False), (5, 13, 1, False), (1001, 11, 8, True), (3, 8, 2, True),
                 (3, 32, 8, True), (3, 17, 2, False), (5, 25, 1, True), (130, 24, 8, True), (2, 31, 3, False),
                 (4, 20, 12, True)]
    if backend_name == "tc":  # tcgen05 kernels: 32 < L <= 256
        cases = [(3, 201, 12, False), (5, 77, 8, True), (2, 197, 12, False), (2, 64, 2, False), (3, 128, 2, True),
                 (2, 256, 2, False), (2, 129, 1, True), (37, 201, 12, False),
                 # chunk geometries of the pipelined backward (L > 128): 10 / 11 / 15 / 9 column groups
                 (2, 145, 2, False), (3, 176, 3, True), (2, 240, 1, False), (2, 130, 2, True)]
    for (n_seq, L, H, causal) in cases:
        D = H * 64
        for dtp in dts:
            qkv = (torch.randn(n_seq * L, 3 * D, device=dev, generator=g)).to(dtp)
            qr = qkv.float().clone().requires_grad_(True)
            o_ref, lse_ref = _attn_ref(qr, n_seq, L, H, causal)
            out = torch.full((n_seq * L, D), float("nan"), device=dev, dtype=dtp)
            lse = torch.full((n_seq, H, L), float("nan"), device=dev)
            ops.attn_fwd(qkv, out, lse, n_seq, L, H, causal, backend)
            tol = 1e-5 if dtp == torch.float32 else 8e-3
            tag = f"attn[{backend_name},{str(dtp)[6:]}] n={n_seq} L={L} H={H} causal={int(causal)}"
            res.add(tag + " out", rel_err(out, o_ref), tol)
            res.add(tag + " lse", max_err(lse, lse_ref), 1e-4 if dtp == torch.float32 else 2e-2)
            dout = (torch.randn(n_seq * L, D, device=dev, generator=g)).to(dtp)
            o_ref.backward(dout.float())
            dqkv = torch.full((n_seq * L, 3 * D), float("nan"), device=dev, dtype=dtp)
            # feed the kernel its own forward outputs, as the training step does
            ops.attn_bwd(qkv, out, dout, lse, dqkv, n_seq, L, H, causal, backend)
            res.add(tag + " dqkv", rel_err(dqkv, qr.grad), 1e-4 if dtp == torch.float32 else 2e-2)


# ----------------------------------------------------------------------------- tokens
def check_tokens(res: Results):
    from psrc_b200 import ops
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(4)
    # im2col vs conv2d
    for (B, R, P, W) in ((3, 224, 16, 64), (2, 224, 14, 32), (2, 64, 16, 128)):
        img = torch.randn(B, 3, R, R, device=dev, generator=g)
        wconv = torch.randn(W, 3, P, P, device=dev, generator=g) * 0.05
        G = R // P
        ref = F.conv2d(img, wconv, stride=P).reshape(B, W, G * G).permute(0, 2, 1).reshape(B * G * G, W)
        for dtp in (torch.float32, torch.bfloat16):
            patches = torch.empty(B * G * G, 3 * P * P, device=dev, dtype=dtp)
            ops.im2col_patches(img, patches, R, P)
            got = patches.float() @ wconv.reshape(W, -1).t()
            res.add(f"im2col B={B} R={R} P={P} {str(dtp)[6:]}", rel_err(got, ref), 2e-5 if dtp == torch.float32 else 6e-3)
    # vision token assembly + ln_pre
    B, npatch, n_ctx, D = 5, 196, 4, 768
    for pdt in (torch.float32, torch.bfloat16):
        pe = torch.randn(B * npatch, D, device=dev, generator=g).to(pdt)
        cls = torch.randn(D, device=dev, generator=g)
        pos = torch.randn(npatch + 1, D, device=dev, generator=g)
        vpt = torch.randn(n_ctx, D, device=dev, generator=g) * 0.02
        gam = torch.randn(D, device=dev, generator=g)
        bet = torch.randn(D, device=dev, generator=g)
        L = 1 + npatch + n_ctx
        x = torch.full((B * L, D), float("nan"), device=dev)
        ops.assemble_vision_tokens(pe, cls, pos, vpt, gam, bet, x, B, npatch, n_ctx, D)
        xr = torch.cat([cls.expand(B, 1, D), pe.float().view(B, npatch, D)], 1) + pos
        xr = torch.cat([xr, vpt.half().float().expand(B, n_ctx, D)], 1)
        ref = F.layer_norm(xr, (D,), gam, bet, 1e-5).reshape(B * L, D)
        res.add(f"assemble_vision_tokens pe={str(pdt)[6:]}", rel_err(x, ref), 1e-5)
    x0 = torch.full((B * (1 + npatch), D), float("nan"), device=dev)
    ops.assemble_vision_tokens(pe, cls, pos, None, gam, bet, x0, B, npatch, 0, D)
    ref0 = F.layer_norm(torch.cat([cls.expand(B, 1, D), pe.float().view(B, npatch, D)], 1) + pos, (D,), gam, bet, 1e-5)
    res.add("assemble_vision_tokens n_ctx=0 (frozen tower)", rel_err(x0, ref0.reshape(-1, D)), 1e-5)
    # text prompt assembly
    Cn, n_ctx, D, Lfull = 11, 4, 512, 77
    prefix = torch.randn(Cn, 1, D, device=dev, generator=g)
    suffix = torch.randn(Cn, Lfull - 1 - n_ctx, D, device=dev, generator=g)
    ctx = torch.randn(n_ctx, D, device=dev, generator=g)
    ctxc = torch.randn(Cn, n_ctx, D, device=dev, generator=g)
    pos = torch.randn(Lfull, D, device=dev, generator=g)
    for L in (77, 16):
        x = torch.full((Cn * L, D), float("nan"), device=dev)
        ops.assemble_text_prompts(prefix, suffix, ctx, pos, x, Cn, L, n_ctx, D)
        ref = (torch.cat([prefix, ctx.expand(Cn, n_ctx, D), suffix], 1) + pos)[:, :L].reshape(Cn * L, D)
        res.add(f"assemble_text_prompts L={L}", max_err(x, ref), 0.0)
        ops.assemble_text_prompts(prefix, suffix, ctxc, pos, x, Cn, L, n_ctx, D, ctx_per_class=True)
        ref = (torch.cat([prefix, ctxc, suffix], 1) + pos)[:, :L].reshape(Cn * L, D)
        res.add(f"assemble_text_prompts L={L} class-specific ctx", max_err(x, ref), 0.0)
    G_ = 3
    bias = torch.randn(G_, D, device=dev, generator=g)
    L = 16
    x = torch.full((G_ * Cn * L, D), float("nan"), device=dev)
    ops.assemble_text_prompts(prefix, suffix, ctx, pos, x, Cn, L, n_ctx, D, ctx_bias=bias, n_groups=G_)
    ref = torch.stack([(torch.cat([prefix, (ctx + bias[i]).expand(Cn, n_ctx, D), suffix], 1) + pos)[:, :L] for i in range(G_)])
    res.add("assemble_text_prompts CoCoOp groups", max_err(x, ref.reshape(-1, D)), 1e-6)
    # insert prompt / prompt grad
    n_seq, L, D, n_ctx = 9, 21, 768, 4
    for start in (1, L - n_ctx):
        x = torch.randn(n_seq * L, D, device=dev, generator=g)
        pr = torch.randn(n_ctx, D, device=dev, generator=g) * 0.02
        ref = x.clone().view(n_seq, L, D)
        ref[:, start:start + n_ctx] = pr.half().float()
        ops.insert_prompt(x, pr, n_seq, L, start, n_ctx, D)
        res.add(f"insert_prompt start={start}", max_err(x, ref.reshape(-1, D)), 0.0)
        dx = torch.randn(n_seq * L, D, device=dev, generator=g)
        gref = dx.view(n_seq, L, D)[:, start:start + n_ctx].sum(0)
        dref = dx.clone().view(n_seq, L, D)
        dref[:, start:start + n_ctx] = 0
        grad = torch.full((n_ctx, D), float("nan"), device=dev)
        ops.prompt_grad(dx, grad, n_seq, L, start, n_ctx, D)
        res.add(f"prompt_grad start={start} grad", rel_err(grad, gref), 1e-6)
        res.add(f"prompt_grad start={start} rows zeroed", max_err(dx, dref.reshape(-1, D)), 0.0)
    # cast + sgd
    src = torch.randn(1000003, device=dev, generator=g)
    dst = torch.empty(1000003, device=dev, dtype=torch.bfloat16)
    ops.cast(src, dst)
    res.add("cast f32->bf16 (odd n)", max_err(dst, src.to(torch.bfloat16)), 0.0)
    src4 = torch.randn(4096, device=dev, generator=g)
    dst4 = torch.empty(4096, device=dev, dtype=torch.bfloat16)
    ops.cast(src4, dst4)
    res.add("cast f32->bf16 (vec4)", max_err(dst4, src4.to(torch.bfloat16)), 0.0)
    p = torch.randn(46080, device=dev, generator=g)
    pt = p.clone().requires_grad_(True)
    opt = torch.optim.SGD([pt], lr=0.0025, momentum=0.9, weight_decay=5e-4)
    buf = torch.zeros_like(p)
    lr = torch.tensor([0.0025], device=dev)
    for step in range(3):
        gr = torch.randn(46080, device=dev, generator=g)
        pt.grad = gr.clone()
        opt.step()
        ops.sgd_step(p, gr, buf, lr, 0.9, 0.0, 5e-4, False, step == 0)
    res.add("sgd_step vs torch.optim.SGD (3 steps)", max_err(p, pt.detach()), 1e-6)


# ----------------------------------------------------------------------------- loss
def check_loss(res: Results):
    from psrc_b200 import ops
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(5)
    for (B, Cn, E, focal) in ((4, 100, 512, False), (128, 1000, 512, False), (32, 100, 512, True), (7, 37, 128, True)):
        logits = (torch.randn(B, Cn, device=dev, generator=g) * 2).requires_grad_(True)
        zs = torch.randn(B, Cn, device=dev, generator=g) * 2
        labels = torch.randint(0, Cn, (B,), device=dev, generator=g)
        that = F.normalize(torch.randn(Cn, E, device=dev, generator=g), dim=-1).requires_grad_(True)
        tzs = F.normalize(torch.randn(Cn, E, device=dev, generator=g), dim=-1)
        fhat = F.normalize(torch.randn(B, E, device=dev, generator=g), dim=-1).requires_grad_(True)
        fzs = F.normalize(torch.randn(B, E, device=dev, generator=g), dim=-1)
        alpha = (torch.rand(Cn, device=dev, generator=g) * 3) if focal else None
        if focal:
            ce = F.cross_entropy(logits, labels, reduction="none")
            ptv = torch.exp(-ce)
            l_ce = (alpha[labels] * (1 - ptv) ** 2 * ce).mean()
        else:
            l_ce = F.cross_entropy(logits, labels)
        l_t = F.l1_loss(that, tzs, reduction="mean") * 25.0
        l_i = F.l1_loss(fhat, fzs, reduction="mean") * 10.0
        l_kl = F.kl_div(F.log_softmax(logits, 1), F.log_softmax(zs, 1), reduction="sum", log_target=True) / logits.numel()
        (l_ce + l_t + l_i + l_kl).backward()
        d_logits = torch.empty(B, Cn, device=dev)
        d_that = torch.empty(Cn, E, device=dev)
        d_fhat = torch.empty(B, E, device=dev)
        losses = torch.empty(4, device=dev)
        scratch = torch.empty(4 * B + 2 * 296, device=dev)
        ops.loss_fwd_bwd(logits.detach(), labels.int(), d_logits, losses, scratch, zs_logits=zs, focal_alpha=alpha,
                         t_hat=that.detach(), t_zs=tzs, d_that=d_that, f_hat=fhat.detach(), f_zs=fzs, d_fhat=d_fhat)
        tag = f"loss B={B} C={Cn} focal={int(focal)}"
        ref_l = torch.stack([l_ce, l_kl, l_t, l_i]).detach()
        res.add(tag + " values", float(((losses - ref_l).abs() / (ref_l.abs() + 1e-6)).max()), 2e-5)
        res.add(tag + " d_logits", rel_err(d_logits, logits.grad), 2e-5)
        res.add(tag + " d_that", rel_err(d_that, that.grad), 1e-6)
        res.add(tag + " d_fhat", rel_err(d_fhat, fhat.grad), 1e-6)
    # CE only (CoOp): no zs / no L1
    B, Cn = 32, 100
    logits = torch.randn(B, Cn, device=dev, generator=g).requires_grad_(True)
    labels = torch.randint(0, Cn, (B,), device=dev, generator=g)
    F.cross_entropy(logits, labels).backward()
    d_logits = torch.empty(B, Cn, device=dev)
    losses = torch.empty(4, device=dev)
    scratch = torch.empty(4 * B + 2 * 296, device=dev)
    ops.loss_fwd_bwd(logits.detach(), labels.int(), d_logits, losses, scratch)
    res.add("loss CE only d_logits", rel_err(d_logits, logits.grad), 2e-5)
    res.add("loss CE only value", abs(float(losses[0]) - float(F.cross_entropy(logits, labels).detach())), 1e-5)


# ----------------------------------------------------------------------------- MaPLe / CoCoOp small layers
def check_small_layers(res: Results):
    from psrc_b200 import ops
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(5)
    rn = lambda *s: torch.randn(*s, device=dev, generator=g)  # noqa: E731
    for tag, (G, R, N, K, relu) in {"maple coupling 9x[2,512]->768": (9, 2, 768, 512, False),
                                     "cocoop linear1 [128,512]->32 relu": (1, 128, 32, 512, True),
                                     "cocoop linear2 [128,32]->512": (1, 128, 512, 32, False),
                                     "ragged [3,5,50]->37": (3, 5, 37, 50, True)}.items():
        x, w, b = rn(G, R, K), rn(G, N, K) * 0.1, rn(G, N)
        y = torch.empty(G, R, N, device=dev)
        ops.small_linear_fwd(x, w, b, y, relu=relu)
        ref = torch.einsum("grk,gnk->grn", x, w) + b[:, None]
        ref = ref.clamp_min(0) if relu else ref
        res.add(f"small_linear_fwd {tag}", rel_err(y, ref), 1e-5)
        dy = rn(G, R, N)
        d = dy * (ref > 0) if relu else dy
        dx0 = rn(G, R, K)
        dx, dw, db = dx0.clone(), torch.empty_like(w), torch.empty_like(b)
        ops.small_linear_bwd(x, w, dy, dx=dx, dw=dw, db=db, y_relu=y if relu else None, accumulate_dx=True)
        res.add(f"small_linear_bwd dX(+=) {tag}", rel_err(dx, dx0 + torch.einsum("grn,gnk->grk", d, w)), 1e-5)
        res.add(f"small_linear_bwd dW {tag}", rel_err(dw, torch.einsum("grn,grk->gnk", d, x)), 1e-5)
        res.add(f"small_linear_bwd db {tag}", rel_err(db, d.sum(1)), 1e-5)
    B, Cn, E = 7, 365, 512
    f, t = rn(B, E), rn(B * Cn, E)
    logits = torch.empty(B, Cn, device=dev)
    ops.group_logits_fwd(f, t, logits, 14.3)
    res.add("group_logits_fwd", rel_err(logits, 14.3 * torch.einsum("be,bce->bc", f, t.view(B, Cn, E))), 1e-5)
    dl, dt = rn(B, Cn), torch.empty(B * Cn, E, device=dev)
    ops.group_logits_bwd(f, dl, dt, 14.3)
    res.add("group_logits_bwd", rel_err(dt.view(B, Cn, E), 14.3 * dl[:, :, None] * f[:, None, :]), 1e-6)
    gg = rn(B, 4, E)
    dctx, dbias = torch.empty(4, E, device=dev), torch.empty(B, E, device=dev)
    ops.ctx_bias_grad(gg, dctx, dbias)
    res.add("ctx_bias_grad dctx", rel_err(dctx, gg.sum(0)), 1e-6)
    res.add("ctx_bias_grad dbias", rel_err(dbias, gg.sum(1)), 1e-6)
    # grouped context assembly (cocoop.py:185-198): sequence s = g*C + c gets ctx + bias[g]
    C2, L, n_ctx, D = 6, 9, 4, 64
    prefix, suffix, ctx, pos, bias = rn(C2, 1, D), rn(C2, 12, D), rn(n_ctx, D), rn(77, D), rn(B, D)
    x = torch.empty(B * C2 * L, D, device=dev)
    ops.assemble_text_prompts(prefix, suffix, ctx, pos, x, C2, L, n_ctx, D, ctx_bias=bias, n_groups=B)
    ref = torch.stack([torch.cat([prefix, (ctx + bias[i]).expand(C2, n_ctx, D), suffix], 1)[:, :L] + pos[:L]
                       for i in range(B)])
    res.add("assemble_text_prompts grouped (+bias)", max_err(x.view(B, C2, L, D), ref), 1e-6)


# ----------------------------------------------------------------------------- fp16 forward operands
def check_fp16_forward(res: Results):
    """The "fp16 forward" storage formats (engine.TowerDtypes): fp16 A / W / 16-bit outputs of the tcgen05 GEMMs (and a
    bf16 output from fp16 operands, the vision towers' q/k/v), fp16 aux of the QuickGELU' epilogue next to bf16
    operands, fp16 LayerNorm / im2col outputs, and the short-sequence attention kernels with fp16 q/k/v/O forward and
    fp16 q/k/v + bf16 dO -> bf16 dqkv backward.  Tolerances are the fp16 ones (3 more mantissa bits than bf16)."""
    from psrc_b200 import ops
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(11)
    h, b16 = torch.float16, torch.bfloat16
    for backend_name, backend in (("pair", ops.GEMM_TCGEN05), ("1cta", ops.GEMM_TCGEN05_1CTA), ("simt", ops.GEMM_SIMT)):
        shapes = [(300, 384, 768), (11000, 512, 512), (2048, 768, 3072)] if backend_name != "simt" else [(300, 384, 72)]
        for (M, N, K) in shapes:
            a = (torch.randn(M, K, device=dev, generator=g) * 0.5).to(h)
            w = (torch.randn(N, K, device=dev, generator=g) * 0.5).to(h)
            bias = torch.randn(N, device=dev, generator=g)
            resid = torch.randn(M, N, device=dev, generator=g)
            ref0 = a.float() @ w.float().t()
            tag = f"gemm[{backend_name},fp16] {M}x{N}x{K}"
            out = torch.full((M, N), float("nan"), device=dev)
            ops.gemm(a, w, out, backend=backend)
            res.add(tag + " plain/f32", rel_err(out, ref0), 1e-5)          # exact products, fp32 accumulation
            for odt, tol in ((h, 6e-4), (b16, 6e-3)):
                o16 = torch.empty(M, N, device=dev, dtype=odt)
                ops.gemm(a, w, o16, bias=bias, backend=backend)
                res.add(tag + f" bias/{str(odt)[6:]}", rel_err(o16, ref0 + bias), tol)
            ro = torch.full((M, N), float("nan"), device=dev)
            ops.gemm(a, w, ro, bias=bias, residual=resid, alpha=0.5, backend=backend)
            res.add(tag + " bias+res", rel_err(ro, 0.5 * ref0 + bias + resid), 1e-5)
            rs = resid.clone()
            ops.gemm(a, w, rs, bias=bias, residual=rs, backend=backend)
            res.add(tag + " bias+res inplace", rel_err(rs, ref0 + bias + resid), 1e-5)
            og, pre = torch.empty(M, N, device=dev, dtype=h), torch.empty(M, N, device=dev, dtype=h)
            ops.gemm(a, w, og, bias=bias, act=ops.ACT_QUICKGELU, out2=pre, backend=backend)
            res.add(tag + " quickgelu", rel_err(og, quickgelu(ref0 + bias)), 2e-3)   # tanh.approx in the fast form
            res.add(tag + " quickgelu pre-act", rel_err(pre, ref0 + bias), 6e-4)
            # backward flavour: bf16 operands, fp16 saved pre-activation
            ab, wb = a.to(b16), w.to(b16)
            u = torch.randn(M, N, device=dev, generator=g).to(h)
            od = torch.empty(M, N, device=dev, dtype=b16)
            ops.gemm(ab, wb, od, act=ops.ACT_QUICKGELU_GRAD, aux=u, backend=backend)
            res.add(tag + " quickgelu_grad (bf16 x bf16, fp16 aux)",
                    rel_err(od, (ab.float() @ wb.float().t()) * quickgelu_grad(u.float())), 6e-3)
    # LayerNorm / gathered LayerNorm / im2col / patch embeddings in fp16
    rows, D = 1003, 768
    x = torch.randn(rows, D, device=dev, generator=g) * 2 + 0.3
    gam, bet = torch.randn(D, device=dev, generator=g), torch.randn(D, device=dev, generator=g)
    y = torch.empty(rows, D, device=dev, dtype=h)
    ops.layernorm_fwd(x, gam, bet, y, torch.empty(rows, device=dev), torch.empty(rows, device=dev))
    res.add("layernorm_fwd fp16 out", rel_err(y, F.layer_norm(x, (D,), gam, bet, 1e-5)), 5e-4)
    n_seq, L = 59, 17
    yg = torch.empty(n_seq, D, device=dev, dtype=h)
    ops.gather_ln_fwd(x[:n_seq * L], None, gam, bet, yg, None, None, n_seq, L, D)
    res.add("gather_ln_fwd fp16 out", rel_err(yg, F.layer_norm(x[:n_seq * L][0::L], (D,), gam, bet, 1e-5)), 5e-4)
    for (B, R, P) in ((3, 224, 16), (2, 224, 14)):
        img = torch.randn(B, 3, R, R, device=dev, generator=g)
        G = R // P
        patches = torch.empty(B * G * G, 3 * P * P, device=dev, dtype=h)
        ops.im2col_patches(img, patches, R, P)
        ref = F.unfold(img, P, stride=P).transpose(1, 2).reshape(B * G * G, 3 * P * P)
        res.add(f"im2col fp16 B={B} R={R} P={P}", max_err(patches, ref.to(h)), 0.0)
    B, npatch, n_ctx = 3, 196, 4
    pe = torch.randn(B * npatch, D, device=dev, generator=g).to(h)
    cls, pos = torch.randn(D, device=dev, generator=g), torch.randn(npatch + 1, D, device=dev, generator=g)
    vpt = torch.randn(n_ctx, D, device=dev, generator=g) * 0.02
    xo = torch.full((B * (1 + npatch + n_ctx), D), float("nan"), device=dev)
    ops.assemble_vision_tokens(pe, cls, pos, vpt, gam, bet, xo, B, npatch, n_ctx, D)
    xr = torch.cat([cls.expand(B, 1, D), pe.float().view(B, npatch, D)], 1) + pos
    xr = torch.cat([xr, vpt.half().float().expand(B, n_ctx, D)], 1)
    res.add("assemble_vision_tokens pe=fp16", rel_err(xo, F.layer_norm(xr, (D,), gam, bet, 1e-5).reshape(-1, D)), 1e-5)
    # short-sequence attention: fp16 forward operands, bf16 gradients
    for (n_seq, L, H, causal) in [(7, 16, 8, True), (4, 9, 2, True), (1001, 11, 8, True), (2, 1, 2, True),
                                  (3, 16, 3, False), (3, 32, 8, True), (3, 17, 2, False), (130, 24, 8, True),
                                  (4, 20, 12, True)]:
        Dm = H * 64
        qkv = torch.randn(n_seq * L, 3 * Dm, device=dev, generator=g).to(h)
        qr = qkv.float().clone().requires_grad_(True)
        o_ref, lse_ref = _attn_ref(qr, n_seq, L, H, causal)
        out = torch.full((n_seq * L, Dm), float("nan"), device=dev, dtype=h)
        lse = torch.full((n_seq, H, L), float("nan"), device=dev)
        ops.attn_fwd(qkv, out, lse, n_seq, L, H, causal)
        tag = f"attn[small_mma,fp16] n={n_seq} L={L} H={H} causal={int(causal)}"
        res.add(tag + " out", rel_err(out, o_ref), 1e-3)
        res.add(tag + " lse", max_err(lse, lse_ref), 2e-3)
        dout = torch.randn(n_seq * L, Dm, device=dev, generator=g).to(b16)
        o_ref.backward(dout.float())
        dqkv = torch.full((n_seq * L, 3 * Dm), float("nan"), device=dev, dtype=b16)
        ops.attn_bwd(qkv, out, dout, lse, dqkv, n_seq, L, H, causal)
        res.add(tag + " dqkv (bf16)", rel_err(dqkv, qr.grad), 2e-2)


def check_attention_row(res: Results):
    """One-query attention of the last block (psrc_attn_row_fwd / _bwd) against the full attention reference
    restricted to the query row: vision (row 0, all keys) and text (EOT row, causal), every storage format."""
    from psrc_b200 import ops
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(13)
    for (n_seq, L, H, causal, use_idx) in [(5, 201, 12, False, False), (3, 197, 12, False, False), (7, 11, 8, True, True),
                                           (4, 77, 8, True, True), (2, 259, 16, False, False), (3, 17, 2, False, False),
                                           (130, 24, 8, True, True), (2, 1, 2, True, True), (3, 33, 2, False, True)]:
        D = H * 64
        idx = torch.randint(0, L, (n_seq,), device=dev, generator=g, dtype=torch.int32) if use_idx else None
        rows = torch.arange(n_seq, device=dev) * L + (idx.long() if use_idx else 0)
        for dtp, gdt, tol_o, tol_g in ((torch.float32, torch.float32, 2e-6, 2e-5), (torch.bfloat16, torch.bfloat16, 6e-3, 1.2e-2),
                                       (torch.float16, torch.bfloat16, 1e-3, 6e-3)):
            qkv = torch.randn(n_seq * L, 3 * D, device=dev, generator=g).to(dtp)
            qr = qkv.float().clone().requires_grad_(True)
            o_ref, lse_ref = _attn_ref(qr, n_seq, L, H, causal)
            out = torch.full((n_seq * L, D), 7.0, device=dev, dtype=dtp)
            lse = torch.full((n_seq, H, L), 7.0, device=dev)
            ops.attn_row_fwd(qkv, out, lse, n_seq, L, H, causal, idx=idx)
            tag = f"attn_row[{str(dtp)[6:]}] n={n_seq} L={L} H={H} causal={int(causal)} idx={int(use_idx)}"
            res.add(tag + " out (query rows)", rel_err(out[rows], o_ref[rows]), tol_o)
            pos = rows - torch.arange(n_seq, device=dev) * L
            ar = torch.arange(n_seq, device=dev)
            res.add(tag + " lse (query rows)", max_err(lse[ar, :, pos], lse_ref[ar, :, pos]), 1e-4 if dtp == torch.float32 else 2e-2)
            mask = torch.ones(n_seq * L, dtype=torch.bool, device=dev)
            mask[rows] = False
            untouched = bool((out[mask].float() == 7.0).all()) if mask.any() else True
            res.add(tag + " other rows untouched", 0.0 if untouched else 1.0, 0.0)
            dout = torch.randn(n_seq * L, D, device=dev, generator=g).to(gdt)     # garbage outside the query rows
            d_ref = torch.zeros(n_seq * L, D, device=dev)
            d_ref[rows] = dout[rows].float()
            o_ref.backward(d_ref)
            dqkv = torch.full((n_seq * L, 3 * D), float("nan"), device=dev, dtype=gdt)
            ops.attn_row_bwd(qkv, dout, lse, dqkv, n_seq, L, H, causal, idx=idx)
            res.add(tag + " dqkv", rel_err(dqkv, qr.grad), tol_g)


GROUPS = {
    "attn_row": check_attention_row,
    "fp16_forward": check_fp16_forward,
    "small_layers": check_small_layers,
    "gemm_simt": lambda r: check_gemm(r, "simt"),
    "gemm_tc": lambda r: check_gemm(r, "tc"),
    "gemm_strided": check_gemm_strided,
    "layernorm": check_layernorm,
    "attn_simt": lambda r: check_attention(r, "simt"),
    "attn_mma": lambda r: check_attention(r, "mma"),
    "attn_tc": lambda r: check_attention(r, "tc"),
    "attn_small_mma": lambda r: check_attention(r, "small_mma"),
    "tokens": check_tokens,
    "loss": check_loss,
}


def run_group(name: str) -> Results:
    res = Results()
    GROUPS[name](res)
    torch.cuda.synchronize()
    return res


if __name__ == "__main__":
    import subprocess
    if "--only" in sys.argv:
        name = sys.argv[sys.argv.index("--only") + 1]
        r = run_group(name)
        sys.exit(0 if r.ok else 1)
    bad = []
    for name in GROUPS:
        print(f"===== {name}", flush=True)
        try:
            p = subprocess.run([sys.executable, os.path.abspath(__file__), "--only", name], timeout=600)
            if p.returncode != 0:
                bad.append(name)
        except subprocess.TimeoutExpired:
            print(f"TIMEOUT {name}", flush=True)
            bad.append(name)
    print("FAILED GROUPS:", bad if bad else "none", flush=True)
    sys.exit(1 if bad else 0)
