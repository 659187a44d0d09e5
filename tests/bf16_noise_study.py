"""CPU study (not a test): which bf16 rounding sites of the kernel pipeline cost prompt-gradient cosine.

Emulates the fast mode's roundings inside the ORACLE's vision tower with autograd functions (forward value
rounding R, gradient rounding G) and reports the worst vision prompt-gradient cosine against the plain fp32
oracle at cfg #1 (B=4, C=100).  `python tests/bf16_noise_study.py [site,site,...]` = sites switched OFF.
Used to decide which tensors the CUDA path keeps in fp32 (DESIGN.md "Precision").
"""
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import clip_oracle as O  # noqa: E402

SITES = ["W", "h1", "qkv", "P", "O", "h2", "u", "g", "dxa", "du", "dh2", "dO", "dS", "dqkv", "dh1",
         "text", "zs", "img", "pe", "hv", "Wp"]
ON = set(SITES)


def bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


class _R(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return bf(x)

    @staticmethod
    def backward(ctx, g):
        return g


class _G(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x.clone()

    @staticmethod
    def backward(ctx, g):
        return bf(g)


def R(x, site):
    return _R.apply(x) if site in ON else x


def G(x, site):
    return _G.apply(x) if site in ON else x


class _Gelu(torch.autograd.Function):
    @staticmethod
    def forward(ctx, u):
        ctx.save_for_backward(bf(u) if "u" in ON else u)
        y = u * torch.sigmoid(1.702 * u)
        return bf(y) if "g" in ON else y

    @staticmethod
    def backward(ctx, gy):
        (u,) = ctx.saved_tensors
        s = torch.sigmoid(1.702 * u)
        du = gy * s * (1 + 1.702 * u * (1 - s))
        return bf(du) if "du" in ON else du


def W(sd, k):
    return bf(sd[k]) if "W" in ON else sd[k]


def block(x, sd, p, H, causal):
    N, L, D = x.shape
    hd = D // H
    h = G(R(O.layer_norm(x, sd[p + "ln_1.weight"], sd[p + "ln_1.bias"]), "h1"), "dh1")
    qkv = G(R(h @ W(sd, p + "attn.in_proj_weight").t() + sd[p + "attn.in_proj_bias"], "qkv"), "dqkv")
    q, k, v = [t.view(N, L, H, hd).transpose(1, 2) for t in qkv.split(D, dim=-1)]
    s = G((q @ k.transpose(-1, -2)) / math.sqrt(hd), "dS")
    if causal:
        s = s + torch.full((L, L), float("-inf")).triu(1)
    a = (R(s.softmax(-1), "P") @ v).transpose(1, 2).reshape(N, L, D)
    a = G(R(a, "O"), "dO")
    x = x + G(a @ W(sd, p + "attn.out_proj.weight").t() + sd[p + "attn.out_proj.bias"], "dxa")
    h = G(R(O.layer_norm(x, sd[p + "ln_2.weight"], sd[p + "ln_2.bias"]), "h2"), "dh2")
    g = _Gelu.apply(h @ W(sd, p + "mlp.c_fc.weight").t() + sd[p + "mlp.c_fc.bias"])
    return x + G(g @ W(sd, p + "mlp.c_proj.weight").t() + sd[p + "mlp.c_proj.bias"], "dxa")


def run(off):
    global ON
    from psrc_b200 import synthetic as S
    from helpers import cos, load_golden
    arch = S.ARCHS["ViT-B/16"]
    fx = load_golden("cfg1_promptsrc")
    sd = S.make_clip_state_dict(arch, seed=0)
    tokens = torch.as_tensor(fx["tokens"]).long()
    B, depth, n_ctx = fx["label"].shape[0], 9, 4
    image, label = S.synthetic_batch(B, tokens.shape[0], arch.image_resolution, seed=1)
    deep = S.make_deep_prompts(arch, n_ctx, n_ctx, depth, depth, seed=100)
    fixed = torch.from_numpy(fx["fixed_embeddings"])
    Lt = int(tokens.argmax(-1).max()) + 1

    def grads(emulate):
        ctx = sd["token_embedding.weight"][tokens[0, 1:1 + n_ctx]].clone().requires_grad_(True)
        vpt = deep["visual.VPT"].clone().requires_grad_(True)
        dv = {i: deep[f"visual.transformer.resblocks.{i}.VPT_shallow"].clone().requires_grad_(True) for i in range(1, depth)}
        dtx = {i: deep[f"transformer.resblocks.{i}.VPT_shallow"].clone().requires_grad_(True) for i in range(1, depth)}
        orig = O.block
        orig_conv, orig_ln = O.F.conv2d, O.layer_norm
        state = {"zs": False}
        if emulate:
            def pick(x, sd_, p, H, causal):
                if p.startswith("visual."):
                    use = ("zs" in ON) if not x.requires_grad else True
                else:
                    use = "text" in ON
                return (block if use else orig)(x, sd_, p, H, causal)
            O.block = pick

            def conv(img, w, stride):
                if "img" in ON:
                    img = bf(img)
                if "W" in ON:
                    w = bf(w)
                y = orig_conv(img, w, stride=stride)
                return bf(y) if "pe" in ON else y
            O.F.conv2d = conv

            def ln(x, w, b):
                y = orig_ln(x, w, b)
                if x.dim() == 2 and "hv" in ON:      # ln_post / (rows gathered) -> act dtype for the projection GEMM
                    y = R(y, "hv")
                return y
            O.layer_norm = ln
            if "Wp" in ON:
                sd_run = dict(sd)
                sd_run["visual.proj"] = bf(sd["visual.proj"])
                sd_run["text_projection"] = bf(sd["text_projection"])
            else:
                sd_run = sd
        else:
            sd_run = sd
        try:
            o = O.promptsrc_step(sd_run, tokens, image, label, ctx, vpt, dv, dtx, fixed, n_ctx, arch.vision_heads,
                                 arch.transformer_heads, truncate=Lt)
            o["loss"].backward()
        finally:
            O.block, O.F.conv2d, O.layer_norm = orig, orig_conv, orig_ln
        return {"vpt": vpt.grad, **{f"vision.{i}": t.grad for i, t in dv.items()}}, o["logits"].detach()

    ON = set()
    ref, ref_logits = grads(False)
    for off_set in off:
        ON = set(SITES) - set(off_set)
        g, lg = grads(True)
        cs = {k: cos(g[k], ref[k]) for k in ref}
        print(f"off={','.join(sorted(off_set)) or '-':40s} worst cos {min(cs.values()):.6f} mean {sum(cs.values()) / len(cs):.6f} "
              f"logits max|d| {float((lg - ref_logits).abs().max()):.2e}", flush=True)


if __name__ == "__main__":
    torch.set_num_threads(os.cpu_count() or 1)
    if len(sys.argv) > 1:
        run([set(a.split(",")) - {"-"} for a in sys.argv[1:]])
    else:
        run([set()] + [{s} for s in SITES])
