"""TEST INFRASTRUCTURE -- CPU restatement (plain PyTorch fp32 + autograd) of the reference hot path.

This is the oracle the CUDA path is checked against.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may import
it; the product never does.  Every function cites the reference lines it restates (paths
relative to the reference root, i.e. PromptSRC/...).

Pinning: `oracle/make_golden.py` runs the UNMODIFIED reference modules in the build container
(`oracle/ref_harness.py`, SURVEY.md section 0 recipe) on seeded synthetic weights/inputs, checks
this restatement against them and commits the reference outputs under `tests/golden/`;
`tests/test_oracle_golden.py` re-checks the restatement against those fixtures anywhere.
The reference itself has no tests or golden vectors, and its arithmetic
lives in PyTorch (pinned torch==1.9.0 in docs/INSTALL.md:17; executed here on torch 2.11).

Definition of "the reference PromptSRC path" (SURVEY.md 8c): fp32 model; prompt tokens
and `fixed_embeddings` (promptsrc.py:201 only) fp16 round-tripped; LOGITS_LOSS_WEIGHT=1.0.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

LN_EPS = 1e-5
# bench.py's torch-on-B200 leg only: route the attention core through F.scaled_dot_product_attention (the fused kernel
# nn.MultiheadAttention's fast path uses on torch 2.x); the default is the explicit softmax(QK^T)V restatement.
USE_SDPA = False


def fp16_rt(x: torch.Tensor) -> torch.Tensor:
    """`.half()` followed by promotion back to fp32 (clip/model.py:238,249,414; F2).
    Autograd casts the gradient through fp16 as well, exactly like the reference graph."""
    return x.to(torch.float16).to(torch.float32)


def layer_norm(x, w, b):
    """clip/model.py:153-159 -- fp32 LayerNorm, biased variance, eps 1e-5."""
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + LN_EPS) * w + b


def quick_gelu(x):
    """clip/model.py:162-164."""
    return x * torch.sigmoid(1.702 * x)


def attention(x, sd, p, n_head, causal):
    """nn.MultiheadAttention as called at clip/model.py:183/227/285 (packed in-proj, 64-dim heads,
    softmax(QK^T/sqrt(64) + mask) V, out-proj); mask = triu(-inf, 1) (model.py:592-598).
    x: [N, L, D] (batch-first restatement of the reference's [L, N, D])."""
    N, L, D = x.shape
    hd = D // n_head
    qkv = x @ sd[p + "attn.in_proj_weight"].t() + sd[p + "attn.in_proj_bias"]
    q, k, v = qkv.split(D, dim=-1)
    q = q.view(N, L, n_head, hd).transpose(1, 2)
    k = k.view(N, L, n_head, hd).transpose(1, 2)
    v = v.view(N, L, n_head, hd).transpose(1, 2)
    if USE_SDPA:
        a = F.scaled_dot_product_attention(q, k, v, is_causal=causal).transpose(1, 2).reshape(N, L, D)
        return a @ sd[p + "attn.out_proj.weight"].t() + sd[p + "attn.out_proj.bias"]
    s = (q @ k.transpose(-1, -2)) / math.sqrt(hd)
    if causal:
        s = s + torch.full((L, L), float("-inf"), device=x.device).triu(1)
    a = (s.softmax(-1) @ v).transpose(1, 2).reshape(N, L, D)
    return a @ sd[p + "attn.out_proj.weight"].t() + sd[p + "attn.out_proj.bias"]


def block(x, sd, p, n_head, causal):
    """ResidualAttentionBlock.forward, clip/model.py:186-188 (= :254-255, :329-330)."""
    x = x + attention(layer_norm(x, sd[p + "ln_1.weight"], sd[p + "ln_1.bias"]), sd, p, n_head, causal)
    h = layer_norm(x, sd[p + "ln_2.weight"], sd[p + "ln_2.bias"])
    h = quick_gelu(h @ sd[p + "mlp.c_fc.weight"].t() + sd[p + "mlp.c_fc.bias"])
    return x + h @ sd[p + "mlp.c_proj.weight"].t() + sd[p + "mlp.c_proj.bias"]


def vision_tower(image, sd, n_head, vpt=None, deep=None, half_prompts=True, prefix="visual."):
    """VisionTransformer.forward, clip/model.py:401-431, with ResidualAttentionBlock_IVLP prompt
    replacement (model.py:232-241): at layer i in `deep`, the last n_ctx tokens are replaced.
    vpt: [n_ctx, D] or None; deep: {layer index: [n_ctx, D]}."""
    rt = fp16_rt if half_prompts else (lambda t: t)
    w = sd[prefix + "conv1.weight"]
    x = F.conv2d(image, w, stride=w.shape[-1])                      # :402
    B, D = x.shape[0], x.shape[1]
    x = x.reshape(B, D, -1).permute(0, 2, 1)                        # :403-404
    x = torch.cat([sd[prefix + "class_embedding"].expand(B, 1, D), x], dim=1)   # :405-409
    x = x + sd[prefix + "positional_embedding"]                     # :410
    n_ctx = 0
    if vpt is not None:
        n_ctx = vpt.shape[0]
        x = torch.cat([x, rt(vpt.expand(B, -1, -1))], dim=1)        # :414-415
    x = layer_norm(x, sd[prefix + "ln_pre.weight"], sd[prefix + "ln_pre.bias"])  # :420
    layers = len([k for k in sd if k.startswith(prefix + "transformer.resblocks") and k.endswith("attn.in_proj_weight")])
    for i in range(layers):
        if deep is not None and i in deep:
            x = torch.cat([x[:, : x.shape[1] - n_ctx], rt(deep[i].expand(B, -1, -1))], dim=1)  # :236-241
        x = block(x, sd, f"{prefix}transformer.resblocks.{i}.", n_head, causal=False)
    x = layer_norm(x[:, 0], sd[prefix + "ln_post.weight"], sd[prefix + "ln_post.bias"])   # :426
    return x @ sd[prefix + "proj"]                                  # :428-429


def text_tower(prompts, eot_idx, sd, n_head, deep=None, n_ctx=0, half_prompts=True, truncate=None):
    """TextEncoder.forward, trainers/promptsrc.py:59-70 (= coop.py:195-205, cocoop.py:54-64), with
    the IVLP text prompt replacement (model.py:243-252): tokens 1..n_ctx replaced at layers in `deep`.
    prompts: [C, 77, D] token embeddings (without positional embedding).  truncate=L runs only the
    first L positions -- exact under the causal mask when L > max(eot_idx)."""
    rt = fp16_rt if half_prompts else (lambda t: t)
    x = prompts + sd["positional_embedding"]                       # :60
    if truncate is not None:
        assert int(eot_idx.max()) < truncate
        x = x[:, :truncate]
    C = x.shape[0]
    layers = len(set(k.split(".")[2] for k in sd if k.startswith("transformer.resblocks")))
    for i in range(layers):
        if deep is not None and i in deep:
            x = torch.cat([x[:, :1], rt(deep[i].expand(C, -1, -1)), x[:, 1 + n_ctx:]], dim=1)  # :246-252
        x = block(x, sd, f"transformer.resblocks.{i}.", n_head, causal=True)
    x = layer_norm(x, sd["ln_final.weight"], sd["ln_final.bias"])  # :64
    return x[torch.arange(C, device=x.device), eot_idx] @ sd["text_projection"]     # :68


def normalize(x):
    """`x / x.norm(dim=-1, keepdim=True)` (promptsrc.py:190-191)."""
    return x / x.norm(dim=-1, keepdim=True)


def focal_ce(logits, labels, alpha=None, gamma=2.0):
    """MultiClassFocalLoss.forward, trainers/coop.py:146-163 (reduction='mean')."""
    ce = F.cross_entropy(logits, labels, reduction="none")
    pt = torch.exp(-ce)
    a = alpha[labels] if alpha is not None else 1.0
    return (a * (1 - pt) ** gamma * ce).mean()


def construct_prompts(prefix, ctx, suffix):
    """VLPromptLearner.forward / construct_prompts, promptsrc.py:144-168 (coop.py:259-276 'end')."""
    if ctx.dim() == 2:
        ctx = ctx.unsqueeze(0).expand(prefix.shape[0], -1, -1)
    return torch.cat([prefix, ctx, suffix], dim=1)


def promptsrc_step(sd, tokens, image, label, ctx, vpt, deep_v, deep_t, fixed_embeddings, n_ctx_text,
                   vision_heads, text_heads, w_text=25.0, w_image=10.0, w_logits=1.0, focal_alpha=None,
                   truncate=None):
    """CustomCLIP.forward (promptsrc.py:183-211) + the loss of PromptSRC.forward_backward (:299-324).
    Returns a dict with every term; call `.backward()` on out["loss"] for the prompt gradients.
    focal_alpha != None swaps F.cross_entropy for the focal CE (promptsrc_pets_focal.py:230-247)."""
    emb = sd["token_embedding.weight"][tokens]                      # promptsrc.py:121
    prefix, suffix = emb[:, :1], emb[:, 1 + n_ctx_text:]            # :136-137
    eot = tokens.argmax(-1)                                         # :68
    scale = sd["logit_scale"].exp()                                 # :185
    prompts = construct_prompts(prefix, ctx, suffix)                # :187
    t = text_tower(prompts, eot, sd, text_heads, deep_t, n_ctx_text, truncate=truncate)   # :188
    f = vision_tower(image, sd, vision_heads, vpt, deep_v)          # :189
    f_hat, t_hat = normalize(f), normalize(t)                       # :190-191
    logits = scale * f_hat @ t_hat.t()                              # :193
    fixed = normalize(fixed_embeddings)                             # :196-197
    with torch.no_grad():
        f_zs = normalize(vision_tower(image, sd, vision_heads))     # :198-200 (ZS_image_encoder: no prompts)
        zs_logits = scale * f_zs @ fp16_rt(fixed).t()               # :201
    if focal_alpha is None:
        ce = F.cross_entropy(logits, label)                         # :204
    else:
        ce = focal_ce(logits, label, focal_alpha)
    l_text = F.l1_loss(t_hat, fixed, reduction="mean") * w_text      # :304-306
    l_img = F.l1_loss(f_hat, f_zs, reduction="mean") * w_image       # :309-311
    l_kl = F.kl_div(F.log_softmax(logits, 1), F.log_softmax(zs_logits, 1), reduction="sum",
                    log_target=True) / logits.numel() * w_logits     # :314-321
    loss = ce + l_kl + l_text + l_img                               # :323-324
    return {"loss": loss, "ce": ce, "kl": l_kl, "l1_text": l_text, "l1_image": l_img, "logits": logits,
            "zs_logits": zs_logits, "text_features": t_hat, "image_features": f_hat, "zs_image_features": f_zs}


def teacher_text_features(sd, teacher_tokens, text_heads, truncate=None):
    """fixed_embeddings: frozen CLIP encode_text("a photo of a {}.") (promptsrc.py:122-132; model.py:605-618)."""
    with torch.no_grad():
        emb = sd["token_embedding.weight"][teacher_tokens]
        return text_tower(emb, teacher_tokens.argmax(-1), sd, text_heads, truncate=truncate)


def construct_prompts_positioned(prefix, ctx, suffix, name_lens, position):
    """PromptLearner.forward of CoOp for class_token_position middle / front (trainers/coop.py:268-291)."""
    C = prefix.shape[0]
    if ctx.dim() == 2:
        ctx = ctx.unsqueeze(0).expand(C, -1, -1)
    half = ctx.shape[1] // 2
    rows = []
    for i in range(C):
        nl = int(name_lens[i])
        cls_i, suf_i = suffix[i:i + 1, :nl], suffix[i:i + 1, nl:]
        if position == "middle":
            rows.append(torch.cat([prefix[i:i + 1], ctx[i:i + 1, :half], cls_i, ctx[i:i + 1, half:], suf_i], dim=1))
        else:
            rows.append(torch.cat([prefix[i:i + 1], cls_i, ctx[i:i + 1], suf_i], dim=1))
    return torch.cat(rows, dim=0)


def clip_logits(sd, tokens, image, ctx, n_ctx, vision_heads, text_heads, truncate=None, ctx_per_class=False,
                position="end"):
    """CoOp CustomCLIP.forward_once, trainers/coop.py:351-365 (class_token_position 'end' :267, 'middle' :268-281,
    'front' :282-291; name_lens = tokens of the class name = EOT index - n_ctx - 2)."""
    emb = sd["token_embedding.weight"][tokens]
    prefix, suffix = emb[:, :1], emb[:, 1 + n_ctx:]                 # coop.py:247-248
    if position == "end":
        prompts = construct_prompts(prefix, ctx, suffix)
    else:
        prompts = construct_prompts_positioned(prefix, ctx, suffix, tokens.argmax(-1) - n_ctx - 2, position)
    f_hat = normalize(vision_tower(image, sd, vision_heads))
    t_hat = normalize(text_tower(prompts, tokens.argmax(-1), sd, text_heads, truncate=truncate))
    return sd["logit_scale"].exp() * f_hat @ t_hat.t()


def coop_step(sd, tokens, image, label, ctx, n_ctx, vision_heads, text_heads, focal_alpha=None, use_focal=False,
              truncate=None, position="end"):
    """coop.py:380-385: criterion_ce(logits, lbl) with CE (:323) or MultiClassFocalLoss (:346)."""
    logits = clip_logits(sd, tokens, image, ctx, n_ctx, vision_heads, text_heads, truncate, position=position)
    loss = focal_ce(logits, label, focal_alpha) if use_focal else F.cross_entropy(logits, label)
    return {"loss": loss, "logits": logits}


def maple_step(sd, tokens, image, label, ctx, compound_text, proj, compound_proj, n_ctx, vision_heads, text_heads,
               focal_alpha=None, use_focal=False, truncate=None):
    """MaPLe: MultiModalPromptLearner.forward (trainers/maple.py:190-204) + CustomCLIP.forward (:240-255).
    ctx [n_ctx, Dt]; compound_text: list of [n_ctx, Dt] (layers 1..depth-1); proj = (W [Dv, Dt], b [Dv]) couples ctx
    to the first-layer vision prompt (:145, :204); compound_proj: list of (W, b), one Linear per deeper layer
    (:155-156, :199-201).  The towers replace prompts exactly like the IVLP blocks, each with `.half()`
    (clip/model.py:287-331, :454-485; TextEncoder maple.py:62-75)."""
    emb = sd["token_embedding.weight"][tokens]
    prefix, suffix = emb[:, :1], emb[:, 1 + n_ctx:]                 # maple.py:168-169
    prompts = construct_prompts(prefix, ctx, suffix)
    shared_ctx = ctx @ proj[0].t() + proj[1]                        # :204 self.proj(self.ctx)
    deep_v = {i + 1: t @ w.t() + b for i, (t, (w, b)) in enumerate(zip(compound_text, compound_proj))}  # :199-201
    deep_t = {i + 1: t for i, t in enumerate(compound_text)}
    t = text_tower(prompts, tokens.argmax(-1), sd, text_heads, deep_t, n_ctx, truncate=truncate)
    f = vision_tower(image, sd, vision_heads, shared_ctx, deep_v)
    logits = sd["logit_scale"].exp() * normalize(f) @ normalize(t).t()
    loss = focal_ce(logits, label, focal_alpha) if use_focal else F.cross_entropy(logits, label)
    return {"loss": loss, "logits": logits}


def cocoop_step(sd, tokens, image, label, ctx, meta, n_ctx, vision_heads, text_heads, focal_alpha=None,
                use_focal=False, truncate=None):
    """CoCoOp: PromptLearner.forward (trainers/cocoop.py:173-198) + CustomCLIP.forward (:235-260).
    meta = (W1 [Hm, E], b1 [Hm], W2 [Dt, Hm], b2 [Dt]) is `meta_net` (:139-143); every image gets its own
    context `ctx + meta_net(f_hat)` and therefore its own pass of the text encoder (:249-252)."""
    w1, b1, w2, b2 = meta
    emb = sd["token_embedding.weight"][tokens]
    prefix, suffix = emb[:, :1], emb[:, 1 + n_ctx:]                 # cocoop.py:162-163
    eot = tokens.argmax(-1)
    f_hat = normalize(vision_tower(image, sd, vision_heads))        # :241-242
    bias = torch.relu(f_hat @ w1.t() + b1) @ w2.t() + b2            # :185
    scale = sd["logit_scale"].exp()
    rows = []
    for i in range(image.shape[0]):                                 # :249-252
        prompts = construct_prompts(prefix, ctx + bias[i], suffix)  # :187-193
        t_hat = normalize(text_tower(prompts, eot, sd, text_heads, truncate=truncate))
        rows.append(scale * f_hat[i] @ t_hat.t())
    logits = torch.stack(rows)                                      # :255
    loss = focal_ce(logits, label, focal_alpha) if use_focal else F.cross_entropy(logits, label)
    return {"loss": loss, "logits": logits}


def sgd_step(params, grads, bufs, lr, momentum=0.9, weight_decay=5e-4):
    """torch.optim.SGD as built by dassl/optim/optimizer.py:105-113 (dampening 0, no nesterov)."""
    out_p, out_b = [], []
    for p, g, b in zip(params, grads, bufs):
        d = g + weight_decay * p
        b = d.clone() if b is None else momentum * b + d
        out_p.append(p - lr * b)
        out_b.append(b)
    return out_p, out_b
