"""End-to-end parity of the CUDA training step against the committed reference outputs
(tests/golden/*.npz, produced by the unmodified reference -- oracle/make_golden.py).

`python tests/step_checks.py` prints every metric; `tests/test_step_gpu.py` asserts them.
Tolerances are BASELINE.json's: logits <= 2e-2 abs in bf16 / 1e-4 in fp32 mode, prompt-gradient
cosine >= 0.999, argmax bit-exact (fp32 mode; bf16 where the reference margin allows).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

from helpers import (COCOOP_REF, cocoop_setup, coop_setup, cos, load_golden, maple_ref_name, maple_setup,  # noqa: E402
                     promptsrc_setup, ref_name)

CASES = {"tiny_promptsrc": ("tiny", 3), "cfg1_promptsrc": ("ViT-B/16", 9),
         # BASELINE configs[2] itself (what bench.py times): 1000 classes, 128 images, depth 9 -- oracle/make_golden_full.py
         "cfg3_promptsrc_ce": ("ViT-B/16", 9), "cfg3_promptsrc_ragged_focal": ("ViT-B/16", 9),
         "cfg3_promptsrc_ref_b8": ("ViT-B/16", 9)}
# BASELINE.json tolerances, stated for the ViT-B/16 configs: fp32 mode 1e-4 on logits, bf16 2e-2, gradient cosine 0.999.
# The north star's gradient gate (cosine >= 0.999) is applied to EVERY trainable tensor and to the concatenated vector.
TOL = {"fp32": dict(logits=1e-4, loss=1e-5, feat=2e-5, cos=0.9999, cos_all=0.9999),
       "bf16": dict(logits=2e-2, loss=1e-2, feat=5e-3, cos=0.999, cos_all=0.999)}
# (Round 1 widened the 16-bit tolerances for the 64-dim 'tiny' architecture; with fp16 forward operands it meets the
# BASELINE numbers like every other case -- measured logits <= 4.2e-3, worst gradient cosine 0.99925 -- so it no
# longer has tolerances of its own.)


def run_promptsrc_case(name: str, prec: str, sgd: bool = True, truncate: bool = True) -> dict:
    from psrc_b200 import ops
    arch_name, depth = CASES[name]
    fx = load_golden(name)
    dev = "cuda"
    from psrc_b200 import synthetic as S
    B = fx["label"].shape[0]
    alpha = S.focal_alpha(fx["shots"].tolist(), fx["tokens"].shape[0]) if int(fx.get("focal", 0)) else None
    step, sd, fixed = promptsrc_setup(arch_name, depth, fx["tokens"], dev, prec, batch=B, focal_alpha=alpha,
                                      truncate=truncate)
    image, label = S.synthetic_batch(B, fx["tokens"].shape[0], S.ARCHS[arch_name].image_resolution, seed=1)
    assert np.array_equal(label.numpy().astype(np.int32), fx["label"]), "synthetic labels drifted from the fixture"
    step.image.copy_(image)
    step.labels.copy_(label.int())
    step.forward_backward()
    torch.cuda.synchronize()
    m = {}
    g = lambda k: torch.from_numpy(fx[k])  # noqa: E731
    if "fixed_embeddings8" in fx:          # full-size fixtures keep every 8th class row of the [C, E] tensors
        fe, te = g("fixed_embeddings8"), g("text_features8")
        m["fixed_embeddings_rel"] = float((fixed.cpu()[::8] - fe).norm() / fe.norm())
        m["text_feat_maxabs"] = float((step.t_hat.cpu()[::8] - te).abs().max())
    else:
        m["fixed_embeddings_rel"] = float((fixed.cpu() - g("fixed_embeddings")).norm() / g("fixed_embeddings").norm())
        m["text_feat_maxabs"] = float((step.t_hat.cpu() - g("text_features")).abs().max())
    m["logits_maxabs"] = float((step.logits.cpu() - g("logits")).abs().max())
    dz = (step.zs_logits.cpu() - g("zs_logits")).abs()
    m["zs_logits_maxabs"] = float(dz.max())
    m["zs_logits_frac_over_1e-4"] = float((dz > 1e-4).float().mean())
    m["image_feat_maxabs"] = float((step.f_hat.cpu() - g("image_features")).abs().max())
    m["zs_image_feat_maxabs"] = float((step.zf_hat.cpu() - g("zs_image_features")).abs().max())
    loss = float(step.losses.sum())
    m["loss"] = loss
    m["loss_rel"] = abs(loss - float(fx["loss"])) / abs(float(fx["loss"]))
    m["ce_rel"] = abs(float(step.losses[0]) - float(fx["ce"])) / abs(float(fx["ce"]))
    if "loss_terms" in fx:                 # ce, kl, 25 * L1(text), 10 * L1(image): each term on its own
        ours, ref = step.losses.cpu().numpy(), fx["loss_terms"]
        rel = np.abs(ours - ref) / np.maximum(np.abs(ref), 1e-12)
        m["loss_terms_rel_max"] = float(np.max(rel[[0, 2, 3]]))
        m["kl_term_rel"] = float(rel[1])
    m.update(_argmax_metrics(step.logits, g("logits"), fx["argmax"], m["logits_maxabs"]))
    worst, worst_key = 2.0, None
    for k in step.params:
        c = cos(step.grads[k], fx["grad/" + ref_name(k)])
        m[f"gradcos/{k}"] = c
        if c < worst:
            worst, worst_key = c, k
    m["gradcos_worst"], m["gradcos_worst_key"] = worst, worst_key
    gn = torch.cat([step.grads[k].flatten().cpu() for k in step.params])
    rn = torch.cat([torch.from_numpy(fx["grad/" + ref_name(k)]).flatten() for k in step.params])
    m["grad_rel"] = float((gn - rn).norm() / rn.norm())
    m["gradcos_all"] = cos(gn, rn)
    if sgd:  # SGD lr 0.0025 / momentum 0.9 / wd 5e-4 (make_golden: promptsrc_forward_backward)
        lr = torch.tensor([0.0025], device=dev)
        buf = torch.zeros_like(step.flat_param)
        before = step.flat_param.clone()
        ops.sgd_step(step.flat_param, step.flat_grad, buf, lr, 0.9, 0.0, 5e-4, False, True)
        torch.cuda.synchronize()
        d_ours = step.flat_param.cpu()  # same order as step.params
        upd_ref = torch.cat([(torch.from_numpy(fx["after/" + ref_name(k)])).flatten() for k in step.params])
        m["param_after_maxabs"] = float((d_ours - upd_ref).abs().max())
        m["param_update_cos"] = cos(d_ours - before.cpu(), upd_ref - before.cpu())
    return m


def _row_tag(fx, kind, n):
    for tag in ("", "4", "8", "32"):
        if f"{kind}{tag}/{n}" in fx:
            return tag
    raise KeyError(f"{kind}/{n}")


def _argmax_metrics(logits, ref_logits, ref_argmax, logits_maxabs):
    """Predicted classes: bit-exact where the reference's top-2 margin exceeds twice the measured logit error
    (SURVEY 7.2); the excluded fraction is reported."""
    am = logits.argmax(1).cpu().numpy().astype(np.int32)
    top2 = ref_logits.topk(2, dim=1).values
    decided = ((top2[:, 0] - top2[:, 1]) > 2 * logits_maxabs).numpy()
    return {"argmax_equal_all": bool(np.array_equal(am, ref_argmax)),
            "argmax_equal_decided": bool(np.array_equal(am[decided], ref_argmax[decided])),
            "argmax_excluded_frac": float(1 - decided.mean()),
            "ref_top2_margin_median": float((top2[:, 0] - top2[:, 1]).median())}


def _generic_metrics(step, fx, name_of, lr, dev, sampled=False):
    """logits / loss / per-tensor gradient cosines / parameters after the reference's SGD step."""
    from psrc_b200 import ops
    m = {}
    m["logits_maxabs"] = float((step.logits.cpu() - torch.from_numpy(fx["logits"])).abs().max())
    loss = float(step.losses.sum())
    m["loss"], m["loss_rel"] = loss, abs(loss - float(fx["loss"])) / abs(float(fx["loss"]))
    worst, worst_key, ours, refs = 2.0, None, [], []
    if "argmax" in fx:
        m.update(_argmax_metrics(step.logits, torch.from_numpy(fx["logits"]), fx["argmax"], m["logits_maxabs"]))
    for k in step.params:
        n = name_of(k)
        g = step.grads[k].cpu()
        tag = _row_tag(fx, "grad", n)      # large tensors: every 4th / 8th / 32nd row is in the fixture
        if tag:
            g = g[::int(tag)]
        r = torch.from_numpy(fx[f"grad{tag}/" + n])
        c = cos(g, r)
        m[f"gradcos/{k}"] = c
        ours.append(g.flatten()), refs.append(r.flatten())
        if c < worst:
            worst, worst_key = c, k
    m["gradcos_worst"], m["gradcos_worst_key"] = worst, worst_key
    m["gradcos_all"] = cos(torch.cat(ours), torch.cat(refs))
    buf = torch.zeros_like(step.flat_param)
    before = {k: step.params[k].detach().cpu().clone() for k in step.params}
    ops.sgd_step(step.flat_param, step.flat_grad, buf, torch.tensor([lr], device=dev), 0.9, 0.0, 5e-4, False, True)
    torch.cuda.synchronize()
    du, dr = [], []
    for k in step.params:
        n = name_of(k)
        a = step.params[k].cpu()
        tag = _row_tag(fx, "after", n)
        r, b0 = torch.from_numpy(fx[f"after{tag}/" + n]), before[k]
        if tag:
            a, b0 = a[::int(tag)], b0[::int(tag)]
        du.append((a - b0).flatten()), dr.append((r - b0).flatten())
    m["param_update_cos"] = cos(torch.cat(du), torch.cat(dr))
    return m


def run_maple_case(name: str, prec: str) -> dict:
    """MaPLe ViT-B/16 (the reference hard-codes these widths) against tests/golden/vitb_maple_*.npz."""
    from psrc_b200 import synthetic as S
    fx = load_golden(name)
    dev, arch = "cuda", S.ARCHS["ViT-B/16"]
    B, C = fx["label"].shape[0], fx["tokens"].shape[0]
    alpha = S.focal_alpha(fx["shots"].tolist(), C) if name.endswith("focal") else None
    step, sd, _ = maple_setup("ViT-B/16", int(fx["depth"]), fx["tokens"], dev, prec, batch=B, n_ctx=int(fx["n_ctx"]),
                              focal_alpha=alpha, ctx_init=torch.from_numpy(fx["init/prompt_learner.ctx"]))
    image, label = S.synthetic_batch(B, C, arch.image_resolution, seed=3)
    assert np.array_equal(label.numpy().astype(np.int32), fx["label"])
    step.image.copy_(image)
    step.labels.copy_(label.int())
    step.forward_backward()
    torch.cuda.synchronize()
    return _generic_metrics(step, fx, maple_ref_name, 0.0035, dev)


def run_cocoop_case(name: str, prec: str) -> dict:
    """CoCoOp (tiny architecture) against tests/golden/tiny_cocoop_*.npz."""
    from psrc_b200 import synthetic as S
    fx = load_golden(name)
    dev, arch = "cuda", S.ARCHS["tiny"]
    B, C = fx["label"].shape[0], fx["tokens"].shape[0]
    alpha = S.focal_alpha(fx["shots"].tolist(), C) if name.endswith("focal") else None
    step, sd, _ = cocoop_setup("tiny", fx["tokens"], dev, prec, batch=B, n_ctx=int(fx["n_ctx"]), focal_alpha=alpha)
    image, label = S.synthetic_batch(B, C, arch.image_resolution, seed=4)
    assert np.array_equal(label.numpy().astype(np.int32), fx["label"])
    step.image.copy_(image)
    step.labels.copy_(label.int())
    step.forward_backward()
    torch.cuda.synchronize()
    return _generic_metrics(step, fx, lambda k: COCOOP_REF[k], 0.002, dev)


def run_coop_case(name: str, prec: str) -> dict:
    """CoOp (tiny architecture; class token position end / middle / front) against tests/golden/tiny_coop_*.npz."""
    from psrc_b200 import synthetic as S
    fx = load_golden(name)
    dev = "cuda"
    B, C = fx["label"].shape[0], fx["tokens"].shape[0]
    focal = name.endswith("focal") or name.endswith("front")
    position = {"tiny_coop_middle": "middle", "tiny_coop_front": "front"}.get(name, "end")
    alpha = S.focal_alpha(fx["shots"].tolist(), C) if focal else None
    step, sd = coop_setup("tiny", fx["tokens"], dev, prec, B, fx["ctx0"], focal_alpha=alpha, position=position)
    image, label = S.synthetic_batch(B, C, 64, seed=2)
    assert np.array_equal(label.numpy().astype(np.int32), fx["label"])
    step.image.copy_(image)
    step.labels.copy_(label.int())
    step.forward_backward()
    torch.cuda.synchronize()
    return _generic_metrics(step, fx, lambda k: "prompt_learner.ctx", 0.002, dev)


def run_coop_full_case(name: str, prec: str) -> dict:
    """BASELINE configs[1]: CoOp ViT-B/16, 16 context tokens, 100 classes with CIFAR-100-LT-shaped counts, batch 32,
    CE / focal / class-specific contexts -- reference outputs (tests/golden/cfg2_coop_*.npz)."""
    from psrc_b200 import synthetic as S
    fx = load_golden(name)
    dev, arch = "cuda", S.ARCHS["ViT-B/16"]
    B, C, n_ctx = fx["label"].shape[0], fx["tokens"].shape[0], int(fx["n_ctx"])
    alpha = S.focal_alpha(fx["shots"].tolist(), C) if int(fx["focal"]) else None
    csc = bool(int(fx["csc"]))
    if csc:   # oracle/make_golden_full.py:csc_ctx_init
        ctx0 = torch.randn(C, n_ctx, arch.transformer_width, generator=torch.Generator().manual_seed(12)) * 0.02
    else:
        ctx0 = fx["init/prompt_learner.ctx"]
    step, sd = coop_setup("ViT-B/16", fx["tokens"], dev, prec, B, ctx0, n_ctx=n_ctx, focal_alpha=alpha, csc=csc)
    image, label = S.synthetic_batch(B, C, arch.image_resolution, seed=2)
    assert np.array_equal(label.numpy().astype(np.int32), fx["label"])
    step.image.copy_(image)
    step.labels.copy_(label.int())
    step.forward_backward()
    torch.cuda.synchronize()
    return _generic_metrics(step, fx, lambda k: "prompt_learner.ctx", 0.002, dev)


def run_cocoop_full_case(name: str, prec: str) -> dict:
    """BASELINE configs[3]: CoCoOp ViT-B/16, 365 classes, focal; 8 images (pinned oracle) / 2 images (reference)."""
    from psrc_b200 import synthetic as S
    fx = load_golden(name)
    dev, arch = "cuda", S.ARCHS["ViT-B/16"]
    B, C = fx["label"].shape[0], fx["tokens"].shape[0]
    alpha = S.focal_alpha(fx["shots"].tolist(), C) if int(fx["focal"]) else None
    step, sd, _ = cocoop_setup("ViT-B/16", fx["tokens"], dev, prec, batch=B, n_ctx=int(fx["n_ctx"]), focal_alpha=alpha)
    image, label = S.synthetic_batch(B, C, arch.image_resolution, seed=4)
    assert np.array_equal(label.numpy().astype(np.int32), fx["label"])
    step.image.copy_(image)
    step.labels.copy_(label.int())
    step.forward_backward()
    torch.cuda.synchronize()
    return _generic_metrics(step, fx, lambda k: COCOOP_REF[k], 0.002, dev)


def run_maple_full_case(name: str, prec: str) -> dict:
    """BASELINE configs[4]: MaPLe ViT-L/14, 1000 classes, depth 9, focal, 16 images (one rank's share of 128 at 8 GPUs)
    against the pinned oracle (the reference hard-codes ViT-B widths)."""
    from psrc_b200 import synthetic as S
    fx = load_golden(name)
    dev, arch = "cuda", S.ARCHS["ViT-L/14"]
    B, C = fx["label"].shape[0], fx["tokens"].shape[0]
    alpha = S.focal_alpha(fx["shots"].tolist(), C) if int(fx["focal"]) else None
    step, sd, _ = maple_setup("ViT-L/14", int(fx["depth"]), fx["tokens"], dev, prec, batch=B, n_ctx=int(fx["n_ctx"]),
                              focal_alpha=alpha)
    image, label = S.synthetic_batch(B, C, arch.image_resolution, seed=3)
    assert np.array_equal(label.numpy().astype(np.int32), fx["label"])
    step.image.copy_(image)
    step.labels.copy_(label.int())
    step.forward_backward()
    torch.cuda.synchronize()
    return _generic_metrics(step, fx, maple_ref_name, 0.0035, dev)


def run_maple_vitl_case(name: str, prec: str) -> dict:
    """BASELINE config #5's architecture (MaPLe on ViT-L/14: vision 1024 x 24 layers x 16 heads, 257 + 2 tokens; text
    768 x 12 heads) against the oracle restatement run on the CPU in this process -- the reference itself cannot
    express it (768 / 512 hard-coded, no ViT-L/14 in clip._MODELS), so there is no golden file."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import clip_oracle as O
    from psrc_b200 import ops, synthetic as S
    dev, arch = "cuda", S.ARCHS["ViT-L/14"]
    depth, n_ctx, B = 3, 2, 2
    tokens = S.synthetic_tokens(10, name_lens=[1, 2, 3, 8, 1, 1, 2, 2, 1, 2])
    C = tokens.shape[0]
    step, sd, _ = maple_setup("ViT-L/14", depth, tokens, dev, prec, batch=B, n_ctx=n_ctx)
    image, label = S.synthetic_batch(B, C, arch.image_resolution, seed=3)
    step.image.copy_(image)
    step.labels.copy_(label.int())
    step.forward_backward()
    torch.cuda.synchronize()
    P = {k: step.params[k].detach().cpu().clone().requires_grad_(True) for k in step.params}
    torch.set_num_threads(os.cpu_count() or 1)
    o = O.maple_step(sd, torch.as_tensor(tokens), image, label, P["ctx"], [P[f"text.{i}"] for i in range(1, depth)],
                     (P["couple.w.0"], P["couple.b.0"]),
                     [(P[f"couple.w.{i}"], P[f"couple.b.{i}"]) for i in range(1, depth)], n_ctx, arch.vision_heads,
                     arch.transformer_heads, truncate=int(torch.as_tensor(tokens).argmax(-1).max()) + 1)
    o["loss"].backward()
    m = {"logits_maxabs": float((step.logits.cpu() - o["logits"].detach()).abs().max())}
    loss = float(step.losses.sum())
    m["loss"], m["loss_rel"] = loss, abs(loss - float(o["loss"].detach())) / abs(float(o["loss"].detach()))
    cs = {k: cos(step.grads[k], P[k].grad) for k in step.params}
    m["gradcos_worst_key"] = min(cs, key=cs.get)
    m["gradcos_worst"] = cs[m["gradcos_worst_key"]]
    m["gradcos_all"] = cos(torch.cat([step.grads[k].flatten().cpu() for k in step.params]),
                           torch.cat([P[k].grad.flatten() for k in step.params]))
    m["param_update_cos"] = m["gradcos_all"]
    m["vision_seq_len"], m["text_width"] = step.Lv, step.Dt
    return m


def check_generic(m: dict, prec: str, name: str = "") -> list[str]:
    t = TOL[prec]
    bad = []
    if m["logits_maxabs"] > t["logits"]:
        bad.append(f"logits max|d| {m['logits_maxabs']:.3e} > {t['logits']}")
    if m["loss_rel"] > t["loss"]:
        bad.append(f"loss rel {m['loss_rel']:.3e} > {t['loss']}")
    if m["gradcos_worst"] < t["cos"]:
        bad.append(f"grad cosine {m['gradcos_worst']:.6f} ({m['gradcos_worst_key']}) < {t['cos']}")
    if m["gradcos_all"] < t["cos_all"]:
        bad.append(f"concatenated gradient cosine {m['gradcos_all']:.6f} < {t['cos_all']}")
    if m["param_update_cos"] < t["cos"]:
        bad.append(f"SGD update cosine {m['param_update_cos']:.6f} < {t['cos']}")
    if "argmax_equal_all" in m:
        if prec == "fp32" and not m["argmax_equal_all"]:
            bad.append("argmax differs in fp32 mode")
        if not m["argmax_equal_decided"]:
            bad.append("argmax differs where the reference margin exceeds twice the logit error")
    return bad


OTHER_CASES = {"vitb_maple_ce": run_maple_case, "vitb_maple_focal": run_maple_case,
               "tiny_cocoop_ce": run_cocoop_case, "tiny_cocoop_focal": run_cocoop_case,
               "vitl_maple_oracle": run_maple_vitl_case,
               "tiny_coop_ce": run_coop_case, "tiny_coop_focal": run_coop_case, "tiny_coop_middle": run_coop_case,
               "tiny_coop_front": run_coop_case,
               # BASELINE configs[1], [3], [4] at full size (oracle/make_golden_full.py)
               "cfg2_coop_ce": run_coop_full_case, "cfg2_coop_focal": run_coop_full_case,
               "cfg2_coop_csc": run_coop_full_case, "cfg4_cocoop_focal": run_cocoop_full_case,
               "cfg4_cocoop_ref_b2": run_cocoop_full_case, "cfg5_maple_vitl": run_maple_full_case}


def check(m: dict, prec: str, name: str = "") -> list[str]:
    t = TOL[prec]
    bad = []
    if m["logits_maxabs"] > t["logits"]:
        bad.append(f"logits max|d| {m['logits_maxabs']:.3e} > {t['logits']}")
    if prec == "fp32":
        # The teacher logits are s * f_zs . half(t_fixed) (promptsrc.py:201): the reference's own fp16 round trip of the
        # init-time teacher text features is a discontinuity -- where our fp32 value and the reference's (equal to
        # ~1e-7 relative) straddle an fp16 rounding boundary (probability ~2^-13 per element, i.e. a few dozen of the
        # 512 000 elements at C = 1000) that element moves by one fp16 ulp (2^-15 .. 2^-14 at |t| ~ 0.03 .. 0.1) and the
        # logits of that class by up to s * |f| * ulp = 100 * 0.15 * 6.1e-5 = 9e-4.  So: 1e-4 everywhere except an
        # isolated <= 1 % of the entries (measured 0.22 % at C = 1000: a flipped class shows in the rows of the few
        # images whose feature is large in that dimension), which stay below 1e-3.
        if m["zs_logits_maxabs"] > 1e-3 or m["zs_logits_frac_over_1e-4"] > 1e-2:
            bad.append(f"zs_logits max|d| {m['zs_logits_maxabs']:.3e} (limit 1e-3), fraction over 1e-4 "
                       f"{m['zs_logits_frac_over_1e-4']:.2e} (limit 1e-2)")
    elif m["zs_logits_maxabs"] > t["logits"]:
        bad.append(f"zs_logits max|d| {m['zs_logits_maxabs']:.3e} > {t['logits']}")
    for k in ("text_feat_maxabs", "image_feat_maxabs", "zs_image_feat_maxabs"):
        if m[k] > t["feat"]:
            bad.append(f"{k} {m[k]:.3e} > {t['feat']}")
    if m["loss_rel"] > t["loss"]:
        bad.append(f"loss rel {m['loss_rel']:.3e} > {t['loss']}")
    if m.get("loss_terms_rel_max", 0.0) > t["loss"]:
        bad.append(f"loss term rel {m['loss_terms_rel_max']:.3e} > {t['loss']}")
    # the KL term compares the logits with the teacher logits and inherits their fp16 rounding flips (see above): a
    # flipped class moves log-softmax of its column by ~1e-4 -- 5x the fp32 loss tolerance for this term only
    if m.get("kl_term_rel", 0.0) > 5 * t["loss"]:
        bad.append(f"KL term rel {m['kl_term_rel']:.3e} > {5 * t['loss']}")
    if m["gradcos_worst"] < t["cos"]:
        bad.append(f"grad cosine {m['gradcos_worst']:.6f} ({m['gradcos_worst_key']}) < {t['cos']}")
    if m["gradcos_all"] < t["cos_all"]:
        bad.append(f"concatenated prompt-gradient cosine {m['gradcos_all']:.6f} < {t['cos_all']}")
    if prec == "fp32" and not m["argmax_equal_all"]:
        bad.append("argmax differs in fp32 mode")
    if not m["argmax_equal_decided"]:
        bad.append("argmax differs where the reference margin exceeds twice the logit error")
    if "param_update_cos" in m and m["param_update_cos"] < t["cos"]:
        bad.append(f"SGD update cosine {m['param_update_cos']:.6f} < {t['cos']}")
    if m["fixed_embeddings_rel"] > 2e-5:
        bad.append(f"teacher text features rel {m['fixed_embeddings_rel']:.3e}")
    return bad


if __name__ == "__main__":
    names = [a for a in sys.argv[1:] if a in CASES] or list(CASES)
    others = [a for a in sys.argv[1:] if a in OTHER_CASES] or ([] if names != list(CASES) else list(OTHER_CASES))
    precs = [a for a in sys.argv[1:] if a in TOL] or ["fp32", "bf16"]
    rc = 0
    for n in others:
        for p in precs:
            m = OTHER_CASES[n](n, p)
            bad = check_generic(m, p, n)
            print(f"===== {n} [{p}] {'PASS' if not bad else 'FAIL: ' + '; '.join(bad)}")
            for k, v in m.items():
                print(f"   {k:32s} {v}")
            rc |= bool(bad)
    if others and not [a for a in sys.argv[1:] if a in CASES] and sys.argv[1:]:
        sys.exit(rc)
    for n in names:
        for p in precs:
            m = run_promptsrc_case(n, p)
            bad = check(m, p, n)
            print(f"===== {n} [{p}] {'PASS' if not bad else 'FAIL: ' + '; '.join(bad)}")
            for k, v in m.items():
                print(f"   {k:32s} {v}")
            rc |= bool(bad)
    sys.exit(rc)


def run_tiny_oracle_case(B: int, name_lens: list, prec: str) -> dict:
    """PromptSRC on the tiny architecture at arbitrary (small) batch / class counts against the oracle restatement run
    on the CPU in this process (edge cases: one image, two classes, one long class name)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import clip_oracle as O
    from psrc_b200 import synthetic as S
    arch = S.ARCHS["tiny"]
    depth, n_ctx = 3, 4
    tokens = S.synthetic_tokens(len(name_lens), name_lens=name_lens)
    sd = S.make_clip_state_dict(arch, seed=0)
    image, label = S.synthetic_batch(B, tokens.shape[0], arch.image_resolution, seed=1)
    tk = torch.as_tensor(tokens)
    fixed = O.teacher_text_features(sd, tk, arch.transformer_heads)
    deep = S.make_deep_prompts(arch, n_ctx, n_ctx, depth, depth, seed=100)
    ctx = sd["token_embedding.weight"][tk[0, 1:1 + n_ctx]].clone().requires_grad_(True)
    vpt = deep["visual.VPT"].clone().requires_grad_(True)
    dv = {i: deep[f"visual.transformer.resblocks.{i}.VPT_shallow"].clone().requires_grad_(True) for i in range(1, depth)}
    dtx = {i: deep[f"transformer.resblocks.{i}.VPT_shallow"].clone().requires_grad_(True) for i in range(1, depth)}
    o = O.promptsrc_step(sd, tk, image, label, ctx, vpt, dv, dtx, fixed, n_ctx, arch.vision_heads, arch.transformer_heads)
    o["loss"].backward()
    og = {"ctx": ctx.grad, "vpt": vpt.grad}
    og.update({f"vision.{i}": t.grad for i, t in dv.items()})
    og.update({f"text.{i}": t.grad for i, t in dtx.items()})
    step, _, _ = promptsrc_setup("tiny", depth, tokens, "cuda", prec, batch=B, sd=sd)
    step.image.copy_(image)
    step.labels.copy_(label.int())
    step.forward_backward()
    torch.cuda.synchronize()
    cs = {k: cos(step.grads[k], og[k]) for k in step.params}
    return {"logits_maxabs": float((step.logits.cpu() - o["logits"].detach()).abs().max()),
            "loss_rel": abs(float(step.losses.sum()) - float(o["loss"].detach())) / abs(float(o["loss"].detach())),
            "gradcos_worst": min(cs.values()), "gradcos_worst_key": min(cs, key=cs.get),
            "argmax_equal": bool((step.logits.argmax(1).cpu() == o["logits"].argmax(1)).all()), "Lt": step.Lt}
