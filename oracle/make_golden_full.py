"""TEST INFRASTRUCTURE -- golden fixtures at BASELINE.json's own configurations (full-size towers, full class counts).

Run in the build container only:  python oracle/make_golden_full.py [case ...]      (see CASES; ~1-4 min of CPU each)

`oracle/make_golden.py` pins the restatement `clip_oracle.py` against the UNMODIFIED reference on small cases and on
ViT-B/16 with <= 100 classes.  This script produces the fixtures for the configurations `bench.py` times and BASELINE.json
names (SURVEY 8d "Synthetic inputs"):

  source = "reference": the unmodified reference modules run the step (PromptSRC.forward_backward / CoOp / CoCoOp
      CustomCLIP.forward + loss.backward + SGD) and the restatement is asserted against them again at this size.
  source = "oracle": the pinned restatement runs it, with the text tower truncated to max(EOT)+1 positions (exact, SURVEY
      F12, asserted against the dense reference in every "reference" case).  Used where the reference cannot express the
      case (PromptSRC + focal CE lives only in non-imported files with hard-coded class counts, F14; MaPLe hard-codes the
      ViT-B widths, F7) or cannot run it in this container's 62 GB (CoCoOp at 8 x 365 sequences x 77 positions keeps
      > 100 GB of fp32 activations for autograd; the reference CoCoOp case therefore runs 2 images x 365 classes).
      PromptSRC at the full cfg #3 (1000 classes x 77 positions, 128 images) DOES fit: 47 GB, and is a "reference" case.

Inputs are regenerated from seeds on the GPU box (psrc_b200.synthetic); the fixture stores tokens, labels and outputs only:
logits, loss terms, every trainable gradient, parameters after one SGD step (large coupling matrices: every 32nd row).
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import clip_oracle as O  # noqa: E402
import ref_harness as H  # noqa: E402
import psrc_b200  # noqa: E402,F401
from psrc_b200 import synthetic as S  # noqa: E402
from make_golden import cos, rel  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

ADJ = ["", "big ", "very old red ", "small green "]


def names_uniform(n):
    return [f"object {i}" for i in range(n)]


def names_ragged(n):
    """Class names of 2-5 tokens ('add a variant with name lengths 1-4 tokens to exercise ragged EOT')."""
    return [f"{ADJ[i % 4]}object {i}" for i in range(n)]


# name: dict(trainer, source, arch, names, C, B, ...)
CASES = {
    # BASELINE configs[2]: PromptSRC ViT-B/16, 1000 classes, 4+4 deep prompts at depth 9, batch 128 -- what bench.py times
    # (the unmodified reference runs this one whole: 143 s, 47 GB peak RSS on the 8-core / 62 GB build container)
    "cfg3_promptsrc_ce": dict(trainer="PromptSRC", source="reference", arch="ViT-B/16", names=names_uniform, C=1000,
                              B=128, depth=9, focal=False),
    "cfg3_promptsrc_ragged_focal": dict(trainer="PromptSRC", source="oracle", arch="ViT-B/16", names=names_ragged, C=1000,
                                        B=128, depth=9, focal=True),
    # the same step on the reference itself at the largest size that fits here (pins the oracle at ViT-B/16, C = 1000)
    "cfg3_promptsrc_ref_b8": dict(trainer="PromptSRC", source="reference", arch="ViT-B/16", names=names_ragged, C=1000, B=8,
                                  depth=9, focal=False),
    # BASELINE configs[1]: CoOp ViT-B/16, 16 context tokens, CIFAR-100-LT-shaped counts (imbalance 100), batch 32
    "cfg2_coop_ce": dict(trainer="CoOp", source="reference", arch="ViT-B/16", names=names_ragged, C=100, B=32,
                         extra={"TRAINER.COOP.LOSS_TYPE": "ce"}),
    "cfg2_coop_focal": dict(trainer="CoOp", source="reference", arch="ViT-B/16", names=names_ragged, C=100, B=32,
                            extra={"TRAINER.COOP.LOSS_TYPE": "focal"}),
    "cfg2_coop_csc": dict(trainer="CoOp", source="reference", arch="ViT-B/16", names=names_ragged, C=100, B=32,
                          extra={"TRAINER.COOP.LOSS_TYPE": "focal", "TRAINER.COOP.CSC": True}),
    # BASELINE configs[3]: CoCoOp ViT-B/16, 365 classes, 8 images per GPU, focal on
    "cfg4_cocoop_focal": dict(trainer="CoCoOp", source="oracle", arch="ViT-B/16", names=names_ragged, C=365, B=8,
                              focal=True),
    "cfg4_cocoop_ref_b2": dict(trainer="CoCoOp", source="reference", arch="ViT-B/16", names=names_ragged, C=365, B=2,
                               focal=True),
    # BASELINE configs[4]: MaPLe ViT-L/14, 1000 classes, depth 9; 16 images = one rank's share of the global 128 at 8 GPUs
    "cfg5_maple_vitl": dict(trainer="MaPLe", source="oracle", arch="ViT-L/14", names=names_uniform, C=1000, B=16, depth=9,
                            focal=True),
}


def _tokens(prefix_words, names):
    return H.tokenize([f"{prefix_words} {n}." for n in names])


def _save(name, fx):
    path = os.path.join(GOLD, name + ".npz")
    np.savez_compressed(path, **fx)
    print(f"[{name}] wrote {path} ({os.path.getsize(path) / 1e6:.2f} MB)")


def _grad_entries(fx, grads, after, init=None, sample=32):
    for n in grads:
        big = grads[n].numel() > 200_000
        tag = (f"{sample}" if big else "")
        sl = slice(None, None, sample) if big else slice(None)
        fx[f"grad{tag}/" + n] = grads[n][sl].numpy()
        fx[f"after{tag}/" + n] = after[n][sl].numpy()
        if init is not None:
            fx[f"init{tag}/" + n] = init[n][sl].numpy()


def promptsrc_case(name, c):
    arch = S.ARCHS[c["arch"]]
    sd = S.make_clip_state_dict(arch, seed=0)
    C, B, depth, n_ctx = c["C"], c["B"], c["depth"], 4
    classnames = c["names"](C)
    tokens = _tokens("a photo of a", classnames)
    image, label = S.synthetic_batch(B, C, arch.image_resolution, seed=1)
    shots = S.lt_counts_pareto(C)
    alpha = torch.tensor(S.focal_alpha(shots, C)) if c["focal"] else None
    deep = S.make_deep_prompts(arch, n_ctx, n_ctx, depth, depth, seed=100)
    trunc = int(tokens.argmax(-1).max()) + 1
    P = {"prompt_learner.ctx": sd["token_embedding.weight"][tokens[0, 1:1 + n_ctx]].clone(),
         "image_encoder.VPT": deep["visual.VPT"].clone()}
    for i in range(1, depth):
        P[f"image_encoder.transformer.resblocks.{i}.VPT_shallow"] = deep[f"visual.transformer.resblocks.{i}.VPT_shallow"].clone()
        P[f"text_encoder.transformer.resblocks.{i}.VPT_shallow"] = deep[f"transformer.resblocks.{i}.VPT_shallow"].clone()

    def oracle_run():
        p = {k: v.clone().requires_grad_(True) for k, v in P.items()}
        fixed = O.teacher_text_features(sd, tokens, arch.transformer_heads, truncate=trunc)
        dv = {i: p[f"image_encoder.transformer.resblocks.{i}.VPT_shallow"] for i in range(1, depth)}
        dt = {i: p[f"text_encoder.transformer.resblocks.{i}.VPT_shallow"] for i in range(1, depth)}
        o = O.promptsrc_step(sd, tokens, image, label, p["prompt_learner.ctx"], p["image_encoder.VPT"], dv, dt, fixed,
                             n_ctx, arch.vision_heads, arch.transformer_heads, focal_alpha=alpha, truncate=trunc)
        o["loss"].backward()
        g = {k: v.grad.detach().clone() for k, v in p.items()}
        o = {k: v.detach() for k, v in o.items()}
        o["fixed_embeddings"] = fixed
        return o, g

    t0 = time.time()
    if c["source"] == "reference":
        assert not c["focal"], "PromptSRC + focal exists only in the non-imported variants (SURVEY F14)"
        cfg = H.make_cfg("PromptSRC", **{"INPUT.SIZE": (arch.image_resolution,) * 2,
                                         "TRAINER.PROMPTSRC.PROMPT_DEPTH_VISION": depth,
                                         "TRAINER.PROMPTSRC.PROMPT_DEPTH_TEXT": depth, "OPTIM.MAX_EPOCH": 20})
        tm, model = H.build_reference_model("PromptSRC", cfg, classnames, sd)
        assert torch.equal(model.tokenized_prompts, tokens)
        ref_sd = model.state_dict()
        with torch.no_grad():
            for k, v in P.items():
                ref_sd[k].copy_(v)
        H.freeze_like_trainer("PromptSRC", model)
        seen = {}

        def grab(mod, inp, out):     # (a forward hook that returns a value would replace the module's output)
            seen.setdefault("out", [t.detach().clone() for t in out])
        hook = model.register_forward_hook(grab)
        loss = H.promptsrc_forward_backward(tm, model, cfg, image, label)[0]["loss"]
        hook.remove()
        ce, t_hat, fixed_n, f_zs, f_hat, zs_logits, logits = seen["out"]
        named = {n: p for n, p in model.named_parameters() if p.requires_grad}
        assert sorted(named) == sorted(P)
        grads = {n: p.grad.detach().clone() for n, p in named.items()}
        after = {n: p.detach().clone() for n, p in named.items()}
        fixed_raw = model.prompt_learner.fixed_embeddings.detach().clone()
        print(f"[{name}] reference forward_backward: loss={loss:.6f} ({time.time() - t0:.1f}s)")
        o, og = oracle_run()
        worst = min(cos(og[n], grads[n]) for n in grads)
        print(f"  oracle(truncate={trunc}) vs reference: loss |d|={abs(float(o['loss']) - loss):.2e} "
              f"logits max|d|={float((o['logits'] - logits).abs().max()):.2e} "
              f"teacher rel={rel(o['fixed_embeddings'], fixed_raw):.2e} grad worst cos={worst:.7f}")
        assert abs(float(o["loss"]) - loss) < 2e-5 * max(1.0, abs(loss))
        assert float((o["logits"] - logits).abs().max()) < 1e-4 and worst > 0.99999
        terms = np.array([float(o[k]) for k in ("ce", "kl", "l1_text", "l1_image")], dtype=np.float32)  # reference sums them
    else:
        o, grads = oracle_run()
        loss, logits, zs_logits = float(o["loss"]), o["logits"], o["zs_logits"]
        t_hat, f_hat, f_zs, fixed_raw, ce = o["text_features"], o["image_features"], o["zs_image_features"], \
            o["fixed_embeddings"], o["ce"]
        keys = list(P)
        newp, _ = O.sgd_step([P[k] for k in keys], [grads[k] for k in keys], [None] * len(keys), 0.0025)
        after = dict(zip(keys, newp))
        terms = np.array([float(o[k]) for k in ("ce", "kl", "l1_text", "l1_image")], dtype=np.float32)
        print(f"[{name}] oracle step: loss={loss:.6f} terms={terms} ({time.time() - t0:.1f}s)")
    fx = {"source": np.array(c["source"]), "tokens": tokens.numpy().astype(np.int32),
          "label": label.numpy().astype(np.int32), "loss": np.float32(loss), "ce": np.float32(float(ce)),
          "loss_terms": terms, "logits": logits.numpy(), "zs_logits": zs_logits.numpy(),
          "text_features8": t_hat[::8].numpy(), "image_features": f_hat.numpy(), "zs_image_features": f_zs.numpy(),
          "fixed_embeddings8": fixed_raw[::8].numpy(), "argmax": logits.argmax(1).numpy().astype(np.int32),
          "shots": np.array(shots, dtype=np.int32), "depth": np.int32(depth), "focal": np.int32(bool(c["focal"]))}
    _grad_entries(fx, grads, after)
    _save(name, fx)


def csc_ctx_init(C, n_ctx, D, seed=12):
    """nn.init.normal_(ctx_vectors, std=0.02) of coop.py:229-233 with an explicit seed."""
    return torch.randn(C, n_ctx, D, generator=torch.Generator().manual_seed(seed)) * 0.02


def coop_case(name, c):
    arch = S.ARCHS[c["arch"]]
    sd = S.make_clip_state_dict(arch, seed=0)
    C, B, n_ctx = c["C"], c["B"], 16
    classnames = c["names"](C)
    shots = S.lt_counts_cifar100(C, 500, 0.01)
    extra = c["extra"]
    over = {"INPUT.SIZE": (arch.image_resolution,) * 2, "TRAINER.COOP.N_CTX": n_ctx, "DATASET.PER_CLASS_SHOTS": shots}
    over.update(extra)
    cfg = H.make_cfg("CoOp", **over)
    torch.manual_seed(11)
    tm, model = H.build_reference_model("CoOp", cfg, classnames, sd)
    H.freeze_like_trainer("CoOp", model)
    image, label = S.synthetic_batch(B, C, arch.image_resolution, seed=2)
    tokens = model.tokenized_prompts.clone()
    csc = bool(extra.get("TRAINER.COOP.CSC", False))
    if csc:   # class-specific contexts (coop.py:222-230): seeded here so the test can regenerate them (csc_ctx_init)
        with torch.no_grad():
            model.prompt_learner.ctx.copy_(csc_ctx_init(C, n_ctx, arch.transformer_width))
    ctx0 = model.prompt_learner.ctx.detach().clone()
    assert ctx0.shape == ((C, n_ctx, arch.transformer_width) if csc else (n_ctx, arch.transformer_width))
    t0 = time.time()
    model.eval()
    with torch.no_grad():
        logits = model(image).clone()                                   # reference eval-mode logits (coop.py:386-390)
    model.train()
    opt = torch.optim.SGD(model.prompt_learner.parameters(), lr=0.002, momentum=0.9, weight_decay=5e-4)
    loss = model(image, label, None)                                    # coop.py:456
    opt.zero_grad(); loss.backward(); opt.step()                        # noqa: E702  coop.py:457-459
    grad = model.prompt_learner.ctx.grad.detach().clone()
    after = model.prompt_learner.ctx.detach().clone()
    use_focal = extra.get("TRAINER.COOP.LOSS_TYPE") == "focal"
    alpha = torch.tensor(S.focal_alpha(shots, C)) if use_focal else None
    ctx = ctx0.clone().requires_grad_(True)
    trunc = int(tokens.argmax(-1).max()) + 1
    o = O.coop_step(sd, tokens, image, label, ctx, n_ctx, arch.vision_heads, arch.transformer_heads, focal_alpha=alpha,
                    use_focal=use_focal, truncate=trunc)
    o["loss"].backward()
    print(f"[{name}] reference loss={float(loss):.6f} ({time.time() - t0:.1f}s); oracle(truncate={trunc}) |d|="
          f"{abs(float(o['loss']) - float(loss)):.2e} logits max|d|={float((o['logits'] - logits).abs().max()):.2e} "
          f"grad cos={cos(ctx.grad, grad):.7f} rel={rel(ctx.grad, grad):.2e}")
    assert abs(float(o["loss"]) - float(loss)) < 2e-5 and cos(ctx.grad, grad) > 0.99999
    assert float((o["logits"] - logits).abs().max()) < 1e-4
    fx = {"source": np.array("reference"), "tokens": tokens.numpy().astype(np.int32),
          "label": label.numpy().astype(np.int32), "loss": np.float32(float(loss)), "logits": logits.numpy(),
          "argmax": logits.argmax(1).numpy().astype(np.int32), "shots": np.array(shots, dtype=np.int32),
          "csc": np.int32(csc), "focal": np.int32(use_focal), "n_ctx": np.int32(n_ctx)}
    if csc:   # [100, 16, 512] = 3.3 MB per tensor: the seeded initial value is regenerated by the test, every 4th class stored
        fx["grad4/prompt_learner.ctx"], fx["after4/prompt_learner.ctx"] = grad[::4].numpy(), after[::4].numpy()
    else:
        _grad_entries(fx, {"prompt_learner.ctx": grad}, {"prompt_learner.ctx": after}, {"prompt_learner.ctx": ctx0})
    _save(name, fx)


def cocoop_case(name, c):
    arch = S.ARCHS[c["arch"]]
    sd = S.make_clip_state_dict(arch, seed=0)
    C, B, n_ctx = c["C"], c["B"], 4
    classnames = c["names"](C)
    shots = S.lt_counts_pareto(C, seed=5)
    tokens = _tokens("a photo of a", classnames)
    image, label = S.synthetic_batch(B, C, arch.image_resolution, seed=4)
    alpha = torch.tensor(S.focal_alpha(shots, C)) if c["focal"] else None
    mp_ = S.make_cocoop_params(arch, seed=300)
    init = {"prompt_learner.ctx": sd["token_embedding.weight"][tokens[0, 1:1 + n_ctx]].clone()}
    init.update({"prompt_learner." + k: v.clone() for k, v in mp_.items()})
    trunc = int(tokens.argmax(-1).max()) + 1

    def oracle_run():
        P = {n: t.clone().requires_grad_(True) for n, t in init.items()}
        meta = tuple(P[f"prompt_learner.meta_net.{k}"] for k in ("linear1.weight", "linear1.bias", "linear2.weight",
                                                                  "linear2.bias"))
        o = O.cocoop_step(sd, tokens, image, label, P["prompt_learner.ctx"], meta, n_ctx, arch.vision_heads,
                          arch.transformer_heads, focal_alpha=alpha, use_focal=c["focal"], truncate=trunc)
        o["loss"].backward()
        return {k: v.detach() for k, v in o.items()}, {n: p.grad.detach().clone() for n, p in P.items()}

    t0 = time.time()
    if c["source"] == "reference":
        cfg = H.make_cfg("CoCoOp", **{"INPUT.SIZE": (arch.image_resolution,) * 2, "TRAINER.COCOOP.N_CTX": n_ctx,
                                      "TRAINER.COCOOP.CTX_INIT": "a photo of a", "TRAINER.COCOOP.PREC": "fp32",
                                      "DATASET.PER_CLASS_SHOTS": shots, "TRAINER.COCOOP.USE_FOCAL_LOSS": c["focal"]})
        torch.manual_seed(17)
        tm, model = H.build_reference_model("CoCoOp", cfg, classnames, sd)
        H.freeze_like_trainer("CoCoOp", model)
        assert torch.equal(model.tokenized_prompts, tokens)
        named = {n: p for n, p in model.named_parameters() if p.requires_grad}
        with torch.no_grad():
            for n, v in init.items():
                assert (named[n] - v).abs().max() < 1e-6 if n.endswith("ctx") else True
                named[n].copy_(v)
        model.eval()
        with torch.no_grad():
            logits = model(image).clone()                               # cocoop.py:257-260 eval branch
        model.train()
        opt = torch.optim.SGD(list(named.values()), lr=0.002, momentum=0.9, weight_decay=5e-4)
        loss = model(image, label)                                      # cocoop.py:323
        opt.zero_grad(); loss.backward(); opt.step()                    # noqa: E702
        grads = {n: p.grad.detach().clone() for n, p in named.items()}
        after = {n: p.detach().clone() for n, p in named.items()}
        loss = float(loss)
        print(f"[{name}] reference loss={loss:.6f} ({time.time() - t0:.1f}s)")
        o, og = oracle_run()
        worst = min(cos(og[n], grads[n]) for n in grads)
        print(f"  oracle(truncate={trunc}) vs reference: loss |d|={abs(float(o['loss']) - loss):.2e} logits max|d|="
              f"{float((o['logits'] - logits).abs().max()):.2e} worst grad cos={worst:.7f}")
        assert abs(float(o["loss"]) - loss) < 2e-5 and worst > 0.99999
        assert float((o["logits"] - logits).abs().max()) < 1e-4
    else:
        o, grads = oracle_run()
        loss, logits = float(o["loss"]), o["logits"]
        keys = list(init)
        newp, _ = O.sgd_step([init[k] for k in keys], [grads[k] for k in keys], [None] * len(keys), 0.002)
        after = dict(zip(keys, newp))
        print(f"[{name}] oracle step: loss={loss:.6f} ({time.time() - t0:.1f}s)")
    fx = {"source": np.array(c["source"]), "tokens": tokens.numpy().astype(np.int32),
          "label": label.numpy().astype(np.int32), "loss": np.float32(loss), "logits": logits.numpy(),
          "argmax": logits.argmax(1).numpy().astype(np.int32), "shots": np.array(shots, dtype=np.int32),
          "n_ctx": np.int32(n_ctx), "focal": np.int32(bool(c["focal"]))}
    _grad_entries(fx, grads, after)
    _save(name, fx)


def maple_case(name, c):
    """MaPLe at ViT-L/14 widths: the restatement with 768 -> vision width, 512 -> text width."""
    assert c["source"] == "oracle"
    arch = S.ARCHS[c["arch"]]
    sd = S.make_clip_state_dict(arch, seed=0)
    C, B, depth, n_ctx = c["C"], c["B"], c["depth"], 2
    classnames = c["names"](C)
    shots = S.lt_counts_pareto(C)
    tokens = _tokens("a photo of a", classnames)
    image, label = S.synthetic_batch(B, C, arch.image_resolution, seed=3)
    alpha = torch.tensor(S.focal_alpha(shots, C)) if c["focal"] else None
    mp_ = S.make_maple_params(arch, n_ctx, depth, seed=200)
    init = {"prompt_learner.ctx": sd["token_embedding.weight"][tokens[0, 1:1 + n_ctx]].clone()}
    init.update({"prompt_learner." + k: v.clone() for k, v in mp_.items()})
    trunc = int(tokens.argmax(-1).max()) + 1
    P = {n: t.clone().requires_grad_(True) for n, t in init.items()}
    nd = depth - 1
    t0 = time.time()
    o = O.maple_step(sd, tokens, image, label, P["prompt_learner.ctx"],
                     [P[f"prompt_learner.compound_prompts_text.{i}"] for i in range(nd)],
                     (P["prompt_learner.proj.weight"], P["prompt_learner.proj.bias"]),
                     [(P[f"prompt_learner.compound_prompt_projections.{i}.weight"],
                       P[f"prompt_learner.compound_prompt_projections.{i}.bias"]) for i in range(nd)],
                     n_ctx, arch.vision_heads, arch.transformer_heads, focal_alpha=alpha, use_focal=c["focal"],
                     truncate=trunc)
    o["loss"].backward()
    grads = {n: p.grad.detach().clone() for n, p in P.items()}
    keys = list(init)
    newp, _ = O.sgd_step([init[k] for k in keys], [grads[k] for k in keys], [None] * len(keys), 0.0035)
    after = dict(zip(keys, newp))
    logits = o["logits"].detach()
    print(f"[{name}] oracle step: loss={float(o['loss']):.6f} ({time.time() - t0:.1f}s)")
    fx = {"source": np.array("oracle"), "tokens": tokens.numpy().astype(np.int32),
          "label": label.numpy().astype(np.int32), "loss": np.float32(float(o["loss"])), "logits": logits.numpy(),
          "argmax": logits.argmax(1).numpy().astype(np.int32), "shots": np.array(shots, dtype=np.int32),
          "n_ctx": np.int32(n_ctx), "depth": np.int32(depth), "focal": np.int32(bool(c["focal"]))}
    _grad_entries(fx, grads, after)
    _save(name, fx)


RUN = {"PromptSRC": promptsrc_case, "CoOp": coop_case, "CoCoOp": cocoop_case, "MaPLe": maple_case}

if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(os.cpu_count() or 1)
    import resource
    for nm in (sys.argv[1:] or list(CASES)):
        if os.environ.get("PSRC_GOLD_SOURCE"):      # e.g. try a case on the reference itself where memory might allow
            CASES[nm] = dict(CASES[nm], source=os.environ["PSRC_GOLD_SOURCE"])
        RUN[CASES[nm]["trainer"]](nm, CASES[nm])
        print(f"[{nm}] peak RSS so far {resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1e6:.1f} GB", flush=True)
