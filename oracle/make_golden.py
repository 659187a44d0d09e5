"""TEST INFRASTRUCTURE -- generate tests/golden/*.npz from the UNMODIFIED reference and pin the oracle.

Run in the build container only (needs /root/reference):  python oracle/make_golden.py [names...]

For every case it (1) builds the reference CustomCLIP through `ref_harness` on the seeded synthetic
CLIP state-dict, (2) runs the reference forward and the reference's own `forward_backward`
(loss.backward + SGD step), (3) stores the reference outputs as the golden fixture, and (4) runs the
restatement `clip_oracle` on the same inputs and asserts that it agrees with the reference.
Also writes the token fixtures the product uses on boxes without the reference tokenizer.
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

GOLD = os.path.join(ROOT, "tests", "golden")

RAGGED_NAMES = ["cat", "golden retriever", "object 7", "a very long class name with many words", "dog",
                "airplane", "sea lion", "fire truck", "x", "traffic light"]

CASES = {
    # name: (trainer, arch, classnames, batch, prompt depth, extra cfg)
    "tiny_promptsrc": ("PromptSRC", "tiny", RAGGED_NAMES, 3, 3, {}),
    "cfg1_promptsrc": ("PromptSRC", "ViT-B/16", [f"object {i}" for i in range(100)], 4, 9, {}),
    "tiny_coop_ce": ("CoOp", "tiny", RAGGED_NAMES, 5, 0, {"TRAINER.COOP.LOSS_TYPE": "ce"}),
    "tiny_coop_focal": ("CoOp", "tiny", RAGGED_NAMES, 5, 0, {"TRAINER.COOP.LOSS_TYPE": "focal"}),
    "tiny_coop_middle": ("CoOp", "tiny", RAGGED_NAMES, 5, 0, {"TRAINER.COOP.LOSS_TYPE": "ce",
                                                              "TRAINER.COOP.CLASS_TOKEN_POSITION": "middle"}),
    "tiny_coop_front": ("CoOp", "tiny", RAGGED_NAMES, 5, 0, {"TRAINER.COOP.LOSS_TYPE": "focal",
                                                             "TRAINER.COOP.CLASS_TOKEN_POSITION": "front"}),
    # MaPLe hard-codes the ViT-B widths 768 / 512 (maple.py:144-156): full-size towers, few classes
    "vitb_maple_ce": ("MaPLe", "ViT-B/16", RAGGED_NAMES, 2, 3, {}),
    "vitb_maple_focal": ("MaPLe", "ViT-B/16", RAGGED_NAMES, 2, 3, {"TRAINER.MAPLE.USE_FOCAL_LOSS": True}),
    # CoCoOp runs unmodified on CPU in fp32
    "tiny_cocoop_ce": ("CoCoOp", "tiny", RAGGED_NAMES, 3, 0, {}),
    "tiny_cocoop_focal": ("CoCoOp", "tiny", RAGGED_NAMES, 3, 0, {"TRAINER.COCOOP.USE_FOCAL_LOSS": True}),
}


def rel(a, b):
    a, b = torch.as_tensor(a).float(), torch.as_tensor(b).float()
    return float((a - b).norm() / (b.norm() + 1e-30))


def cos(a, b):
    a, b = torch.as_tensor(a).flatten().double(), torch.as_tensor(b).flatten().double()
    return float(a @ b / (a.norm() * b.norm() + 1e-300))


def run_promptsrc(name, arch_name, classnames, batch, depth):
    arch = S.ARCHS[arch_name]
    sd = S.make_clip_state_dict(arch, seed=0)
    n_ctx = 4
    cfg = H.make_cfg("PromptSRC", **{"INPUT.SIZE": (arch.image_resolution, arch.image_resolution),
                                     "TRAINER.PROMPTSRC.PROMPT_DEPTH_VISION": depth,
                                     "TRAINER.PROMPTSRC.PROMPT_DEPTH_TEXT": depth,
                                     "OPTIM.MAX_EPOCH": 20})
    tm, model = H.build_reference_model("PromptSRC", cfg, classnames, sd)
    deep = S.make_deep_prompts(arch, n_ctx, n_ctx, depth, depth, seed=100)
    ref_sd = model.state_dict()
    with torch.no_grad():
        for k, v in deep.items():
            key = ("image_encoder." + k[len("visual."):]) if k.startswith("visual.") else ("text_encoder." + k)
            assert key in ref_sd, key
            ref_sd[key].copy_(v)
    image, label = S.synthetic_batch(batch, len(classnames), arch.image_resolution, seed=1)
    tokens = model.tokenized_prompts.clone()
    H.freeze_like_trainer("PromptSRC", model)
    model.train()
    with H.cpu_patches(half_roundtrip=True):
        out7 = model(image, label)
    ce, t_hat, fixed_n, f_zs, f_hat, zs_logits, logits = [o.detach().clone() for o in out7]
    names = [n for n, p in model.named_parameters() if p.requires_grad]
    before = {n: p.detach().clone() for n, p in model.named_parameters() if p.requires_grad}
    t0 = time.time()
    loss = H.promptsrc_forward_backward(tm, model, cfg, image, label)[0]["loss"]
    dt = time.time() - t0
    grads = {n: p.grad.detach().clone() for n, p in model.named_parameters() if p.requires_grad}
    after = {n: p.detach().clone() for n, p in model.named_parameters() if p.requires_grad}
    fixed_raw = model.prompt_learner.fixed_embeddings.detach().clone()
    print(f"[{name}] reference forward_backward: loss={loss:.6f} ({dt:.2f}s), {len(names)} trainable tensors")

    # ---- pin the restatement
    ctx = before["prompt_learner.ctx"].clone().requires_grad_(True)
    vpt = before["image_encoder.VPT"].clone().requires_grad_(True)
    dv = {i: before[f"image_encoder.transformer.resblocks.{i}.VPT_shallow"].clone().requires_grad_(True) for i in range(1, depth)}
    dtx = {i: before[f"text_encoder.transformer.resblocks.{i}.VPT_shallow"].clone().requires_grad_(True) for i in range(1, depth)}
    fixed_o = O.teacher_text_features(sd, tokens, arch.transformer_heads)
    print(f"  oracle teacher features vs reference fixed_embeddings: rel {rel(fixed_o, fixed_raw):.2e}")
    assert rel(fixed_o, fixed_raw) < 1e-5
    max_eot = int(tokens.argmax(-1).max())
    for trunc in (None, max_eot + 1):
        for t_ in [ctx, vpt, *dv.values(), *dtx.values()]:
            t_.grad = None
        o = O.promptsrc_step(sd, tokens, image, label, ctx, vpt, dv, dtx, fixed_raw, n_ctx, arch.vision_heads,
                             arch.transformer_heads, truncate=trunc)
        o["loss"].backward()
        og = {"prompt_learner.ctx": ctx.grad, "image_encoder.VPT": vpt.grad}
        og.update({f"image_encoder.transformer.resblocks.{i}.VPT_shallow": t_.grad for i, t_ in dv.items()})
        og.update({f"text_encoder.transformer.resblocks.{i}.VPT_shallow": t_.grad for i, t_ in dtx.items()})
        worst_cos = min(cos(og[n], grads[n]) for n in names)
        worst_rel = max(rel(og[n], grads[n]) for n in names)
        print(f"  oracle(truncate={trunc}) vs reference: loss |d|={abs(float(o['loss']) - loss):.2e} "
              f"logits max|d|={float((o['logits'] - logits).abs().max()):.2e} "
              f"zs_logits max|d|={float((o['zs_logits'] - zs_logits).abs().max()):.2e} "
              f"grad worst cos={worst_cos:.7f} worst rel={worst_rel:.2e}")
        assert abs(float(o["loss"]) - loss) < 2e-5 * max(1.0, abs(loss))
        assert float((o["logits"] - logits).abs().max()) < 1e-4
        assert worst_cos > 0.99999
    fx = {"tokens": tokens.numpy().astype(np.int32), "label": label.numpy().astype(np.int32),
          "loss": np.float32(loss), "ce": ce.numpy(), "logits": logits.numpy(), "zs_logits": zs_logits.numpy(),
          "text_features": t_hat.numpy(), "image_features": f_hat.numpy(), "zs_image_features": f_zs.numpy(),
          "fixed_embeddings": fixed_raw.numpy(), "argmax": logits.argmax(1).numpy().astype(np.int32),
          "classnames": np.array(classnames)}
    for n in names:
        fx["grad/" + n] = grads[n].numpy()
        fx["after/" + n] = after[n].numpy()
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **fx)


def run_coop(name, arch_name, classnames, batch, extra):
    arch = S.ARCHS[arch_name]
    sd = S.make_clip_state_dict(arch, seed=0)
    n_ctx = 16
    shots = S.lt_counts_cifar100(len(classnames), 50, 0.1)
    over = {"INPUT.SIZE": (arch.image_resolution, arch.image_resolution), "TRAINER.COOP.N_CTX": n_ctx,
            "DATASET.PER_CLASS_SHOTS": shots}
    over.update(extra)
    cfg = H.make_cfg("CoOp", **over)
    torch.manual_seed(11)  # ctx is randomly initialised by the reference (coop.py:232-233)
    tm, model = H.build_reference_model("CoOp", cfg, classnames, sd)
    H.freeze_like_trainer("CoOp", model)
    image, label = S.synthetic_batch(batch, len(classnames), arch.image_resolution, seed=2)
    tokens = model.tokenized_prompts.clone()
    ctx0 = model.prompt_learner.ctx.detach().clone()
    model.eval()
    with torch.no_grad():
        logits_ref = model(image)                                   # the REFERENCE's logits of the initial prompts
    model.train()
    opt = torch.optim.SGD(model.prompt_learner.parameters(), lr=0.002, momentum=0.9, weight_decay=5e-4)
    loss = model(image, label, None)                                # coop.py:456
    opt.zero_grad(); loss.backward(); opt.step()                    # coop.py:457-459
    model.eval()
    with torch.no_grad():
        logits_after = model(image)                                 # coop.py:465-467 (post-step logits)
    grad = model.prompt_learner.ctx.grad.detach().clone()
    after = model.prompt_learner.ctx.detach().clone()
    use_focal = extra.get("TRAINER.COOP.LOSS_TYPE") == "focal"
    alpha = torch.tensor(S.focal_alpha(shots, len(classnames))) if use_focal else None
    ctx = ctx0.clone().requires_grad_(True)
    position = extra.get("TRAINER.COOP.CLASS_TOKEN_POSITION", "end")
    assert model.prompt_learner.name_lens == (tokens.argmax(-1) - n_ctx - 2).tolist()   # name length from the EOT index
    o = O.coop_step(sd, tokens, image, label, ctx, n_ctx, arch.vision_heads, arch.transformer_heads,
                    focal_alpha=alpha, use_focal=use_focal, position=position)
    o["loss"].backward()
    print(f"[{name}] reference loss={float(loss):.6f}; oracle |d|={abs(float(o['loss']) - float(loss)):.2e} "
          f"grad cos={cos(ctx.grad, grad):.7f} rel={rel(ctx.grad, grad):.2e}")
    assert abs(float(o["loss"]) - float(loss)) < 2e-5 and cos(ctx.grad, grad) > 0.99999
    assert float((o["logits"].detach() - logits_ref).abs().max()) < 1e-4
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), tokens=tokens.numpy().astype(np.int32),
                        label=label.numpy().astype(np.int32), loss=np.float32(float(loss)), ctx0=ctx0.numpy(),
                        logits=logits_ref.numpy(), logits_after=logits_after.numpy(),
                        shots=np.array(shots, dtype=np.int32), classnames=np.array(classnames),
                        **{"grad/prompt_learner.ctx": grad.numpy(), "after/prompt_learner.ctx": after.numpy()})


def run_maple(name, arch_name, classnames, batch, depth, extra):
    arch = S.ARCHS[arch_name]
    sd = S.make_clip_state_dict(arch, seed=0)
    n_ctx = 2
    shots = S.lt_counts_cifar100(len(classnames), 50, 0.1)
    over = {"INPUT.SIZE": (arch.image_resolution, arch.image_resolution), "TRAINER.MAPLE.N_CTX": n_ctx,
            "TRAINER.MAPLE.PROMPT_DEPTH": depth, "TRAINER.MAPLE.CTX_INIT": "a photo of a",
            "DATASET.PER_CLASS_SHOTS": shots}
    over.update(extra)
    cfg = H.make_cfg("MaPLe", **over)
    torch.manual_seed(13)  # compound prompts and the coupling Linears are randomly initialised (maple.py:150-156)
    tm, model = H.build_reference_model("MaPLe", cfg, classnames, sd)
    # the two harness-side fixes without which the shipped MaPLe does not run
    model.prompt_learner.proj.float()
    model.image_encoder.transformer.init = model.text_encoder.transformer.init = False
    H.freeze_like_trainer("MaPLe", model)
    with torch.no_grad():  # seeded coupling parameters, reproducible without the reference (synthetic.make_maple_params)
        mp_ = S.make_maple_params(arch, n_ctx, depth, seed=200)
        named = dict(model.prompt_learner.named_parameters())
        for k, v in mp_.items():
            named[k].copy_(v)
    image, label = S.synthetic_batch(batch, len(classnames), arch.image_resolution, seed=3)
    tokens = model.tokenized_prompts.clone()
    pl = model.prompt_learner
    init = {n: p.detach().clone() for n, p in model.named_parameters() if p.requires_grad}
    model.eval()
    with torch.no_grad():
        logits_ref = model(image)                                   # the REFERENCE's logits of the initial prompts
    model.train()
    opt = torch.optim.SGD([p for p in model.parameters() if p.requires_grad], lr=0.0035, momentum=0.9, weight_decay=5e-4)
    loss = model(image, label)                                      # maple.py:320
    opt.zero_grad(); loss.backward(); opt.step()                    # :321-323
    grads = {n: p.grad.detach().clone() for n, p in model.named_parameters() if p.requires_grad}
    after = {n: p.detach().clone() for n, p in model.named_parameters() if p.requires_grad}
    use_focal = bool(extra.get("TRAINER.MAPLE.USE_FOCAL_LOSS", False))
    alpha = torch.tensor(S.focal_alpha(shots, len(classnames))) if use_focal else None
    P = {n: t.clone().requires_grad_(True) for n, t in init.items()}
    nd = depth - 1
    o = O.maple_step(sd, tokens, image, label, P["prompt_learner.ctx"],
                     [P[f"prompt_learner.compound_prompts_text.{i}"] for i in range(nd)],
                     (P["prompt_learner.proj.weight"], P["prompt_learner.proj.bias"]),
                     [(P[f"prompt_learner.compound_prompt_projections.{i}.weight"],
                       P[f"prompt_learner.compound_prompt_projections.{i}.bias"]) for i in range(nd)],
                     n_ctx, arch.vision_heads, arch.transformer_heads, focal_alpha=alpha, use_focal=use_focal)
    o["loss"].backward()
    worst = min(cos(P[n].grad, grads[n]) for n in grads)
    print(f"[{name}] reference loss={float(loss):.6f}; oracle |d|={abs(float(o['loss']) - float(loss)):.2e} "
          f"worst grad cos={worst:.7f} trainable={sorted(grads)}")
    assert abs(float(o["loss"]) - float(loss)) < 2e-5 and worst > 0.99999
    assert float((o["logits"].detach() - logits_ref).abs().max()) < 1e-4
    fx = {"tokens": tokens.numpy().astype(np.int32), "label": label.numpy().astype(np.int32),
          "loss": np.float32(float(loss)), "logits": logits_ref.numpy(),
          "shots": np.array(shots, dtype=np.int32), "n_ctx": np.int32(n_ctx), "depth": np.int32(depth)}
    for n in grads:
        if n.endswith(".weight"):  # [768, 512] coupling matrices: every 8th row keeps the fixture small
            fx["grad8/" + n], fx["after8/" + n] = grads[n][::8].numpy(), after[n][::8].numpy()
        else:
            fx["grad/" + n], fx["after/" + n] = grads[n].numpy(), after[n].numpy()
    fx["init/prompt_learner.ctx"] = init["prompt_learner.ctx"].numpy()
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **fx)


def run_cocoop(name, arch_name, classnames, batch, extra):
    arch = S.ARCHS[arch_name]
    sd = S.make_clip_state_dict(arch, seed=0)
    n_ctx = 4
    shots = S.lt_counts_cifar100(len(classnames), 50, 0.1)
    over = {"INPUT.SIZE": (arch.image_resolution, arch.image_resolution), "TRAINER.COCOOP.N_CTX": n_ctx,
            "TRAINER.COCOOP.CTX_INIT": "a photo of a", "TRAINER.COCOOP.PREC": "fp32",
            "DATASET.PER_CLASS_SHOTS": shots}
    over.update(extra)
    cfg = H.make_cfg("CoCoOp", **over)
    torch.manual_seed(17)
    tm, model = H.build_reference_model("CoCoOp", cfg, classnames, sd)
    H.freeze_like_trainer("CoCoOp", model)
    with torch.no_grad():  # seeded meta-net, reproducible without the reference (synthetic.make_cocoop_params)
        mp_ = S.make_cocoop_params(arch, seed=300)
        named = dict(model.prompt_learner.named_parameters())
        for k, v in mp_.items():
            named[k].copy_(v)
    image, label = S.synthetic_batch(batch, len(classnames), arch.image_resolution, seed=4)
    tokens = model.tokenized_prompts.clone()
    init = {n: p.detach().clone() for n, p in model.named_parameters() if p.requires_grad}
    model.eval()
    with torch.no_grad():
        logits_ref = model(image)                                   # the REFERENCE's logits of the initial prompts
    model.train()
    opt = torch.optim.SGD([p for p in model.parameters() if p.requires_grad], lr=0.002, momentum=0.9, weight_decay=5e-4)
    loss = model(image, label)                                      # cocoop.py:323
    opt.zero_grad(); loss.backward(); opt.step()                    # :324-326
    grads = {n: p.grad.detach().clone() for n, p in model.named_parameters() if p.requires_grad}
    after = {n: p.detach().clone() for n, p in model.named_parameters() if p.requires_grad}
    use_focal = bool(extra.get("TRAINER.COCOOP.USE_FOCAL_LOSS", False))
    alpha = torch.tensor(S.focal_alpha(shots, len(classnames))) if use_focal else None
    P = {n: t.clone().requires_grad_(True) for n, t in init.items()}
    meta = tuple(P[f"prompt_learner.meta_net.{k}"] for k in ("linear1.weight", "linear1.bias", "linear2.weight",
                                                              "linear2.bias"))
    o = O.cocoop_step(sd, tokens, image, label, P["prompt_learner.ctx"], meta, n_ctx, arch.vision_heads,
                      arch.transformer_heads, focal_alpha=alpha, use_focal=use_focal)
    o["loss"].backward()
    worst = min(cos(P[n].grad, grads[n]) for n in grads)
    print(f"[{name}] reference loss={float(loss):.6f}; oracle |d|={abs(float(o['loss']) - float(loss)):.2e} "
          f"worst grad cos={worst:.7f} trainable={sorted(grads)}")
    assert abs(float(o["loss"]) - float(loss)) < 2e-5 and worst > 0.99999
    assert float((o["logits"].detach() - logits_ref).abs().max()) < 1e-4
    fx = {"tokens": tokens.numpy().astype(np.int32), "label": label.numpy().astype(np.int32),
          "loss": np.float32(float(loss)), "logits": logits_ref.numpy(),
          "shots": np.array(shots, dtype=np.int32), "n_ctx": np.int32(n_ctx)}
    for n in grads:
        fx["init/" + n], fx["grad/" + n], fx["after/" + n] = init[n].numpy(), grads[n].numpy(), after[n].numpy()
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **fx)


def write_token_fixtures():
    os.makedirs(S.DATA_DIR, exist_ok=True)
    toks = H.tokenize([f"a photo of a object {i}." for i in range(1000)]).numpy().astype(np.int32)
    np.savez_compressed(S.DATA_DIR / "tokens_object1000.npz", tokens=toks)
    xs = H.tokenize([" ".join(["X"] * 16) + f" object {i}." for i in range(1000)]).numpy().astype(np.int32)
    np.savez_compressed(S.DATA_DIR / "tokens_x16_object1000.npz", tokens=xs)
    print("token fixtures: EOT index range", toks.argmax(-1).min(), toks.argmax(-1).max(), "| X16:",
          xs.argmax(-1).min(), xs.argmax(-1).max())


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(os.cpu_count() or 1)
    want = sys.argv[1:] or ["tokens", *CASES]
    if "tokens" in want:
        write_token_fixtures()
    for nm in want:
        if nm == "tokens":
            continue
        trainer, arch_name, classnames, batch, depth, extra = CASES[nm]
        if trainer == "PromptSRC":
            run_promptsrc(nm, arch_name, classnames, batch, depth)
        elif trainer == "CoOp":
            run_coop(nm, arch_name, classnames, batch, extra)
        elif trainer == "MaPLe":
            run_maple(nm, arch_name, classnames, batch, depth, extra)
        elif trainer == "CoCoOp":
            run_cocoop(nm, arch_name, classnames, batch, extra)
    print("golden fixtures written to", GOLD)
