"""Checks of the Dassl-facing drop-in (psrc_b200.trainers) against the oracle: fused trainer step, autograd
bridge through CustomCLIP.forward, eval logits with cached text features, checkpoint round trip, CoOp step.
`python tests/trainer_checks.py` prints the metrics on a GPU box; tests/test_trainers_gpu.py asserts them."""
from __future__ import annotations

import os
import sys
import tempfile

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

from helpers import FakeDM, cos, trainer_cfg  # noqa: E402

NAME_LENS = [1, 2, 3, 8, 1, 1, 2, 2, 1, 2]
DEPTH, N_CTX, B = 3, 4, 3


def make_trainer(name="PromptSRC", prec="fp32", device="cuda", **over):
    import psrc_b200  # noqa: F401
    from psrc_b200 import synthetic as S, trainers as T
    tokens = torch.as_tensor(S.synthetic_tokens(len(NAME_LENS), name_lens=NAME_LENS))
    cfg = trainer_cfg(name, batch=B, depth=DEPTH, prec=prec, **over)
    dv = DEPTH if name == "PromptSRC" else 0
    T.CLIP_MODEL_FACTORY = lambda cfg_, nm: T.synthetic_clip("tiny", 0, N_CTX if dv else 0, N_CTX, dv, DEPTH if dv else 1)
    T.TOKEN_FACTORY = lambda cfg_, names: tokens
    cls = T.TRAINER_REGISTRY.get(name)
    if T.HAVE_DASSL:  # Dassl's SimpleTrainer.__init__ builds a DataManager: bypass it, keep every other hook
        tr = cls.__new__(cls)
        T.TrainerX.__mro__[-2].__init__(tr)  # TrainerBase.__init__
        tr.cfg, tr.dm, tr.device = cfg, FakeDM([f"c{i}" for i in range(len(NAME_LENS))]), torch.device(device)
        tr.start_epoch = tr.epoch = 0
        tr.max_epoch = cfg.OPTIM.MAX_EPOCH
        tr.check_cfg(cfg)
        tr.build_model()
    else:
        tr = cls(cfg, dm=FakeDM([f"c{i}" for i in range(len(NAME_LENS))]), device=device)
    tr.batch_idx, tr.num_batches = 0, 2
    return tr, cfg, tokens


def oracle_promptsrc(tokens, image, label, steps=1, lrs=(1e-5,)):
    """`steps` reference PromptSRC iterations (oracle restatement + torch SGD) from the synthetic init."""
    import clip_oracle as O
    from psrc_b200 import synthetic as S
    arch = S.ARCHS["tiny"]
    sd = S.make_clip_state_dict(arch, seed=0)
    deep = S.make_deep_prompts(arch, N_CTX, N_CTX, DEPTH, DEPTH, seed=100)
    fixed = O.teacher_text_features(sd, tokens, arch.transformer_heads)
    P = {"prompt_learner.ctx": sd["token_embedding.weight"][tokens[0, 1:1 + N_CTX]].clone(),
         "image_encoder.VPT": deep["visual.VPT"].clone()}
    for i in range(1, DEPTH):
        P[f"image_encoder.transformer.resblocks.{i}.VPT_shallow"] = deep[f"visual.transformer.resblocks.{i}.VPT_shallow"].clone()
        P[f"text_encoder.transformer.resblocks.{i}.VPT_shallow"] = deep[f"transformer.resblocks.{i}.VPT_shallow"].clone()
    for v in P.values():
        v.requires_grad_(True)
    opt = torch.optim.SGD(list(P.values()), lr=lrs[0], momentum=0.9, weight_decay=5e-4)
    out = []
    for s in range(steps):
        for g in opt.param_groups:
            g["lr"] = lrs[min(s, len(lrs) - 1)]
        dv = {i: P[f"image_encoder.transformer.resblocks.{i}.VPT_shallow"] for i in range(1, DEPTH)}
        dt = {i: P[f"text_encoder.transformer.resblocks.{i}.VPT_shallow"] for i in range(1, DEPTH)}
        o = O.promptsrc_step(sd, tokens, image, label, P["prompt_learner.ctx"], P["image_encoder.VPT"], dv, dt, fixed,
                             N_CTX, arch.vision_heads, arch.transformer_heads)
        opt.zero_grad()
        o["loss"].backward()
        grads = {k: v.grad.clone() for k, v in P.items()}
        opt.step()
        out.append({"loss": float(o["loss"].detach()), "logits": o["logits"].detach().clone(), "grads": grads,
                    "params": {k: v.detach().clone() for k, v in P.items()}, "o": o})
    return out, sd, fixed


def check_promptsrc_trainer(prec="fp32"):
    """Two fused trainer steps (eager, then CUDA-graph replay) vs two oracle iterations."""
    from psrc_b200 import synthetic as S
    tr, cfg, tokens = make_trainer("PromptSRC", prec)
    image, label = S.synthetic_batch(B, tokens.shape[0], 64, seed=1)
    lr0 = tr.optim.param_groups[0]["lr"]
    ref, _, _ = oracle_promptsrc(tokens, image, label, steps=3, lrs=(lr0,))
    m = {}
    batch = {"img": image.pin_memory(), "label": label.pin_memory()}
    for s in range(3):
        tr.batch_idx = 0
        out = tr.forward_backward(batch)
        m[f"loss_abs_{s}"] = abs(out["loss"] - ref[s]["loss"])
    named = dict(tr.model.named_parameters())
    m["graph_captured"] = getattr(tr.model._step, "_graph", None) is not None
    worst = 0.0
    for k, v in ref[2]["params"].items():
        worst = max(worst, float((named[k].detach().cpu() - v).abs().max()))
    m["param_after3_maxabs"] = worst
    upd = torch.cat([(named[k].detach().cpu() - ref[0]["params"][k]).flatten() for k in ref[2]["params"]])
    upd_ref = torch.cat([(ref[2]["params"][k] - ref[0]["params"][k]).flatten() for k in ref[2]["params"]])
    m["update_cos"] = cos(upd, upd_ref)
    # epoch end: scheduler stepped through update_lr (constant warm-up 1e-5 -> base lr at epoch 1)
    tr.batch_idx = tr.num_batches - 1
    tr.forward_backward(batch)
    m["lr_after_epoch"] = tr.optim.param_groups[0]["lr"]
    tr.batch_idx = 0
    tr.forward_backward(batch)           # the new lr reaches the device-resident scalar the captured graph reads
    m["lr_device"] = float(tr.model._step.lr_dev.item())
    return m


def check_bridge(prec="fp32"):
    """CustomCLIP.forward 7-tuple + the reference trainer's own loss composition (promptsrc.py:299-324) through
    autograd: gradients must equal the oracle's."""
    from psrc_b200 import synthetic as S
    tr, cfg, tokens = make_trainer("PromptSRC", prec)
    image, label = S.synthetic_batch(B, tokens.shape[0], 64, seed=1)
    ref, _, _ = oracle_promptsrc(tokens, image, label, steps=1)
    model = tr.model
    model.train()
    loss_ce, t_hat, t_zs, f_zs, f_hat, zs_logits, logits = model(image.cuda(), label.cuda())
    l_text = F.l1_loss(t_hat, t_zs, reduction="mean") * 25
    l_img = F.l1_loss(f_hat, f_zs, reduction="mean") * 10
    l_kl = F.kl_div(F.log_softmax(logits, 1), F.log_softmax(zs_logits, 1), reduction="sum", log_target=True) / logits.numel()
    loss = loss_ce + l_kl + l_text + l_img
    tr.optim.zero_grad()
    loss.backward()
    m = {"loss_abs": abs(float(loss) - ref[0]["loss"]),
         "logits_maxabs": float((logits.detach().cpu() - ref[0]["logits"]).abs().max())}
    named = dict(model.named_parameters())
    m["gradcos_worst"] = min(cos(named[k].grad, g) for k, g in ref[0]["grads"].items())
    m["n_trainable"] = sum(p.requires_grad for p in model.parameters())
    # torch-optimizer-style step through FlatSGD (grads were re-created by autograd after zero_grad)
    tr.optim.step()
    worst = max(float((named[k].detach().cpu() - v).abs().max()) for k, v in ref[0]["params"].items())
    m["param_after_maxabs"] = worst
    return m


def check_eval_and_checkpoint(prec="fp32"):
    from psrc_b200 import synthetic as S
    import clip_oracle as O
    tr, cfg, tokens = make_trainer("PromptSRC", prec)
    arch = S.ARCHS["tiny"]
    image, _ = S.synthetic_batch(7, tokens.shape[0], 64, seed=5)     # 7 = one full test batch of 5 + a ragged 2
    sd = S.make_clip_state_dict(arch, seed=0)
    m = {}
    # checkpoint round trip: save, perturb, load -> parameters restored; keys carry the reference names
    with tempfile.TemporaryDirectory() as d:
        tr.save_model(0, d)
        keys = set(torch.load(os.path.join(d, "VLPromptLearner", "model.pth.tar-1"), weights_only=False)["state_dict"])
        before = tr.model._step.flat_param.clone()
        tr.model._step.flat_param.add_(1.0)
        tr.load_model(d, epoch=1)
        m["ckpt_restore_maxabs"] = float((tr.model._step.flat_param - before).abs().max())
    # slim checkpoint (TRAINER.B200.SLIM_CHECKPOINT): trainable tensors only, still loadable
    with tempfile.TemporaryDirectory() as d:
        tr.cfg.TRAINER.B200["SLIM_CHECKPOINT"] = True
        tr.save_model(0, d)
        tr.cfg.TRAINER.B200["SLIM_CHECKPOINT"] = False
        f = os.path.join(d, "VLPromptLearner", "model.pth.tar-1")
        slim = torch.load(f, weights_only=False)["state_dict"]
        m["slim_keys"] = sorted(slim)
        m["slim_bytes"], m["full_state_numel"] = os.path.getsize(f), sum(v.numel() for v in tr.model.state_dict().values())
        before = tr.model._step.flat_param.clone()
        tr.model._step.flat_param.mul_(0.5)
        tr.load_model(d, epoch=1)
        m["slim_restore_maxabs"] = float((tr.model._step.flat_param - before).abs().max())
        m["state_dict_method_restored"] = len(tr.model.state_dict()) > len(slim)
    need = {"prompt_learner.ctx", "prompt_learner.token_prefix", "prompt_learner.token_suffix", "image_encoder.VPT",
            "image_encoder.transformer.resblocks.1.VPT_shallow", "text_encoder.transformer.resblocks.2.VPT_shallow",
            "image_encoder.conv1.weight", "text_encoder.text_projection", "logit_scale",
            "text_encoder.transformer.resblocks.0.attn.in_proj_weight"}
    m["missing_state_keys"] = sorted(need - keys)
    # eval logits (initial prompts) vs the oracle's prompted forward
    tr.set_model_mode("eval")
    with torch.no_grad():
        lg = tr.model_inference(image.cuda()).cpu()
    deep = S.make_deep_prompts(arch, N_CTX, N_CTX, DEPTH, DEPTH, seed=100)
    dv = {i: deep[f"visual.transformer.resblocks.{i}.VPT_shallow"] for i in range(1, DEPTH)}
    dt = {i: deep[f"transformer.resblocks.{i}.VPT_shallow"] for i in range(1, DEPTH)}
    emb = sd["token_embedding.weight"][tokens]
    ctx = sd["token_embedding.weight"][tokens[0, 1:1 + N_CTX]]
    with torch.no_grad():
        t = O.normalize(O.text_tower(O.construct_prompts(emb[:, :1], ctx, emb[:, 1 + N_CTX:]), tokens.argmax(-1), sd,
                                     arch.transformer_heads, dt, N_CTX))
        f = O.normalize(O.vision_tower(image, sd, arch.vision_heads, deep["visual.VPT"], dv))
        ref_logits = sd["logit_scale"].exp() * f @ t.t()
    m["eval_logits_maxabs"] = float((lg - ref_logits).abs().max())
    m["eval_argmax_equal"] = bool((lg.argmax(1) == ref_logits.argmax(1)).all())
    m["eval_shape_ok"] = tuple(lg.shape) == (7, tokens.shape[0])
    return m



def check_gpa(prec="fp32"):
    """USE_GPA = True over 3 epochs x 2 batches (promptsrc.py:335-360): after every epoch the prompts enter the running
    Gaussian-weighted sum (`state_dict_weighting` / `state_dict_add`, :363-378), after the last one the sum replaces
    the model.  Restated here on the oracle's parameters after the same SGD steps with the trainer's own lr schedule;
    the frozen tensors are scaled by weights that sum to 1, i.e. unchanged, so only the prompts are compared."""
    import numpy as np
    from psrc_b200 import synthetic as S
    epochs, nb = 3, 2
    tr, cfg, tokens = make_trainer("PromptSRC", prec, **{"TRAINER.PROMPTSRC.USE_GPA": True, "OPTIM.MAX_EPOCH": epochs})
    tr.num_batches = nb
    image, label = S.synthetic_batch(B, tokens.shape[0], 64, seed=1)
    batch = {"img": image.pin_memory(), "label": label.pin_memory()}
    lrs, losses = [], []
    for e in range(epochs):
        for b in range(nb):
            tr.epoch, tr.batch_idx = e, b
            lrs.append(tr.optim.param_groups[0]["lr"])
            losses.append(tr.forward_backward(batch)["loss"])
    ref, _, _ = oracle_promptsrc(tokens, image, label, steps=epochs * nb, lrs=tuple(lrs))
    mean, std = cfg.TRAINER.PROMPTSRC.GPA_MEAN, cfg.TRAINER.PROMPTSRC.GPA_STD
    g = np.array([1 / (std * np.sqrt(2 * np.pi)) * np.exp(-0.5 * ((a - mean) / std) ** 2) for a in range(1, epochs + 1)])
    g = g / g.sum()
    gpa = {k: sum(float(g[e]) * ref[(e + 1) * nb - 1]["params"][k] for e in range(epochs)) for k in ref[0]["params"]}
    named = dict(tr.model.named_parameters())
    m = {"loss_abs_max": max(abs(l - r["loss"]) for l, r in zip(losses, ref)),
         "lrs": lrs,
         "gpa_param_maxabs": max(float((named[k].detach().cpu() - v).abs().max()) for k, v in gpa.items()),
         "last_vs_gpa_maxabs": max(float((ref[-1]["params"][k] - v).abs().max()) for k, v in gpa.items()),
         "step_counter": tr.step_counter}
    # the evaluation path must see the aggregated prompts (cached text features dropped by params_changed)
    import clip_oracle as O
    arch = S.ARCHS["tiny"]
    sd = S.make_clip_state_dict(arch, seed=0)
    tr.set_model_mode("eval")
    img7, _ = S.synthetic_batch(7, tokens.shape[0], 64, seed=5)
    with torch.no_grad():
        lg = tr.model_inference(img7.cuda()).cpu()
        dv = {i: gpa[f"image_encoder.transformer.resblocks.{i}.VPT_shallow"] for i in range(1, DEPTH)}
        dt = {i: gpa[f"text_encoder.transformer.resblocks.{i}.VPT_shallow"] for i in range(1, DEPTH)}
        emb = sd["token_embedding.weight"][tokens]
        t = O.normalize(O.text_tower(O.construct_prompts(emb[:, :1], gpa["prompt_learner.ctx"], emb[:, 1 + N_CTX:]),
                                     tokens.argmax(-1), sd, arch.transformer_heads, dt, N_CTX))
        f = O.normalize(O.vision_tower(img7, sd, arch.vision_heads, gpa["image_encoder.VPT"], dv))
        m["eval_logits_maxabs"] = float((lg - sd["logit_scale"].exp() * f @ t.t()).abs().max())
    return m


def check_coop_eval_follows_prompt_learner_load():
    """Dassl reaches CoOp's prompts through the registered `prompt_learner` only.  eval -> load other
    prompts through `prompt_learner.load_state_dict` (what `load_model` does before the final test) -> eval must score
    the LOADED prompts, not the cached text features of the previous ones."""
    from psrc_b200 import synthetic as S
    tr, cfg, tokens = make_trainer("CoOp", "fp32")
    img7, _ = S.synthetic_batch(7, tokens.shape[0], 64, seed=5)
    tr.set_model_mode("eval")
    with torch.no_grad():
        lg0 = tr.model_inference(img7.cuda()).clone()
        sd_pl = {k: v.clone() for k, v in tr.model.prompt_learner.state_dict().items()}
        sd_pl["ctx"] = sd_pl["ctx"] + 0.05 * torch.randn_like(sd_pl["ctx"])
        tr.model.prompt_learner.load_state_dict(sd_pl)
        lg1 = tr.model_inference(img7.cuda()).clone()
        # ground truth for the loaded prompts: a fresh text pass
        tr.model._infer.text_valid = False
        lg2 = tr.model_inference(img7.cuda()).clone()
    return {"changed_maxabs": float((lg1 - lg0).abs().max()), "stale_maxabs": float((lg1 - lg2).abs().max())}


def _seed_trainer_params(tr, values):
    """Overwrite the trainer's (randomly initialised) prompt-learner parameters with seeded values."""
    named = dict(tr.model.prompt_learner.named_parameters())
    with torch.no_grad():
        for k, v in values.items():
            named[k].copy_(v.to(named[k].device))


def check_maple(prec="fp32", focal=False):
    """MaPLe trainer: 3 fused steps (eager, then graph replays) vs 3 oracle iterations with torch SGD; eval logits."""
    import clip_oracle as O
    from psrc_b200 import synthetic as S
    shots = [50, 40, 30, 20, 10, 5, 5, 3, 2, 1]
    tr, cfg, tokens = make_trainer("MaPLe", prec, **{"TRAINER.MAPLE.USE_FOCAL_LOSS": focal,
                                                     "DATASET.PER_CLASS_SHOTS": shots})
    arch, n_ctx = S.ARCHS["tiny"], 2
    sd = S.make_clip_state_dict(arch, seed=0)
    mp_ = S.make_maple_params(arch, n_ctx, DEPTH, seed=200)
    _seed_trainer_params(tr, mp_)
    image, label = S.synthetic_batch(B, tokens.shape[0], 64, seed=1)
    P = {k: v.clone().requires_grad_(True) for k, v in mp_.items()}
    P["ctx"] = sd["token_embedding.weight"][tokens[0, 1:1 + n_ctx]].clone().requires_grad_(True)
    lr = tr.optim.param_groups[0]["lr"]
    opt = torch.optim.SGD(list(P.values()), lr=lr, momentum=0.9, weight_decay=5e-4)
    alpha = torch.tensor(S.focal_alpha(shots, len(shots))) if focal else None
    nd = DEPTH - 1
    m, batch = {}, {"img": image.pin_memory(), "label": label.pin_memory()}
    for s_ in range(3):
        o = O.maple_step(sd, tokens, image, label, P["ctx"], [P[f"compound_prompts_text.{i}"] for i in range(nd)],
                         (P["proj.weight"], P["proj.bias"]),
                         [(P[f"compound_prompt_projections.{i}.weight"], P[f"compound_prompt_projections.{i}.bias"])
                          for i in range(nd)], n_ctx, arch.vision_heads, arch.transformer_heads, focal_alpha=alpha,
                         use_focal=focal)
        opt.zero_grad()
        o["loss"].backward()
        opt.step()
        tr.batch_idx = 0
        out = tr.forward_backward(batch)
        m[f"loss_abs_{s_}"] = abs(out["loss"] - float(o["loss"].detach()))
    named = dict(tr.model.prompt_learner.named_parameters())
    m["param_after3_maxabs"] = max(float((named[k].detach().cpu() - v.detach()).abs().max()) for k, v in P.items())
    m["graph_captured"] = getattr(tr.model._step, "_graph", None) is not None
    m["registered"] = tr.get_model_names()
    m["n_trainable"] = sum(p.requires_grad for p in tr.model.parameters())
    keys = set(tr.model.state_dict())
    m["missing_state_keys"] = sorted({"prompt_learner.ctx", "prompt_learner.proj.weight", "prompt_learner.proj.bias",
                                      "prompt_learner.compound_prompts_text.0",
                                      "prompt_learner.compound_prompt_projections.1.weight",
                                      "prompt_learner.token_prefix", "image_encoder.conv1.weight"} - keys)
    # eval: logits of the UPDATED prompts through the cached-text path vs the oracle
    tr.set_model_mode("eval")
    img7, _ = S.synthetic_batch(7, tokens.shape[0], 64, seed=5)
    with torch.no_grad():
        lg = tr.model_inference(img7.cuda()).cpu()
        o = O.maple_step(sd, tokens, img7, torch.zeros(7, dtype=torch.long), P["ctx"],
                         [P[f"compound_prompts_text.{i}"] for i in range(nd)], (P["proj.weight"], P["proj.bias"]),
                         [(P[f"compound_prompt_projections.{i}.weight"], P[f"compound_prompt_projections.{i}.bias"])
                          for i in range(nd)], n_ctx, arch.vision_heads, arch.transformer_heads)
    m["eval_logits_maxabs"] = float((lg - o["logits"]).abs().max())
    m["eval_argmax_equal"] = bool((lg.argmax(1) == o["logits"].argmax(1)).all())
    return m


def check_cocoop(prec="fp32", focal=False):
    """CoCoOp trainer: 3 fused steps vs 3 oracle iterations; eval logits (per-image text passes, ragged test batch)."""
    import clip_oracle as O
    from psrc_b200 import synthetic as S
    shots = [50, 40, 30, 20, 10, 5, 5, 3, 2, 1]
    tr, cfg, tokens = make_trainer("CoCoOp", prec, **{"TRAINER.COCOOP.USE_FOCAL_LOSS": focal,
                                                      "DATASET.PER_CLASS_SHOTS": shots})
    arch, n_ctx = S.ARCHS["tiny"], 4
    sd = S.make_clip_state_dict(arch, seed=0)
    mp_ = S.make_cocoop_params(arch, seed=300)
    _seed_trainer_params(tr, mp_)
    image, label = S.synthetic_batch(B, tokens.shape[0], 64, seed=1)
    P = {k: v.clone().requires_grad_(True) for k, v in mp_.items()}
    P["ctx"] = sd["token_embedding.weight"][tokens[0, 1:1 + n_ctx]].clone().requires_grad_(True)
    meta = lambda: tuple(P["meta_net." + k] for k in ("linear1.weight", "linear1.bias", "linear2.weight",  # noqa: E731
                                                      "linear2.bias"))
    lr = tr.optim.param_groups[0]["lr"]
    opt = torch.optim.SGD(list(P.values()), lr=lr, momentum=0.9, weight_decay=5e-4)
    alpha = torch.tensor(S.focal_alpha(shots, len(shots))) if focal else None
    m, batch = {}, {"img": image.pin_memory(), "label": label.pin_memory()}
    for s_ in range(3):
        o = O.cocoop_step(sd, tokens, image, label, P["ctx"], meta(), n_ctx, arch.vision_heads,
                          arch.transformer_heads, focal_alpha=alpha, use_focal=focal)
        opt.zero_grad()
        o["loss"].backward()
        opt.step()
        tr.batch_idx = 0
        out = tr.forward_backward(batch)
        m[f"loss_abs_{s_}"] = abs(out["loss"] - float(o["loss"].detach()))
    named = dict(tr.model.prompt_learner.named_parameters())
    m["param_after3_maxabs"] = max(float((named[k].detach().cpu() - v.detach()).abs().max()) for k, v in P.items())
    m["graph_captured"] = getattr(tr.model._step, "_graph", None) is not None
    m["registered"] = tr.get_model_names()
    m["n_trainable"] = sum(p.requires_grad for p in tr.model.parameters())
    tr.set_model_mode("eval")
    img7, _ = S.synthetic_batch(7, tokens.shape[0], 64, seed=5)     # 7 = two chunks of the train batch (3) + a ragged 1
    with torch.no_grad():
        lg = tr.model_inference(img7.cuda()).cpu()
        o = O.cocoop_step(sd, tokens, img7, torch.zeros(7, dtype=torch.long), P["ctx"], meta(), n_ctx,
                          arch.vision_heads, arch.transformer_heads)
    m["eval_logits_maxabs"] = float((lg - o["logits"]).abs().max())
    m["eval_argmax_equal"] = bool((lg.argmax(1) == o["logits"].argmax(1)).all())
    m["eval_shape_ok"] = tuple(lg.shape) == (7, tokens.shape[0])
    return m


def check_teacher_ensemble():
    """Frozen-teacher precompute through the kernels: single template and a 3-template ensemble
    (promptsrc_pets_focal.py:165-175) against the oracle's frozen text tower."""
    import clip_oracle as O
    from psrc_b200 import synthetic as S, trainers as T
    arch = S.ARCHS["tiny"]
    sd = S.make_clip_state_dict(arch, seed=0)
    lens = [[1, 2, 3, 8, 1, 1, 2, 2, 1, 2], [2, 2, 3, 5, 1, 4, 2, 2, 1, 2], [1, 1, 1, 9, 3, 1, 2, 6, 1, 2]]
    templ = torch.stack([torch.as_tensor(S.synthetic_tokens(10, name_lens=nl, seed=7 + i)) for i, nl in enumerate(lens)])
    ref = torch.stack([O.teacher_text_features(sd, t_, arch.transformer_heads) for t_ in templ]).mean(0)
    cfg = trainer_cfg("PromptSRC", batch=B, depth=DEPTH, prec="fp32")
    clip_model = T.synthetic_clip("tiny", 0, N_CTX, N_CTX, DEPTH, DEPTH)
    model = T.PromptSRCCustomCLIP(cfg, [f"c{i}" for i in range(10)], clip_model, tokenized_prompts=templ[0],
                                  teacher_tokens=templ, device="cuda")
    got = model.prompt_learner.fixed_embeddings.cpu()
    single = T.PromptSRCCustomCLIP(cfg, [f"c{i}" for i in range(10)], clip_model, tokenized_prompts=templ[0],
                                   teacher_tokens=templ[0], device="cuda").prompt_learner.fixed_embeddings.cpu()
    ref1 = O.teacher_text_features(sd, templ[0], arch.transformer_heads)
    return {"ensemble_rel": float((got - ref).norm() / ref.norm()), "single_rel": float((single - ref1).norm() / ref1.norm())}


def check_text_ahead():
    """The text-tower prefix enqueued ahead of the next step (step.PromptedClipStep._replay) must not change results:
    a trainer with it and one without take the same steps, including across an in-place parameter edit made from
    outside (what load_state_dict on a sub-module does), which has to invalidate the prefix computed ahead."""
    from psrc_b200 import synthetic as S
    out = {}
    finals, losses = [], []
    for ahead in (True, False):
        tr, cfg, tokens = make_trainer("PromptSRC", "fp32")
        tr.model._step.text_ahead = ahead
        image, label = S.synthetic_batch(B, tokens.shape[0], 64, seed=1)
        batch = {"img": image.pin_memory(), "label": label.pin_memory()}
        ls = []
        for s in range(6):
            if s == 4:
                with torch.no_grad():
                    tr.model.prompt_learner.ctx.mul_(1.05)
            tr.batch_idx = 0
            ls.append(tr.forward_backward(batch)["loss"])
        out[f"ahead_flag_{ahead}"] = bool(tr.model._step._text_ahead)
        finals.append(torch.cat([p.detach().flatten().cpu() for _, p in sorted(tr.model.named_parameters()) if p.requires_grad]))
        losses.append(ls)
    out["param_maxabs"] = float((finals[0] - finals[1]).abs().max())
    out["loss_maxabs"] = max(abs(a - b) for a, b in zip(*losses))
    out["loss_jump_after_edit"] = abs(losses[0][4] - losses[0][3])
    # the look-ahead loader (next batch copied over under the current step) against plain per-batch calls
    from psrc_b200 import trainers as T
    finals, losses = [], []
    for look in (True, False):
        tr, cfg, tokens = make_trainer("PromptSRC", "fp32")
        batches = []
        for i in range(5):
            image, label = S.synthetic_batch(B, tokens.shape[0], 64, seed=10 + i)
            batches.append({"img": image.pin_memory(), "label": label.pin_memory()})
        ls = []
        for b in (T.LookaheadLoader(batches) if look else batches):
            tr.batch_idx = 0
            ls.append(tr.forward_backward(b)["loss"])
        finals.append(torch.cat([p.detach().flatten().cpu() for _, p in sorted(tr.model.named_parameters()) if p.requires_grad]))
        losses.append(ls)
        if look:
            out["staged_steps_used"] = getattr(tr.model._step, "image_next", None) is not None
    out["lookahead_param_maxabs"] = float((finals[0] - finals[1]).abs().max())
    out["lookahead_loss_maxabs"] = max(abs(a - b) for a, b in zip(*losses))
    out["lookahead_losses_distinct"] = len(set(round(x, 6) for x in losses[0])) > 1
    return out


def check_coop(prec="fp32", loss_type="ce"):
    import clip_oracle as O
    from psrc_b200 import synthetic as S
    shots = [50, 40, 30, 20, 10, 5, 5, 3, 2, 1]
    tr, cfg, tokens = make_trainer("CoOp", prec, **{"TRAINER.COOP.LOSS_TYPE": loss_type, "DATASET.PER_CLASS_SHOTS": shots})
    arch = S.ARCHS["tiny"]
    sd = S.make_clip_state_dict(arch, seed=0)
    image, label = S.synthetic_batch(B, tokens.shape[0], 64, seed=1)
    ctx0 = tr.model.prompt_learner.ctx.detach().cpu().clone()
    ctx = ctx0.clone().requires_grad_(True)
    alpha = torch.tensor(S.focal_alpha(shots, len(shots)))
    o = O.coop_step(sd, tokens, image, label, ctx, N_CTX, arch.vision_heads, arch.transformer_heads,
                    focal_alpha=alpha, use_focal=loss_type == "focal")
    o["loss"].backward()
    lr = tr.optim.param_groups[0]["lr"]
    ctx1 = ctx0 - lr * (ctx.grad + 5e-4 * ctx0)
    out = tr.forward_backward({"img": image, "label": label})
    m = {"loss_abs": abs(out["loss"] - float(o["loss"])),
         "ctx_after_maxabs": float((tr.model.prompt_learner.ctx.detach().cpu() - ctx1).abs().max()),
         "gradcos": cos(tr.model._step.grads["ctx"], ctx.grad),
         "registered": tr.get_model_names()}
    if loss_type == "ce":
        with torch.no_grad():
            lg = O.clip_logits(sd, tokens, image, ctx1, N_CTX, arch.vision_heads, arch.transformer_heads)
        m["acc_abs"] = abs(out["acc"] - float((lg.argmax(1) == label).float().mean() * 100))
    return m


if __name__ == "__main__":
    for prec in ("fp32", "fp16"):
        for fn in (check_promptsrc_trainer, check_bridge, check_eval_and_checkpoint, check_coop):
            print(f"== {fn.__name__} [{prec}]", fn(prec))
    print("== check_coop focal [fp32]", check_coop("fp32", "focal"))
    for prec in ("fp32", "fp16"):
        for focal in (False, True):
            print(f"== check_maple focal={focal} [{prec}]", check_maple(prec, focal))
            print(f"== check_cocoop focal={focal} [{prec}]", check_cocoop(prec, focal))
