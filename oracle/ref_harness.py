"""TEST INFRASTRUCTURE -- runs the UNMODIFIED reference (read-only /root/reference) in-process.

Only usable in the build container (the GPU box has no /root/reference).  It is how the
oracle restatement (`clip_oracle.py`) is pinned and how `tests/golden/` is generated
(`make_golden.py`).  Nothing here is imported by the product.

Harness-side patches (SURVEY.md section 0 recipe; none touches the reference tree):
  1. sys.path <- PromptSRC, Dassl.pytorch
  2. stub modules for absent third-party imports: ftfy, gdown, wilds, timm, yacs
  3. trainers.<x>.load_clip_to_cpu -> clip.model.build_model(deepcopy(state_dict), design_details)
     (the shipped loader downloads weights, clip/clip.py:39-68)
  4. Tensor.cuda / Module.cuda -> identity (CPU box; trainers/promptsrc.py:116,127,201)
  5. PromptSRC only: Tensor.half -> dtype-preserving fp16 round trip.  With the model in fp32
     (clip/model.py:699 convert_weights commented out) torch.cat(fp32, fp16) promotes to fp32, so
     this is literally what the executing lines compute, and it lets promptsrc.py:201 run (F3).
  6. PromptSRC: cfg keys TRAINER.PROMPTSRC.LOGITS_LOSS_WEIGHT = 1.0, USE_GPA = False (read at
     promptsrc.py:321,335 but never defined by train.py:125-141).
"""
from __future__ import annotations

import contextlib
import copy
import os
import sys
import types

import torch

REF_ROOT = os.environ.get("PSRC_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "PromptSRC", "trainers"))


class _CfgNode(dict):
    """Minimal stand-in for yacs.config.CfgNode (attribute access, clone, merge_from_list)."""

    def __init__(self, init=None, *a, **k):
        super().__init__()
        if init:
            for key, v in init.items():
                self[key] = _CfgNode(v) if isinstance(v, dict) and not isinstance(v, _CfgNode) else v

    def __getattr__(self, name):
        try:
            return self[name]
        except KeyError:
            raise AttributeError(name)

    def __setattr__(self, name, value):
        self[name] = value

    def clone(self):
        return copy.deepcopy(self)

    def freeze(self):
        pass

    def defrost(self):
        pass

    def is_frozen(self):
        return False

    def merge_from_list(self, lst):
        for k, v in zip(lst[0::2], lst[1::2]):
            node = self
            parts = k.split(".")
            for p in parts[:-1]:
                node = node[p]
            node[parts[-1]] = v

    def merge_from_file(self, path):
        raise NotImplementedError("yaml configs are not needed by the harness")


def _install_stubs():
    def mod(name, **attrs):
        if name in sys.modules:
            return sys.modules[name]
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    mod("ftfy", fix_text=lambda s: s)
    mod("gdown", download=lambda *a, **k: None)
    w = mod("wilds")
    w.get_dataset = lambda *a, **k: None
    t = mod("timm")
    t.create_model = lambda *a, **k: None
    try:
        import yacs.config  # noqa: F401
    except Exception:
        y = mod("yacs")
        yc = mod("yacs.config", CfgNode=_CfgNode)
        y.config = yc
    try:
        import tensorboard  # noqa: F401
    except Exception:
        pass


_ready = False


def setup():
    global _ready
    if _ready:
        return
    if not available():
        raise RuntimeError(f"reference not found under {REF_ROOT}")
    _install_stubs()
    for p in (os.path.join(REF_ROOT, "PromptSRC"), os.path.join(REF_ROOT, "Dassl.pytorch")):
        if p not in sys.path:
            sys.path.insert(0, p)
    _ready = True


def make_cfg(trainer: str, **over):
    """Reference defaults (dassl get_cfg_default + train.py:extend_cfg) + overrides 'A.B.C'=value."""
    setup()
    from dassl.config import get_cfg_default
    import importlib
    train = importlib.import_module("train")
    cfg = get_cfg_default()
    train.extend_cfg(cfg)
    cfg.TRAINER.NAME = trainer
    cfg.MODEL.BACKBONE.NAME = "ViT-B/16"
    cfg.TRAINER.PROMPTSRC.PREC = "fp32"
    cfg.TRAINER.COOP.PREC = "fp32"
    cfg.TRAINER.COCOOP.PREC = "fp32"
    cfg.TRAINER.MAPLE.PREC = "fp32"
    # F4: keys the shipped PromptSRC reads but train.py never defines
    cfg.TRAINER.PROMPTSRC.LOGITS_LOSS_WEIGHT = 1.0
    cfg.TRAINER.PROMPTSRC.USE_GPA = False
    lst = []
    for k, v in over.items():
        lst += [k, v]
    cfg.merge_from_list(lst)
    return cfg


@contextlib.contextmanager
def cpu_patches(half_roundtrip: bool):
    """Recipe items 4 and 5, active only inside the context."""
    orig_tcuda, orig_mcuda, orig_half = torch.Tensor.cuda, torch.nn.Module.cuda, torch.Tensor.half
    torch.Tensor.cuda = lambda self, *a, **k: self
    torch.nn.Module.cuda = lambda self, *a, **k: self
    if half_roundtrip:
        torch.Tensor.half = lambda self, *a, **k: self.to(torch.float16).to(self.dtype)
    try:
        yield
    finally:
        torch.Tensor.cuda, torch.nn.Module.cuda, torch.Tensor.half = orig_tcuda, orig_mcuda, orig_half


def _patched_loader(trainer_mod, state_dict, trainer_name):
    """Replacement for trainers.<x>.load_clip_to_cpu (recipe item 3), same design_details logic."""
    from clip import model as clip_model_mod

    def load_clip_to_cpu(cfg, zero_shot_model=False):
        if trainer_name == "PromptSRC":
            if not zero_shot_model:
                dd = {"trainer": "IVLP",
                      "vision_depth": cfg.TRAINER.PROMPTSRC.PROMPT_DEPTH_VISION,
                      "language_depth": cfg.TRAINER.PROMPTSRC.PROMPT_DEPTH_TEXT,
                      "vision_ctx": cfg.TRAINER.PROMPTSRC.N_CTX_VISION,
                      "language_ctx": cfg.TRAINER.PROMPTSRC.N_CTX_TEXT}   # promptsrc.py:30-34
            else:
                dd = {"trainer": "IVLP", "vision_depth": 0, "language_depth": 0,
                      "vision_ctx": 0, "language_ctx": 0}                  # promptsrc.py:38-41
        elif trainer_name == "CoOp":
            dd = {"trainer": "CoOp", "vision_depth": 0, "language_depth": 0, "vision_ctx": 0, "language_ctx": 0}
        elif trainer_name == "CoCoOp":
            dd = {"trainer": "CoCoOp", "vision_depth": 0, "language_depth": 0, "vision_ctx": 0, "language_ctx": 0}
        elif trainer_name == "MaPLe":
            dd = {"trainer": "MaPLe", "vision_depth": 0, "language_depth": 0, "vision_ctx": 0,
                  "language_ctx": 0, "maple_length": cfg.TRAINER.MAPLE.N_CTX}
        else:
            raise ValueError(trainer_name)
        sd = copy.deepcopy(state_dict)
        import io
        with contextlib.redirect_stdout(io.StringIO()):  # build_model prints the missing VPT keys
            m = clip_model_mod.build_model(sd, dd)
        return m

    trainer_mod.load_clip_to_cpu = load_clip_to_cpu


def tokenize(texts, context_length=77):
    setup()
    from clip import clip
    return clip.tokenize(texts, context_length)


def build_reference_model(trainer_name: str, cfg, classnames, state_dict):
    """Reference CustomCLIP (trainers/<x>.py) on CPU with the given CLIP state-dict."""
    setup()
    import importlib
    modname = {"PromptSRC": "trainers.promptsrc", "CoOp": "trainers.coop", "CoCoOp": "trainers.cocoop",
               "MaPLe": "trainers.maple"}[trainer_name]
    tm = importlib.import_module(modname)
    _patched_loader(tm, state_dict, trainer_name)
    import io
    with cpu_patches(half_roundtrip=(trainer_name == "PromptSRC")), contextlib.redirect_stdout(io.StringIO()):
        clip_model = tm.load_clip_to_cpu(cfg)
        clip_model.float()
        model = tm.CustomCLIP(cfg, classnames, clip_model)
    return tm, model


def freeze_like_trainer(trainer_name: str, model):
    """requires_grad pattern of build_model (promptsrc.py:235-247; coop.py:417-419; cocoop/maple alike)."""
    if trainer_name in ("PromptSRC", "MaPLe"):
        for name, p in model.named_parameters():
            if "prompt_learner" not in name:
                p.requires_grad_("VPT" in name)
            elif "ZS_image_encoder" in name:
                p.requires_grad_(False)
    else:
        for name, p in model.named_parameters():
            p.requires_grad_("prompt_learner" in name)


def promptsrc_forward_backward(tm, model, cfg, image, label, lr=0.0025, momentum=0.9, wd=5e-4, steps=1):
    """Drive the reference's own PromptSRC.forward_backward (promptsrc.py:285-360) on a bare trainer."""
    freeze_like_trainer("PromptSRC", model)
    tr = object.__new__(tm.PromptSRC)
    tr.cfg = cfg
    tr.model = model
    tr.device = torch.device("cpu")
    tr.optim = torch.optim.SGD(model.parameters(), lr=lr, momentum=momentum, weight_decay=wd)  # optimizer.py:105-113
    tr.scaler = None
    tr.batch_idx, tr.num_batches = 0, 10 ** 9  # never the last batch: no update_lr / GPA inside the call
    tr.step_counter = 1
    model.train()
    out = []
    with cpu_patches(half_roundtrip=True):
        for _ in range(steps):
            out.append(tr.forward_backward({"img": image, "label": label}))
    return out
