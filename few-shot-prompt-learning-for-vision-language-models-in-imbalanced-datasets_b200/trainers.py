"""Dassl-facing drop-in for the prompted-CLIP trainers.

`CustomCLIP` keeps the reference module signature (`__init__(cfg, classnames, clip_model)`, `forward(image,
label=None)`, attributes `prompt_learner / tokenized_prompts / image_encoder / text_encoder / logit_scale /
dtype`, state-dict keys) and the trainers keep the Dassl `TrainerX` hooks (`check_cfg`, `build_model`,
`forward_backward`, `model_inference`, `parse_batch_train`, `load_model`), registered under the reference's
names so `train.py --trainer PromptSRC|MaPLe|CoOp|CoCoOp` selects them:

    PromptSRC   PromptSRC/trainers/promptsrc.py:170-213 (CustomCLIP), :216-420 (trainer)
    CoOp        PromptSRC/trainers/coop.py:302-390 (CustomCLIP), :396-500 (trainer)
    MaPLe       PromptSRC/trainers/maple.py:111-255 (prompt learner, CustomCLIP), :257-357 (trainer)
    CoCoOp      PromptSRC/trainers/cocoop.py:103-260 (prompt learner, CustomCLIP), :262-372 (trainer)

All arithmetic runs in the sm_100a kernels through `step.PromptedClipStep`; there is no PyTorch fallback.
The trainable prompt tensors are the reference's own `nn.Parameter`s (same names), re-pointed at views of the
step's flat fp32 parameter / gradient buffers, so `state_dict()`, `load_state_dict()`, Dassl checkpoints and
torch LR schedulers work unchanged while the optimizer update is ONE fused kernel (`FlatSGD`).

When Dassl is importable (the reference checkout on `sys.path`), `TrainerX`, `TRAINER_REGISTRY`,
`build_lr_scheduler` are Dassl's own; otherwise (GPU box: the reference does not travel) a minimal stand-in with
the same method names is used so the same trainer code runs in tests and the benchmark.
"""
from __future__ import annotations

import math
import os.path as osp
from collections import OrderedDict

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import synthetic as S
from .step import PromptedClipStep, StepConfig, VisionInference, encode_text_frozen

# ------------------------------------------------------------------ Dassl (real or stand-in)
try:  # pragma: no cover - depends on the reference checkout being importable
    from dassl.engine import TRAINER_REGISTRY, TrainerX
    from dassl.optim import build_lr_scheduler as _dassl_build_sched
    HAVE_DASSL = True
except Exception:  # noqa: BLE001 - any import problem means "no Dassl here"
    HAVE_DASSL = False
    _dassl_build_sched = None

    class _Registry:
        """dassl/utils/registry.py:36-57 -- class name is the key; duplicates raise unless force=True."""

        def __init__(self):
            self._map = {}

        def register(self, obj=None, force=False):
            def deco(cls):
                if cls.__name__ in self._map and not force:
                    raise KeyError(f"{cls.__name__} already registered")
                self._map[cls.__name__] = cls
                return cls
            return deco(obj) if obj is not None else deco

        def get(self, name):
            return self._map[name]

        def registered_names(self):
            return list(self._map)

    TRAINER_REGISTRY = _Registry()

    class TrainerX:
        """The subset of dassl/engine/trainer.py (TrainerBase :74-293, SimpleTrainer :296-486, TrainerX :596-639)
        the prompt trainers touch.  `dm` only needs `.dataset.classnames`."""

        def __init__(self, cfg, dm=None, device=None):
            self._models, self._optims, self._scheds = OrderedDict(), OrderedDict(), OrderedDict()
            self.check_cfg(cfg)
            self.device = torch.device(device if device is not None else "cuda")
            self.start_epoch = self.epoch = 0
            self.max_epoch = cfg.OPTIM.MAX_EPOCH
            self.cfg, self.dm = cfg, dm
            self.batch_idx, self.num_batches = 0, 1
            self.build_model()

        def check_cfg(self, cfg):
            pass

        def register_model(self, name="model", model=None, optim=None, sched=None):
            assert name not in self._models, "Found duplicate model names"
            self._models[name], self._optims[name], self._scheds[name] = model, optim, sched

        def get_model_names(self, names=None):
            return list(self._models) if names is None else list(names)

        def set_model_mode(self, mode="train", names=None):
            for n in self.get_model_names(names):
                self._models[n].train() if mode == "train" else self._models[n].eval()

        def update_lr(self, names=None):
            for n in self.get_model_names(names):
                if self._scheds[n] is not None:
                    self._scheds[n].step()

        def get_current_lr(self, names=None):
            return self._optims[self.get_model_names(names)[0]].param_groups[0]["lr"]

        def model_inference(self, input):
            return self.model(input)

        def parse_batch_train(self, batch):
            return batch["img"].to(self.device), batch["label"].to(self.device)

        def parse_batch_test(self, batch):                                   # dassl/engine/trainer.py:491-498
            return batch["img"].to(self.device), batch["label"].to(self.device)

        def save_model(self, epoch, directory, is_best=False, val_result=None, model_name=""):
            import os
            for n in self.get_model_names():
                os.makedirs(osp.join(directory, n), exist_ok=True)
                torch.save({"state_dict": self._models[n].state_dict(), "epoch": epoch + 1,
                            "optimizer": self._optims[n].state_dict() if self._optims[n] else None,
                            "scheduler": self._scheds[n].state_dict() if self._scheds[n] else None,
                            "val_result": val_result},
                           osp.join(directory, n, model_name or f"model.pth.tar-{epoch + 1}"))


class LookaheadLoader:
    """Wraps Dassl's train loader (dassl/data/data_manager.py:13-52): yields every batch dict with the FOLLOWING batch
    under "_next", so the trainer can start that batch's host-to-device copy behind the launch of the current step.
    Batches, order and length are unchanged; the DataLoader's own workers already run ahead of the consumer."""

    def __init__(self, loader):
        self.loader = loader

    def __len__(self):
        return len(self.loader)

    def __getattr__(self, name):
        return getattr(self.loader, name)

    def __iter__(self):
        it = iter(self.loader)
        cur = next(it, None)
        while cur is not None:
            nxt = next(it, None)
            if isinstance(cur, dict):
                cur = dict(cur)
                cur["_next"] = nxt
            yield cur
            cur = nxt


def _next_of(batch):
    nxt = batch.get("_next") if isinstance(batch, dict) else None
    return (nxt["img"], nxt["label"]) if isinstance(nxt, dict) else None


def _get(node, path, default=None):
    """cfg.A.B.C with a default (the reference reads keys its own extend_cfg never defines)."""
    for k in path.split("."):
        if node is None:
            return default
        node = node.get(k, None) if hasattr(node, "get") else getattr(node, k, None)
    return default if node is None else node


# ------------------------------------------------------------------ module tree from a flat state-dict
class ParamTree(nn.Module):
    """nn.Module tree whose parameter / buffer names are the dotted keys of `tensors` -- a stand-in for the
    reference `clip.model.CLIP` (model.py:488-561) when only its state-dict is available (tests, bench)."""

    def __init__(self, tensors: dict, trainable=()):
        super().__init__()
        for key, t in tensors.items():
            node, parts = self, key.split(".")
            for p in parts[:-1]:
                if p not in node._modules:
                    node.add_module(p, ParamTree({}))
                node = node._modules[p]
            t = t.detach().clone().float()
            node.register_parameter(parts[-1], nn.Parameter(t, requires_grad=key in trainable))

    def __getitem__(self, i):  # resblocks[i]
        return self._modules[str(i)]


def synthetic_clip(arch, seed=0, n_ctx_vision=0, n_ctx_text=0, depth_vision=0, depth_text=1, prompt_seed=100):
    """Random-init CLIP with the reference's parameter names, including the IVLP design's VPT tensors
    (model.py:215-217, :388-390) when depths > 0."""
    arch = S.ARCHS[arch] if isinstance(arch, str) else arch
    sd = S.make_clip_state_dict(arch, seed=seed)
    deep = S.make_deep_prompts(arch, n_ctx_vision, n_ctx_text, depth_vision, depth_text, seed=prompt_seed)
    m = ParamTree({**sd, **deep}, trainable=set(deep))
    m.dtype = torch.float32
    m.visual.input_resolution = arch.image_resolution
    return m


# ------------------------------------------------------------------ optimizer / scheduler
class FlatSGD(torch.optim.Optimizer):
    """torch.optim.SGD semantics (dassl/optim/optimizer.py:105-113: momentum 0.9, wd 5e-4, dampening 0, no
    nesterov) as ONE fused kernel over the step's flat parameter buffer.  It is a torch Optimizer so Dassl's
    LR schedulers, `update_lr` and checkpointing (`state_dict`) drive it unchanged."""

    def __init__(self, step: PromptedClipStep, named_params, lr, momentum=0.9, weight_decay=5e-4, dampening=0.0,
                 nesterov=False):
        self._step = step
        self._named = list(named_params)                       # [(step key, nn.Parameter)]
        super().__init__([p for _, p in self._named], dict(lr=lr, momentum=momentum, weight_decay=weight_decay,
                                                           dampening=dampening, nesterov=nesterov))
        if float(dampening) != 0.0:
            # torch seeds the momentum buffer with the raw first gradient (`buf = g`), which a zero-initialised buffer
            # reproduces only for dampening 0 -- the value of every prompt-learning yaml (optimizer.py:105-113)
            raise NotImplementedError("the fused SGD kernel supports OPTIM.SGD_DAMPNING = 0 only")
        step.init_optimizer(lr=lr, momentum=momentum, weight_decay=weight_decay)
        self._apply_hp(momentum, weight_decay, nesterov)
        self._lr_host = float(lr)
        off = 0
        for k, p in self._named:                               # momentum buffers: views, like torch's state
            n = p.numel()
            self.state[p]["momentum_buffer"] = step.mom_buf[off:off + n].view_as(p)
            off += n

    def _apply_hp(self, momentum, weight_decay, nesterov):
        """Hyper-parameters of the fused kernel (also the ones baked into captured graphs)."""
        st = self._step
        hp, nes = (float(momentum), 0.0, float(weight_decay)), bool(nesterov)
        if hp != st.sgd_hp or nes != st.sgd_nesterov:
            st.sgd_hp, st.sgd_nesterov = hp, nes
            st.drop_graphs()                                   # re-captured on the next fused step
        self._nesterov = nes

    def sync_lr(self):
        lr = float(self.param_groups[0]["lr"])
        if lr != self._lr_host:
            self._step.set_lr(lr)
            self._lr_host = lr

    def sync_grads(self):
        """After an autograd backward, `p.grad` may be a fresh tensor (zero_grad(set_to_none)); bring it into
        the flat gradient buffer the fused kernel reads."""
        for k, p in self._named:
            g = self._step.grads[k]
            if p.grad is None:
                g.zero_()
            elif p.grad.data_ptr() != g.data_ptr():
                g.copy_(p.grad)

    @torch.no_grad()
    def step(self, closure=None):
        self.sync_lr()
        self.sync_grads()
        st = self._step
        m, d, wd = st.sgd_hp
        from . import ops
        ops.sgd_step(st.flat_param, st.flat_grad, st.mom_buf, st.lr_dev, m, d, wd, self._nesterov, False)

    def load_state_dict(self, state_dict):
        bufs = [s.get("momentum_buffer") for s in state_dict.get("state", {}).values()]
        groups = state_dict["param_groups"]
        for g, sg in zip(self.param_groups, groups):
            for k in ("lr", "momentum", "weight_decay", "dampening", "nesterov", "initial_lr"):
                if k in sg:
                    g[k] = sg[k]
        for (k, p), b in zip(self._named, bufs):
            if b is not None:
                self.state[p]["momentum_buffer"].copy_(b)
        g0 = self.param_groups[0]
        if float(g0.get("dampening", 0.0)) != 0.0:
            raise NotImplementedError("the fused SGD kernel supports dampening 0 only")
        self._apply_hp(g0["momentum"], g0["weight_decay"], g0.get("nesterov", False))   # resume with other settings
        self._lr_host = None
        self.sync_lr()


class ConstantWarmupCosine(torch.optim.lr_scheduler.LRScheduler):
    """Stand-in for Dassl's ConstantWarmupScheduler over CosineAnnealingLR (dassl/optim/lr_scheduler.py:35-54,
    :128-141; PromptSRC yaml: WARMUP_EPOCH 1, WARMUP_CONS_LR 1e-5), stepped once per epoch by `update_lr`."""

    def __init__(self, optimizer, max_epoch, warmup_epoch=1, cons_lr=1e-5):
        self.max_epoch, self.warmup_epoch, self.cons_lr = max_epoch, warmup_epoch, cons_lr
        super().__init__(optimizer)

    def get_lr(self):
        e = self.last_epoch
        if e < self.warmup_epoch:
            return [self.cons_lr for _ in self.base_lrs]
        # WARMUP_RECOUNT (default True, lr_scheduler.py:131-132): the cosine successor starts counting when the
        # warm-up ends, so epoch `warmup_epoch` runs at the base lr
        e -= self.warmup_epoch
        return [0.5 * b * (1 + math.cos(math.pi * e / self.max_epoch)) for b in self.base_lrs]


def build_scheduler(optim, optim_cfg):
    if _dassl_build_sched is not None:
        try:
            return _dassl_build_sched(optim, optim_cfg)
        except TypeError:
            # Dassl's warm-up schedulers pass `verbose` positionally (lr_scheduler.py:22), which torch >= 2.7
            # rejects: fall through to the restatement below (same lr sequence, verified on torch 1.9 semantics)
            pass
    warm = _get(optim_cfg, "WARMUP_EPOCH", 0) or 0
    if _get(optim_cfg, "LR_SCHEDULER", "cosine") != "cosine":
        raise NotImplementedError("without Dassl only the cosine scheduler of the prompt-learning yamls is provided")
    return ConstantWarmupCosine(optim, optim_cfg.MAX_EPOCH, warm, _get(optim_cfg, "WARMUP_CONS_LR", 1e-5))


# ------------------------------------------------------------------ tokenisation (reference tokenizer, host side)
def _reference_tokenize(texts):
    """`clip.tokenize` of the reference (clip/clip.py:185-221) -- integer work stays the reference's own code
    (SURVEY a-18).  Needs PromptSRC/ on sys.path; tests and the benchmark pass `tokenized_prompts` instead."""
    try:
        from clip import clip as _clip
    except Exception as e:  # noqa: BLE001
        raise RuntimeError("CustomCLIP needs the reference tokenizer (PromptSRC/clip on sys.path) or explicit "
                           "`tokenized_prompts=`") from e
    return torch.cat([_clip.tokenize(t) for t in texts])


class _TextEncoder(nn.Module):
    """TextEncoder(clip_model) (promptsrc.py:49-57): same attribute / state-dict names, arithmetic in the kernels."""

    def __init__(self, clip_model):
        super().__init__()
        self.transformer = clip_model.transformer
        self.positional_embedding = clip_model.positional_embedding
        self.ln_final = clip_model.ln_final
        self.text_projection = clip_model.text_projection
        self.dtype = clip_model.dtype


class _PromptLearner(nn.Module):
    """VLPromptLearner (promptsrc.py:72-168) / PromptLearner (coop.py:208-296): `ctx` + the frozen token
    embeddings either side of it.  `forward()` returns the assembled prompts like the reference (a plain
    torch.cat -- only used by external callers; the training step assembles them inside a kernel)."""

    def __init__(self, ctx, token_prefix, token_suffix, tokenized_prompts, name_lens=None):
        super().__init__()
        self.ctx = nn.Parameter(ctx)
        self.register_buffer("token_prefix", token_prefix)
        self.register_buffer("token_suffix", token_suffix)
        self.tokenized_prompts = tokenized_prompts
        self.n_cls, self.n_ctx = token_prefix.shape[0], ctx.shape[-2]
        self.name_lens = name_lens

    def forward(self):
        ctx = self.ctx
        if ctx.dim() == 2:
            ctx = ctx.unsqueeze(0).expand(self.n_cls, -1, -1)
        return torch.cat([self.token_prefix, ctx, self.token_suffix], dim=1)


class _Bridge(torch.autograd.Function):
    """Autograd view of the kernel step: (prompt parameters) -> (logits, t_hat, f_hat).  backward() runs the
    prompt-only backward kernels from the upstream gradients, whatever loss the caller composed."""

    @staticmethod
    def forward(ctx, model, image, *params):
        st = model._step
        st.image.copy_(image)
        st.forward_train()
        ctx.model = model
        ctx.set_materialize_grads(False)      # unused outputs arrive as None, not as zero tensors
        return st.logits.clone(), st.t_hat.clone(), st.f_hat.clone()

    @staticmethod
    def backward(ctx, g_logits, g_that, g_fhat):
        st = ctx.model._step
        st.backward_from(g_logits, g_that, g_fhat)
        return (None, None) + tuple(st.grads[k].clone() for k, _ in ctx.model._named_prompts)


class _CustomCLIPBase(nn.Module):
    """Shared machinery of the CustomCLIP variants: builds the kernel step from a reference-keyed CLIP module,
    binds the prompt nn.Parameters to the flat buffers, provides the fused / bridged / eval paths."""

    def _make_prompt_learner(self, scfg, ctx_init, prefix, suffix, tokens):
        return _PromptLearner(ctx_init.float(), prefix, suffix, tokens)

    def _named_for_step(self, scfg):
        """[(step key, reference-named nn.Parameter)] in the order of the step's flat buffer."""
        named = [("ctx", self.prompt_learner.ctx)]
        for i in range(1, scfg.depth_text):
            named.append((f"text.{i}", self.text_encoder.transformer.resblocks[i].VPT_shallow))
        if scfg.n_ctx_vision > 0:
            named.append(("vpt", self.image_encoder.VPT))
            for i in range(1, scfg.depth_vision):
                named.append((f"vision.{i}", self.image_encoder.transformer.resblocks[i].VPT_shallow))
        return named

    def _build(self, cfg, clip_model, tokens, step_kwargs, ctx_init, fixed_embeddings=None, device=None):
        dev = torch.device(device if device is not None else "cuda")
        from . import ops as _ops
        if dev.type != "cuda" and not getattr(_ops, "_FAKE_KERNELS", False):   # fakes: CPU host-logic tests only
            raise RuntimeError("psrc_b200.CustomCLIP runs on a CUDA device only (there is no CPU path)")
        sd = {k: v.detach().float().cpu() for k, v in clip_model.state_dict().items()}
        arch = S.arch_from_state_dict(sd)
        prec_cfg = step_kwargs.pop("prec_cfg", "fp16")
        prec = "fp32" if prec_cfg == "fp32" else "bf16"      # "fp16"/"amp" -> 16-bit tensor-core mode
        world, rank, pg = 1, 0, None
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            world, rank, pg = dist.get_world_size(), dist.get_rank(), dist.group.WORLD
        # Dassl drops the last short batch only when the train set holds at least one full batch
        # (data_manager.py:37-48): a few-shot set SMALLER than BATCH_SIZE arrives as one short batch every iteration,
        # and the reference trains on it -- the trainer passes that size in (`train_batch`, _PromptTrainer._train_batch_size)
        batch_size = int(step_kwargs.pop("train_batch", None) or cfg.DATALOADER.TRAIN_X.BATCH_SIZE)
        if batch_size % world != 0:
            raise RuntimeError(f"the training batch of {batch_size} images does not split over {world} ranks: every rank "
                               "takes an equal contiguous share of the batch Dassl's loader delivers "
                               "(choose BATCH_SIZE / the shot count as a multiple of the world size)")
        scfg = StepConfig(arch=arch, n_cls=tokens.shape[0], batch=batch_size // world,
                          prec=prec, world_size=world, rank=rank, process_group=pg, **step_kwargs)
        self._step = PromptedClipStep(scfg, sd, tokens, dev, fixed_embeddings=fixed_embeddings)
        self._infer = None
        self._test_batch = int(_get(cfg, "DATALOADER.TEST.BATCH_SIZE", 100))
        self._use_graph = bool(_get(cfg, "TRAINER.B200.CUDA_GRAPH", True)) and dev.type == "cuda"
        self._eager_steps = 0
        emb = sd["token_embedding.weight"][tokens]
        n_ctx = scfg.n_ctx_text
        self.prompt_learner = self._make_prompt_learner(scfg, ctx_init, emb[:, :1].clone(), emb[:, 1 + n_ctx:].clone(),
                                                        tokens)
        self.tokenized_prompts = tokens
        self.image_encoder = clip_model.visual
        self.text_encoder = _TextEncoder(clip_model)
        self.logit_scale = clip_model.logit_scale
        self.dtype = clip_model.dtype
        self.n_cls = tokens.shape[0]
        self.to(dev)
        # ---- bind the reference-named prompt parameters to the flat buffers
        st = self._step
        named = self._named_for_step(scfg)
        assert [k for k, _ in named] == list(st.params), (list(st.params), [k for k, _ in named])
        with torch.no_grad():
            for k, p in named:
                st.params[k].copy_(p.detach().to(dev).reshape(st.params[k].shape))
                p.data = st.params[k]
                p.grad = st.grads[k]
                p.requires_grad_(True)
        self._named_prompts = named
        # everything else is frozen, as the reference trainers' build_model leaves it (promptsrc.py:238-247,
        # coop.py:419-421): a CLIP module that arrives with requires_grad=True everywhere (load_clip_to_cpu) must not
        # show up as trainable in named_parameters() / checkpoints' optimizer state
        prompt_ids = {id(p) for _, p in named}
        for p in self.parameters():
            if id(p) not in prompt_ids:
                p.requires_grad_(False)
        return st

    # -- reference behaviour: everything except the prompts is frozen (promptsrc.py:238-247, coop.py:419-421)
    def prompt_parameters(self):
        return [p for _, p in self._named_prompts]

    def train(self, mode: bool = True):
        if getattr(self, "_infer", None) is not None:
            self._infer.text_valid = False       # prompts may change between evaluations
        return super().train(mode)

    def _is_inference(self, label):
        """Dassl's `set_model_mode("eval")` reaches only the REGISTERED module -- for CoOp / CoCoOp that is
        `prompt_learner` (coop.py:431, cocoop.py:304), so `self.training` stays True during `test()`.  The reference
        then returns logits because `label is None`; `model_inference` runs under `torch.no_grad()`
        (dassl/engine/trainer.py:445-449), which is what selects the forward-only path here."""
        return (not self.training) or (label is None and not torch.is_grad_enabled())

    def _bridged(self, image):
        return _Bridge.apply(self, image, *self.prompt_parameters())

    def _eval_logits(self, image):
        if self._step.per_image:                 # CoCoOp: text features depend on the image, nothing to cache
            return self._step.per_image_logits_for(image.float())
        if self._infer is None:
            self._infer = VisionInference(self._step, self._test_batch)
        # The cached text features follow the prompts.  Dassl reaches CoOp's / CoCoOp's prompts through the REGISTERED
        # sub-module only (`prompt_learner.eval()`, `prompt_learner.load_state_dict()` in load_model -- e.g. the
        # best-val checkpoint before the final test), which never passes through this module's train() /
        # load_state_dict() overrides; every such torch-level write bumps the tensors' version counters.
        ver = sum(p._version for _, p in self._named_prompts)
        if ver != getattr(self, "_eval_param_version", None):
            self._infer.text_valid = False
            self._eval_param_version = ver
        return self._infer.logits_for(image.float())

    def make_optimizer(self, optim_cfg):
        return FlatSGD(self._step, self._named_prompts, lr=float(optim_cfg.LR),
                       momentum=float(_get(optim_cfg, "MOMENTUM", 0.9)),
                       weight_decay=float(_get(optim_cfg, "WEIGHT_DECAY", 5e-4)),
                       dampening=float(_get(optim_cfg, "SGD_DAMPNING", 0.0)),
                       nesterov=bool(_get(optim_cfg, "SGD_NESTEROV", False)))

    def train_step_fused(self, image, label, optim: FlatSGD, next_batch=None):
        """The trainer's fast path: H2D of the batch, ONE CUDA-graph replay of forward + fused losses +
        prompt-only backward + SGD, and the loss terms left on the device (`.item()` is the caller's sync)."""
        st = self._step
        w = st.cfg.world_size
        if isinstance(image, dict):          # TRAINER.B200.DEVICE_INPUT: uint8 windows + crop geometry
            return self._train_step_fused_u8(image, label, optim, next_batch)
        if w > 1 and image.shape[0] == st.cfg.batch * w:
            # Dassl's loader has no DistributedSampler: every rank is handed the full BATCH_SIZE (same seed, same
            # order) -- each rank trains on its contiguous share, as the sharded step expects
            if not getattr(self, "_ranks_share_batch", False):
                self.check_ranks_share_the_batch(label)          # first step only: one 16-byte all-gather
                self._ranks_share_batch = True
            lo = st.cfg.rank * st.cfg.batch
            image, label = image[lo:lo + st.cfg.batch], label[lo:lo + st.cfg.batch]
            next_batch = None
        if image.shape[0] != st.cfg.batch:
            raise RuntimeError(f"training batch {image.shape[0]} != configured per-rank batch {st.cfg.batch} "
                               "(Dassl's train loader drops the last short batch, data_manager.py:42)")
        self._before_fused_step(optim)
        if image.is_cuda or st.device.type != "cuda":     # (CPU: host-logic tests over restated kernels only)
            st.image.copy_(image, non_blocking=True)
            st.labels.copy_(label.to(torch.int32), non_blocking=True)
            st.run()
        elif getattr(st, "_staged_key", None) == id(image):
            st.run_staged()                  # this batch was copied over while the previous step was running
        else:
            st.run_host(image, label)        # H2D on the copy stream, under the text-tower forward
        if next_batch is not None and not next_batch[0].is_cuda and next_batch[0].shape[0] == st.cfg.batch \
                and st.device.type == "cuda":
            st.stage_host(*next_batch)       # the following batch starts its H2D under this step's compute
        if self._infer is not None:
            self._infer.text_valid = False
        return st.losses

    def _before_fused_step(self, optim):
        st = self._step
        optim.sync_lr()
        if self._use_graph and getattr(st, "_graph", None) is None and getattr(st, "_seg_graphs", None) is None \
                and self._eager_steps >= 1:
            st.capture(warmup=1)
        self._eager_steps += 1
        # any torch-level in-place write to a trainable tensor (load_state_dict on a sub-module, a foreign optimizer,
        # re-initialisation) bumps its version counter, the fused SGD kernel does not -- a text prefix enqueued
        # ahead with the old values must then be recomputed
        ver = sum(p._version for _, p in self._named_prompts)
        if ver != getattr(self, "_param_version", None):
            st.invalidate_text()
            self._param_version = ver

    def _train_step_fused_u8(self, item, label, optim, next_batch=None):
        """`train_step_fused` for batches collated from `input_pipeline.DeviceTransform` items ({"u8", "geom_f",
        "geom_i"}): the uint8 windows cross PCIe, `psrc_crop_resize_normalize` writes the step's fp32 image buffer, then
        the same graph replay.  The following batch's copies are started under this step (LookaheadLoader)."""
        st = self._step
        w, b = st.cfg.world_size, st.cfg.batch
        if w > 1 and item["u8"].shape[0] == b * w:
            if not getattr(self, "_ranks_share_batch", False):
                self.check_ranks_share_the_batch(label)
                self._ranks_share_batch = True
            lo = st.cfg.rank * b
            item = {k: v[lo:lo + b].contiguous() for k, v in item.items()}
            label, next_batch = label[lo:lo + b], None
        if item["u8"].shape[0] != b:
            raise RuntimeError(f"training batch {item['u8'].shape[0]} != configured per-rank batch {b} "
                               "(Dassl's train loader drops the last short batch, data_manager.py:42)")
        self._before_fused_step(optim)
        st.input_kwargs = getattr(self, "input_kwargs", {})
        ok_next = next_batch is not None and isinstance(next_batch[0], dict) and next_batch[0]["u8"].shape[0] == b
        st.run_host_u8(item, label, next_host=next_batch if ok_next else None)
        if self._infer is not None:
            self._infer.text_valid = False
        return st.losses

    def images_from_u8(self, item):
        """{"u8", "geom_f", "geom_i"} (any batch size: evaluation, the autograd-bridged forward) -> normalised fp32
        [B, 3, R, R] on the device, one launch of `psrc_crop_resize_normalize`."""
        from . import ops as _ops
        from .input_pipeline import CLIP_MEAN, CLIP_STD, INTERPOLATION, max_taps
        st, kw = self._step, getattr(self, "input_kwargs", {})
        interp = INTERPOLATION[kw.get("interpolation", "bicubic")]
        dev, R = st.device, st.cfg.arch.image_resolution
        u8, gf, gi = (item[k].to(dev, non_blocking=True).contiguous() for k in ("u8", "geom_f", "geom_i"))
        out = torch.empty(u8.shape[0], 3, R, R, device=dev)
        _ops.crop_resize_normalize(u8, gf, gi, out, interp, int(kw.get("quantize", 2)),
                                   max_taps(item["geom_f"].cpu().numpy(), interp), kw.get("mean", CLIP_MEAN),
                                   kw.get("std", CLIP_STD))
        return out

    def check_ranks_share_the_batch(self, label):
        """Slicing the full batch by rank is only a partition of it when every rank's Dassl loader drew the SAME batch.
        Dassl seeds its samplers only for cfg.SEED >= 0 (dassl/utils/tools.py:set_random_seed, called from train.py
        when a seed is given; the default is -1): unseeded ranks shuffle differently, and an epoch then visits some
        samples twice and others never.  Compare a fingerprint of the labels across ranks and refuse to go on."""
        import torch.distributed as dist
        st = self._step
        lab = label.detach().to("cpu", torch.int64).flatten()
        fp = torch.stack([lab.sum(), (lab * torch.arange(1, lab.numel() + 1)).sum()]).to(st.device)
        got = [torch.empty_like(fp) for _ in range(st.cfg.world_size)]
        dist.all_gather(got, fp, group=st.cfg.process_group)
        if any(not torch.equal(g, got[0]) for g in got):
            raise RuntimeError("the ranks received different training batches, so slicing the batch by rank would not "
                               "partition it: give every rank the same seed >= 0 (train.py --seed / cfg.SEED; Dassl "
                               "leaves its samplers unseeded for the default -1)")

    def step_loss(self, losses) -> float:
        """`loss.item()` of the fused step (promptsrc.py:330): read from the pinned copy made right after the step, so the
        host does not wait for the next step's text-tower prefix that is already enqueued behind it."""
        return self._step.loss_value()

    def params_changed(self):
        """Call after writing the prompt buffer from outside the fused step (checkpoint load, GPA, a foreign
        optimizer): drops the text prefix enqueued ahead and the evaluation path's cached text features."""
        if getattr(self, "_step", None) is not None:
            self._step.invalidate_text()
        if getattr(self, "_infer", None) is not None:
            self._infer.text_valid = False

    def load_state_dict(self, *args, **kwargs):
        out = super().load_state_dict(*args, **kwargs)
        self.params_changed()                 # prompt parameters changed behind the step's back
        return out


def _focal_alpha_from_cfg(cfg, n_cls):
    """alpha_c = sum(n) / (C * n_c) from DATASET.PER_CLASS_SHOTS (coop.py:326-346, cocoop.py:219-230)."""
    per_class = _get(cfg, "DATASET.PER_CLASS_SHOTS", None)
    if per_class is None or len(per_class) == 0:
        return [1.0] * n_cls
    if isinstance(per_class, str):
        per_class = list(map(int, per_class.strip("[]").split(",")))
    return S.focal_alpha(list(per_class), n_cls)


class PromptSRCCustomCLIP(_CustomCLIPBase):
    """CustomCLIP of trainers/promptsrc.py:170-213.  Extra keyword arguments (all optional) exist because the
    reference obtains them from the network / its tokenizer: `tokenized_prompts` int64 [C, 77],
    `teacher_tokens` (the "a photo of a {}." prompts, promptsrc.py:124-129; or `[T, C, 77]` for the template-ensemble
    teacher of promptsrc_pets_focal.py:165-175), `fixed_embeddings`."""

    def __init__(self, cfg, classnames, clip_model, tokenized_prompts=None, teacher_tokens=None,
                 fixed_embeddings=None, device=None, train_batch=None):
        super().__init__()
        T = cfg.TRAINER.PROMPTSRC
        assert T.PROMPT_DEPTH_TEXT >= 1, "Language prompt depth should be >= 1 (promptsrc.py:77-80)"
        n_ctx, ctx_init = int(T.N_CTX_TEXT), T.CTX_INIT
        names = [n.replace("_", " ") for n in classnames]
        sd_emb = clip_model.state_dict()["token_embedding.weight"].detach().float().cpu()
        if ctx_init and n_ctx <= 4:                                          # promptsrc.py:90-98
            ctx_init = ctx_init.replace("_", " ")
            prefix_words = ctx_init
        else:                                                                # :99-103
            prefix_words = " ".join(["X"] * n_ctx)
        if tokenized_prompts is None:
            tokenized_prompts = _reference_tokenize([prefix_words + " " + n + "." for n in names])   # :109-111
        tokens = torch.as_tensor(tokenized_prompts).long().cpu()
        if ctx_init and n_ctx <= 4:
            ctx_vectors = sd_emb[tokens[0, 1:1 + n_ctx]].clone()             # embedding of the init words
        else:
            ctx_vectors = torch.empty(n_ctx, sd_emb.shape[1])
            nn.init.normal_(ctx_vectors, std=0.02)
        dev = torch.device(device if device is not None else "cuda")
        if fixed_embeddings is None:                                         # :116-132 frozen-CLIP teacher
            if teacher_tokens is None:
                same = (ctx_init or "").strip().lower() == "a photo of a"
                teacher_tokens = tokens if same else _reference_tokenize([f"a photo of a {n}." for n in names])
            sd = {k: v.detach().float().cpu() for k, v in clip_model.state_dict().items()}
            tt_ = torch.as_tensor(teacher_tokens).long()
            if tt_.dim() == 3:
                # template ensemble of the teacher (promptsrc_pets_focal.py:165-175: one encode_text pass per template,
                # mean of the un-normalised features over templates); teacher_tokens is [T, C, 77]
                fixed_embeddings = torch.stack([encode_text_frozen(sd, t_, dev, prec="fp32") for t_ in tt_]).mean(0)
            else:
                fixed_embeddings = encode_text_frozen(sd, tt_, dev, prec="fp32")
        depth_v = int(T.PROMPT_DEPTH_VISION)
        kw = dict(prec_cfg=T.PREC, n_ctx_text=n_ctx, n_ctx_vision=int(T.N_CTX_VISION) if depth_v >= 1 else 0,
                  depth_text=int(T.PROMPT_DEPTH_TEXT), depth_vision=depth_v, zero_shot_branch=True,
                  w_text=float(T.TEXT_LOSS_WEIGHT), w_image=float(T.IMAGE_LOSS_WEIGHT),
                  w_logits=float(_get(T, "LOGITS_LOSS_WEIGHT", 1.0)))      # key missing upstream
        if _get(T, "LOSS_TYPE", "ce") == "focal":                            # promptsrc_pets_focal.py:230-247
            kw["focal_alpha"] = _focal_alpha_from_cfg(cfg, len(classnames))
        kw["train_batch"] = train_batch
        self._build(cfg, clip_model, tokens, kw, ctx_vectors, fixed_embeddings=fixed_embeddings, device=dev)
        self.prompt_learner.fixed_embeddings = fixed_embeddings.to(dev)
        self.total_epochs = cfg.OPTIM.MAX_EPOCH
        self.cfg = cfg

    def forward(self, image, label=None):
        if isinstance(image, dict):
            image = self.images_from_u8(image)
        if not self.prompt_learner.training:
            return self._eval_logits(image)                                  # promptsrc.py:212-213
        st = self._step
        logits, t_hat, f_hat = self._bridged(image.float())
        loss_ce = F.cross_entropy(logits, label) if st.focal_alpha is None else _focal(logits, label, st)
        return (loss_ce, t_hat, st.fixed_hat.clone(), st.zf_hat.clone(), f_hat, st.zs_logits.clone(), logits)


def _focal(logits, label, st):
    ce = F.cross_entropy(logits, label, reduction="none")                   # coop.py:146-163
    pt = torch.exp(-ce)
    return (st.focal_alpha[label] * (1 - pt) ** st.cfg.focal_gamma * ce).mean()


class CoOpCustomCLIP(_CustomCLIPBase):
    """CustomCLIP of trainers/coop.py:302-390 for LOSS_TYPE ce / focal, CLASS_TOKEN_POSITION end / middle / front."""

    def __init__(self, cfg, classnames, clip_model, tokenized_prompts=None, device=None, train_batch=None):
        super().__init__()
        T = cfg.TRAINER.COOP
        self.cfg = cfg
        self.loss_type = _get(T, "LOSS_TYPE", "ce")
        if self.loss_type not in ("ce", "focal"):
            raise NotImplementedError("LOSS_TYPE simclr (coop.py:316-321) is outside the accelerated path")
        position = _get(T, "CLASS_TOKEN_POSITION", "end")
        if position not in ("end", "middle", "front"):
            raise ValueError("Unknown class_token_position")                 # coop.py:293
        n_ctx, ctx_init, csc = int(T.N_CTX), _get(T, "CTX_INIT", ""), bool(_get(T, "CSC", False))
        names = [n.replace("_", " ") for n in classnames]
        sd_emb = clip_model.state_dict()["token_embedding.weight"].detach().float().cpu()
        if ctx_init:                                                         # coop.py:218-226
            ctx_init = ctx_init.replace("_", " ")
            n_ctx = len(ctx_init.split(" "))
            prefix_words = ctx_init
        else:
            prefix_words = " ".join(["X"] * n_ctx)
        if tokenized_prompts is None:
            tokenized_prompts = _reference_tokenize([prefix_words + " " + n + "." for n in names])
        tokens = torch.as_tensor(tokenized_prompts).long().cpu()
        if ctx_init:
            ctx_vectors = sd_emb[tokens[0, 1:1 + n_ctx]].clone()
        else:                                                                # :228-236
            shape = (len(names), n_ctx, sd_emb.shape[1]) if csc else (n_ctx, sd_emb.shape[1])
            ctx_vectors = torch.empty(*shape)
            nn.init.normal_(ctx_vectors, std=0.02)
        kw = dict(prec_cfg=T.PREC, n_ctx_text=n_ctx, n_ctx_vision=0, depth_text=1, depth_vision=0,
                  ctx_per_class=csc, zero_shot_branch=False, class_token_position=position)
        if self.loss_type == "focal":
            kw["focal_alpha"] = _focal_alpha_from_cfg(cfg, len(classnames))
        kw["train_batch"] = train_batch
        self._build(cfg, clip_model, tokens, kw, ctx_vectors, device=device)

    def forward_once(self, image):
        return self._bridged(image.float())[0]

    def forward(self, img1, lbl=None, img2=None):
        if isinstance(img1, dict):
            img1 = self.images_from_u8(img1)
        if self._is_inference(lbl):
            return self._eval_logits(img1)
        logits = self.forward_once(img1)
        if lbl is None:
            return logits
        st = self._step
        return F.cross_entropy(logits, lbl) if st.focal_alpha is None else _focal(logits, lbl, st)



class _MaPLePromptLearner(nn.Module):
    """MultiModalPromptLearner (maple.py:111-204): `ctx`, the coupling `proj`, the deeper text prompts and one
    coupling Linear per deeper layer -- same attribute / state-dict names.  The reference hard-codes the ViT-B widths
    (768, 512); here they follow the CLIP model (so ViT-L/14 works).  `forward()` returns the reference's
    4-tuple with plain torch ops (external callers only; the training step runs the batched kernel)."""

    def __init__(self, ctx, vis_width, depth, token_prefix, token_suffix, tokenized_prompts):
        super().__init__()
        import copy
        n_ctx, ctx_dim = ctx.shape
        self.proj = nn.Linear(ctx_dim, vis_width)                               # maple.py:144 (kept fp32)
        self.ctx = nn.Parameter(ctx)
        self.compound_prompts_text = nn.ParameterList(
            [nn.Parameter(torch.empty(n_ctx, ctx_dim)) for _ in range(depth - 1)])   # :149-151
        for p_ in self.compound_prompts_text:
            nn.init.normal_(p_, std=0.02)
        single = nn.Linear(ctx_dim, vis_width)
        self.compound_prompt_projections = nn.ModuleList([copy.deepcopy(single) for _ in range(depth - 1)])  # :155-156
        self.register_buffer("token_prefix", token_prefix)
        self.register_buffer("token_suffix", token_suffix)
        self.tokenized_prompts = tokenized_prompts
        self.n_cls, self.n_ctx, self.compound_prompts_depth = token_prefix.shape[0], n_ctx, depth

    def forward(self):
        ctx = self.ctx.unsqueeze(0).expand(self.n_cls, -1, -1)
        prompts = torch.cat([self.token_prefix, ctx, self.token_suffix], dim=1)
        visual = [layer(self.compound_prompts_text[i]) for i, layer in enumerate(self.compound_prompt_projections)]
        return prompts, self.proj(self.ctx), self.compound_prompts_text, visual


class MaPLeCustomCLIP(_CustomCLIPBase):
    """CustomCLIP of trainers/maple.py:206-255 (CE or focal loss, :221-238)."""

    def __init__(self, cfg, classnames, clip_model, tokenized_prompts=None, device=None, train_batch=None):
        super().__init__()
        T = cfg.TRAINER.MAPLE
        self.cfg = cfg
        depth = int(T.PROMPT_DEPTH)
        assert depth >= 1, "PROMPT_DEPTH must be >= 1"                          # maple.py:122
        n_ctx, ctx_init = int(T.N_CTX), _get(T, "CTX_INIT", "")
        names = [n.replace("_", " ") for n in classnames]
        sd_emb = clip_model.state_dict()["token_embedding.weight"].detach().float().cpu()
        if ctx_init and n_ctx <= 4:                                             # :127-133
            prefix_words = ctx_init.replace("_", " ")
        else:
            prefix_words = " ".join(["X"] * n_ctx)
        if tokenized_prompts is None:
            tokenized_prompts = _reference_tokenize([prefix_words + " " + n + "." for n in names])
        tokens = torch.as_tensor(tokenized_prompts).long().cpu()
        if ctx_init and n_ctx <= 4:
            ctx_vectors = sd_emb[tokens[0, 1:1 + n_ctx]].clone()
        else:
            ctx_vectors = torch.empty(n_ctx, sd_emb.shape[1])
            nn.init.normal_(ctx_vectors, std=0.02)
        self.use_focal_loss = bool(_get(T, "USE_FOCAL_LOSS", False))
        kw = dict(prec_cfg=T.PREC, n_ctx_text=n_ctx, n_ctx_vision=n_ctx, depth_text=depth, depth_vision=depth,
                  coupling=True, zero_shot_branch=False)
        if self.use_focal_loss:
            kw["focal_alpha"] = _focal_alpha_from_cfg(cfg, len(classnames))
        kw["train_batch"] = train_batch
        self._build(cfg, clip_model, tokens, kw, ctx_vectors, device=device)

    def _make_prompt_learner(self, scfg, ctx_init, prefix, suffix, tokens):
        return _MaPLePromptLearner(ctx_init.float(), scfg.arch.vision_width, scfg.depth_text, prefix, suffix, tokens)

    def _named_for_step(self, scfg):
        pl, d = self.prompt_learner, scfg.depth_text
        named = [("ctx", pl.ctx)] + [(f"text.{i}", pl.compound_prompts_text[i - 1]) for i in range(1, d)]
        lin = [pl.proj] + list(pl.compound_prompt_projections)
        named += [(f"couple.w.{i}", lin[i].weight) for i in range(d)]
        named += [(f"couple.b.{i}", lin[i].bias) for i in range(d)]
        return named

    def forward(self, image, label=None):
        if isinstance(image, dict):
            image = self.images_from_u8(image)
        if self._is_inference(label):
            return self._eval_logits(image)
        logits = self._bridged(image.float())[0]
        if label is None:
            return logits                                                       # maple.py:252-255
        st = self._step
        return F.cross_entropy(logits, label) if st.focal_alpha is None else _focal(logits, label, st)


class _CoCoOpPromptLearner(nn.Module):
    """PromptLearner of trainers/cocoop.py:103-198: `ctx` + `meta_net` (linear1 -> ReLU -> linear2), same names.
    `forward(im_features)` returns the [B, C, 77, D] instance-conditioned prompts with torch ops (external callers
    only -- the training step never materialises them: the context shift is applied inside the assembly kernel)."""

    def __init__(self, ctx, vis_dim, token_prefix, token_suffix, tokenized_prompts):
        super().__init__()
        ctx_dim = ctx.shape[1]
        self.ctx = nn.Parameter(ctx)
        self.meta_net = nn.Sequential(OrderedDict([("linear1", nn.Linear(vis_dim, vis_dim // 16)),
                                                   ("relu", nn.ReLU(inplace=True)),
                                                   ("linear2", nn.Linear(vis_dim // 16, ctx_dim))]))   # :139-143
        self.register_buffer("token_prefix", token_prefix)
        self.register_buffer("token_suffix", token_suffix)
        self.tokenized_prompts = tokenized_prompts
        self.n_cls, self.n_ctx = token_prefix.shape[0], ctx.shape[0]

    def forward(self, im_features):
        ctx_shifted = self.ctx.unsqueeze(0) + self.meta_net(im_features).unsqueeze(1)       # :185-188
        return torch.stack([torch.cat([self.token_prefix, c.unsqueeze(0).expand(self.n_cls, -1, -1),
                                       self.token_suffix], dim=1) for c in ctx_shifted], dim=0)


class CoCoOpCustomCLIP(_CustomCLIPBase):
    """CustomCLIP of trainers/cocoop.py:200-260 (CE or focal loss, :214-233).  The meta-net stays fp32 whatever
    PREC says (the reference's `.half()` at :146-147 cannot run on its fp32 model)."""

    def __init__(self, cfg, classnames, clip_model, tokenized_prompts=None, device=None, train_batch=None):
        super().__init__()
        T = cfg.TRAINER.COCOOP
        self.cfg = cfg
        n_ctx, ctx_init = int(T.N_CTX), _get(T, "CTX_INIT", "")
        names = [n.replace("_", " ") for n in classnames]
        sd_emb = clip_model.state_dict()["token_embedding.weight"].detach().float().cpu()
        if ctx_init:                                                            # cocoop.py:122-131
            ctx_init = ctx_init.replace("_", " ")
            n_ctx = len(ctx_init.split(" "))
            prefix_words = ctx_init
        else:
            prefix_words = " ".join(["X"] * n_ctx)
        if tokenized_prompts is None:
            tokenized_prompts = _reference_tokenize([prefix_words + " " + n + "." for n in names])
        tokens = torch.as_tensor(tokenized_prompts).long().cpu()
        if ctx_init:
            ctx_vectors = sd_emb[tokens[0, 1:1 + n_ctx]].clone()
        else:
            ctx_vectors = torch.empty(n_ctx, sd_emb.shape[1])
            nn.init.normal_(ctx_vectors, std=0.02)
        self.use_focal_loss = bool(_get(T, "USE_FOCAL_LOSS", False))
        vis_dim = clip_model.state_dict()["text_projection"].shape[1]           # = visual.output_dim (:115)
        kw = dict(prec_cfg=T.PREC, n_ctx_text=n_ctx, n_ctx_vision=0, depth_text=1, depth_vision=0,
                  per_image_text=True, meta_hidden=vis_dim // 16, zero_shot_branch=False)
        if self.use_focal_loss:
            kw["focal_alpha"] = _focal_alpha_from_cfg(cfg, len(classnames))
        kw["train_batch"] = train_batch
        self._build(cfg, clip_model, tokens, kw, ctx_vectors, device=device)

    def _make_prompt_learner(self, scfg, ctx_init, prefix, suffix, tokens):
        return _CoCoOpPromptLearner(ctx_init.float(), scfg.arch.embed_dim, prefix, suffix, tokens)

    def _named_for_step(self, scfg):
        m = self.prompt_learner.meta_net
        return [("ctx", self.prompt_learner.ctx), ("meta.w1", m.linear1.weight), ("meta.b1", m.linear1.bias),
                ("meta.w2", m.linear2.weight), ("meta.b2", m.linear2.bias)]

    def forward(self, image, label=None):
        if isinstance(image, dict):
            image = self.images_from_u8(image)
        if self._is_inference(label):
            return self._eval_logits(image)
        logits = self._bridged(image.float())[0]
        if label is None:
            return logits                                                       # cocoop.py:257-260
        st = self._step
        return F.cross_entropy(logits, label) if st.focal_alpha is None else _focal(logits, label, st)


# ------------------------------------------------------------------ trainers
def _load_clip(cfg, trainer_name):
    """`load_clip_to_cpu(cfg)` of the reference trainer module (promptsrc.py:19-46, coop.py:169-190): used when the
    trainer runs inside the reference checkout.  Tests / bench set `CLIP_MODEL_FACTORY` instead (no network)."""
    if CLIP_MODEL_FACTORY is not None:
        return CLIP_MODEL_FACTORY(cfg, trainer_name)
    import importlib
    mod = importlib.import_module(f"trainers.{trainer_name.lower()}")
    return mod.load_clip_to_cpu(cfg)


CLIP_MODEL_FACTORY = None   # callable(cfg, trainer_name) -> reference-keyed CLIP module (fp32)
TOKEN_FACTORY = None        # callable(cfg, classnames) -> int64 [C, 77] tokenised prompts, or None for the reference tokenizer


class _PromptTrainer(TrainerX):
    MODEL_NAME = "prompt_learner"

    def _tokens(self, classnames):
        return TOKEN_FACTORY(self.cfg, classnames) if TOKEN_FACTORY is not None else None

    def _train_batch_size(self):
        """The batch size Dassl's train loader will deliver: BATCH_SIZE, or -- a train set SMALLER than BATCH_SIZE -- the
        whole set as one short batch per epoch, which is not dropped (data_manager.py:37-48).  None = cfg's BATCH_SIZE."""
        bs = int(self.cfg.DATALOADER.TRAIN_X.BATCH_SIZE)
        train_x = getattr(getattr(self.dm, "dataset", None), "train_x", None)
        return len(train_x) if train_x is not None and 0 < len(train_x) < bs else None

    def _init_weights(self, module):
        """`if cfg.MODEL.INIT_WEIGHTS: load_pretrained_weights(self.model, ...)` of every reference build_model
        (promptsrc.py:253-254, coop.py:424-426; dassl/utils/torchtools.py:267-311: matching names and sizes are loaded,
        the rest ignored).  The prompt tensors land in the flat buffer through their nn.Parameter views.  The frozen CLIP
        tensors were already converted to the kernel layout when the module was built, so a file that would CHANGE one of
        them is refused instead of being half-applied."""
        path = _get(self.cfg, "MODEL.INIT_WEIGHTS", "")
        if not path:
            return
        ckpt = torch.load(path, map_location="cpu", weights_only=False)
        sd = ckpt["state_dict"] if "state_dict" in ckpt else ckpt
        own = module.state_dict()
        trainable = {k for k, p_ in module.named_parameters() if p_.requires_grad}
        take, frozen_changed = {}, []
        for k, v in sd.items():
            k = k[7:] if k.startswith("module.") else k
            if k in own and own[k].size() == v.size():
                if k in trainable:
                    take[k] = v
                elif not torch.equal(own[k].detach().cpu().to(v.dtype), v.cpu()):
                    frozen_changed.append(k)
        if frozen_changed:
            raise NotImplementedError(f"MODEL.INIT_WEIGHTS would overwrite frozen CLIP tensors ({frozen_changed[:3]} ...) "
                                      "after their conversion to the kernel layout; load them into the CLIP model instead")
        if not take:
            import warnings
            warnings.warn(f"Cannot load {path} (check the key names manually)")
        module.load_state_dict(take, strict=False)
        self.model.params_changed()
        print(f'Successfully loaded pretrained weights from "{path}" ({len(take)} trainable tensors)')

    def _finish_build(self, registered_module):
        cfg = self.cfg
        self._init_weights(registered_module)          # the module the reference passes (coop.py:424-426: prompt_learner)
        self.optim = self.model.make_optimizer(cfg.OPTIM)                    # SGD of optim/optimizer.py:105-113
        self.sched = build_scheduler(self.optim, cfg.OPTIM)
        self.register_model(self.MODEL_NAME, registered_module, self.optim, self.sched)
        self.model.input_kwargs = self._input_kwargs()                  # cfg.INPUT.* for the on-device input pipeline
        loader = getattr(self, "train_loader_x", None)                  # built before the model (trainer.py:325)
        if loader is not None and not isinstance(loader, LookaheadLoader) and _get(cfg, "TRAINER.B200.LOOKAHEAD", True):
            self.train_loader_x = LookaheadLoader(loader)

    def build_data_loader(self):
        """`TRAINER.B200.DEVICE_INPUT = True`: Dassl's own DataManager (datasets, samplers, loaders, worker processes
        unchanged -- dassl/engine/trainer.py:341-358) with `input_pipeline.DeviceTransform` as `custom_tfm_train` /
        `custom_tfm_test` (data_manager.py:64-85): the workers draw the reference's augmentation parameters and return
        the uint8 window, the resampling runs on the GPU.  `TRAINER.B200.DEVICE_INPUT_CANVAS` (default 512) bounds the
        window; see input_pipeline.py for what happens to larger ones.  Default False = the reference's PIL chain."""
        cfg = self.cfg
        if not bool(_get(cfg, "TRAINER.B200.DEVICE_INPUT", False)):
            return super().build_data_loader()
        import dassl.engine.trainer as _dtr            # the name SimpleTrainer.build_data_loader itself resolves
        from .input_pipeline import DeviceTransform
        c = int(_get(cfg, "TRAINER.B200.DEVICE_INPUT_CANVAS", 512))
        dm = _dtr.DataManager(cfg, custom_tfm_train=DeviceTransform.from_cfg(cfg, True, canvas=(c, c)),
                              custom_tfm_test=DeviceTransform.from_cfg(cfg, False, canvas=(c, c)))
        self.train_loader_x, self.train_loader_u = dm.train_loader_x, dm.train_loader_u
        self.val_loader, self.test_loader = dm.val_loader, dm.test_loader
        self.num_classes, self.num_source_domains, self.lab2cname = dm.num_classes, dm.num_source_domains, dm.lab2cname
        self.dm = dm

    def _input_kwargs(self):
        cfg = self.cfg
        kw = {"interpolation": str(_get(cfg, "INPUT.INTERPOLATION", "bicubic"))}
        if "normalize" in list(_get(cfg, "INPUT.TRANSFORMS", ["normalize"])):
            mean, std = _get(cfg, "INPUT.PIXEL_MEAN", None), _get(cfg, "INPUT.PIXEL_STD", None)
            if mean is not None and std is not None:
                kw.update(mean=tuple(float(v) for v in mean), std=tuple(float(v) for v in std))
        else:
            kw.update(mean=(0.0, 0.0, 0.0), std=(1.0, 1.0, 1.0))
        return kw

    def model_inference(self, input):
        return self.model(input)

    def parse_batch_train(self, batch):
        # no `.to(device)` round trip: the fused step copies the (pinned) host batch straight into its buffers
        return batch["img"], batch["label"]

    def parse_batch_test(self, batch):
        if isinstance(batch["img"], dict):             # DEVICE_INPUT: the model turns the uint8 windows into images
            return batch["img"], batch["label"].to(self.device)
        return super().parse_batch_test(batch)

    def save_model(self, epoch, directory, is_best=False, val_result=None, model_name=""):
        """Dassl's save_model (trainer.py:118-145) writes the registered module's whole state_dict -- for PromptSRC /
        MaPLe the complete frozen CLIP (~0.6 GB at ViT-B/16) next to 46 080 trained scalars.  With
        `TRAINER.B200.SLIM_CHECKPOINT = True` only the trainable tensors are written; the file keeps the
        reference's layout and key names, and both this package and the reference load it (`load_model` uses
        strict=False and the frozen weights come from the CLIP checkpoint anyway).  Default False: byte-for-byte the
        reference's behaviour."""
        if not bool(_get(self.cfg, "TRAINER.B200.SLIM_CHECKPOINT", False)):
            return super().save_model(epoch, directory, is_best=is_best, val_result=val_result, model_name=model_name)
        full = {}
        for name in self.get_model_names():
            m = self._models[name]
            full[name] = m.state_dict
            keep = {k for k, p_ in m.named_parameters() if p_.requires_grad}
            m.state_dict = (lambda sd_fn, keep_: (lambda *a, **k: {kk: v for kk, v in sd_fn(*a, **k).items()
                                                                    if kk in keep_}))(m.state_dict, keep)
        try:
            return super().save_model(epoch, directory, is_best=is_best, val_result=val_result, model_name=model_name)
        finally:
            for name, fn in full.items():
                del self._models[name].state_dict          # back to the class's method

    def load_model(self, directory, epoch=None):                             # promptsrc.py:390-420
        if not directory:
            print("Note that load_model() is skipped as no pretrained model is given")
            return
        model_file = "model-best.pth.tar" if epoch is None else "model.pth.tar-" + str(epoch)
        for name in self.get_model_names():
            path = osp.join(directory, name, model_file)
            if not osp.exists(path):
                raise FileNotFoundError('Model not found at "{}"'.format(path))
            ckpt = torch.load(path, map_location="cpu", weights_only=False)
            sd = ckpt["state_dict"]
            for k in ("prompt_learner.token_prefix", "prompt_learner.token_suffix", "token_prefix", "token_suffix"):
                sd.pop(k, None)                                              # ignore fixed token vectors
            print('Loading weights to {} from "{}" (epoch = {})'.format(name, path, ckpt["epoch"]))
            self._models[name].load_state_dict(sd, strict=False)


@TRAINER_REGISTRY.register(force=True)
class PromptSRC(_PromptTrainer):
    """Drop-in for trainers/promptsrc.py:216-420."""
    MODEL_NAME = "VLPromptLearner"

    def check_cfg(self, cfg):
        assert cfg.TRAINER.PROMPTSRC.PREC in ["fp16", "fp32", "amp"]

    def build_model(self):
        cfg = self.cfg
        classnames = self.dm.dataset.classnames
        clip_model = _load_clip(cfg, "PromptSRC")
        clip_model.float()
        self.model = PromptSRCCustomCLIP(cfg, classnames, clip_model, tokenized_prompts=self._tokens(classnames),
                                         device=self.device, train_batch=self._train_batch_size())
        self._finish_build(self.model)
        self.total_epochs = cfg.OPTIM.MAX_EPOCH
        self.step_counter = 1
        # Gaussian weights of the prompt aggregation (promptsrc.py:268-273)
        import numpy as np
        mean, std = cfg.TRAINER.PROMPTSRC.GPA_MEAN, cfg.TRAINER.PROMPTSRC.GPA_STD
        g = np.array([(1 / (std * np.sqrt(2 * np.pi))) * np.exp(-0.5 * ((a - mean) / std) ** 2)
                      for a in range(1, cfg.OPTIM.MAX_EPOCH + 1)])
        self.gauss = g / g.sum()
        self.previous_model_gpa = None

    def forward_backward(self, batch):
        image, label = self.parse_batch_train(batch)
        losses = self.model.train_step_fused(image, label, self.optim, _next_of(batch))
        loss_summary = {"loss": self.model.step_loss(losses)}                  # promptsrc.py:330
        last = (self.batch_idx + 1) == self.num_batches
        if _get(self.cfg, "TRAINER.PROMPTSRC.USE_GPA", False):               # promptsrc.py:335-360
            if last:
                self.update_lr()
                self.step_counter += 1
                self._gpa_accumulate(float(self.gauss[self.step_counter - 2]))
            if self.step_counter == self.model.total_epochs + 1:
                print("Using GPA model for final inference...")
                self.model._step.flat_param.copy_(self.previous_model_gpa)
                self.model.params_changed()              # text prefix enqueued ahead / cached eval text: pre-GPA prompts
        elif last:
            self.update_lr()
        return loss_summary

    def _gpa_accumulate(self, w):
        """previous_model_gpa += w * state (promptsrc.py:341-347) on the trainable prompt buffer only: the frozen
        tensors are multiplied by weights that sum to 1, i.e. unchanged."""
        from . import ops
        st = self.model._step
        if self.previous_model_gpa is None:
            self.previous_model_gpa = torch.zeros_like(st.flat_param)
        ops.axpy(self.previous_model_gpa, st.flat_param, w)


@TRAINER_REGISTRY.register(force=True)
class CoOp(_PromptTrainer):
    """Drop-in for trainers/coop.py:396-500 (LOSS_TYPE ce / focal)."""
    MODEL_NAME = "prompt_learner"

    def check_cfg(self, cfg):
        assert cfg.TRAINER.COOP.PREC in ["fp16", "fp32", "amp"]

    def build_model(self):
        cfg = self.cfg
        classnames = self.dm.dataset.classnames
        clip_model = _load_clip(cfg, "CoOp")
        clip_model.float()
        self.model = CoOpCustomCLIP(cfg, classnames, clip_model, tokenized_prompts=self._tokens(classnames),
                                    device=self.device, train_batch=self._train_batch_size())
        self._finish_build(self.model.prompt_learner)                        # coop.py:431

    def forward_backward(self, batch):
        image, label = self.parse_batch_train(batch)
        losses = self.model.train_step_fused(image, label, self.optim, _next_of(batch))
        loss_summary = {"loss": self.model.step_loss(losses)}
        if self.model.loss_type == "ce":                                     # coop.py:461-469: accuracy of the UPDATED prompts
            st = self.model._step
            with torch.no_grad():
                st.text_features()
                st.compute_logits()                                          # image features are prompt-independent here
                pred = st.logits.argmax(1)
                loss_summary["acc"] = float((pred == st.labels.long()).float().mean().item() * 100.0)
        if (self.batch_idx + 1) == self.num_batches:
            self.update_lr()
        return loss_summary


@TRAINER_REGISTRY.register(force=True)
class MaPLe(_PromptTrainer):
    """Drop-in for trainers/maple.py:257-357."""
    MODEL_NAME = "MultiModalPromptLearner"

    def check_cfg(self, cfg):
        assert cfg.TRAINER.MAPLE.PREC in ["fp16", "fp32", "amp"]

    def build_model(self):
        cfg = self.cfg
        classnames = self.dm.dataset.classnames
        clip_model = _load_clip(cfg, "MaPLe")
        clip_model.float()
        self.model = MaPLeCustomCLIP(cfg, classnames, clip_model, tokenized_prompts=self._tokens(classnames),
                                     device=self.device, train_batch=self._train_batch_size())
        self._finish_build(self.model)                                       # maple.py:301 registers the whole model

    def forward_backward(self, batch):
        image, label = self.parse_batch_train(batch)
        losses = self.model.train_step_fused(image, label, self.optim, _next_of(batch))
        loss_summary = {"loss": self.model.step_loss(losses)}                  # maple.py:325
        if (self.batch_idx + 1) == self.num_batches:
            self.update_lr()
        return loss_summary


@TRAINER_REGISTRY.register(force=True)
class CoCoOp(_PromptTrainer):
    """Drop-in for trainers/cocoop.py:262-372."""
    MODEL_NAME = "prompt_learner"

    def check_cfg(self, cfg):
        assert cfg.TRAINER.COCOOP.PREC in ["fp16", "fp32", "amp"]

    def build_model(self):
        cfg = self.cfg
        classnames = self.dm.dataset.classnames
        clip_model = _load_clip(cfg, "CoCoOp")
        clip_model.float()
        self.model = CoCoOpCustomCLIP(cfg, classnames, clip_model, tokenized_prompts=self._tokens(classnames),
                                      device=self.device, train_batch=self._train_batch_size())
        self._finish_build(self.model.prompt_learner)                        # cocoop.py:304

    def forward_backward(self, batch):
        image, label = self.parse_batch_train(batch)
        losses = self.model.train_step_fused(image, label, self.optim, _next_of(batch))
        loss_summary = {"loss": self.model.step_loss(losses)}                  # cocoop.py:328
        if (self.batch_idx + 1) == self.num_batches:
            self.update_lr()
        return loss_summary
