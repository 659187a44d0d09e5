"""Shared test plumbing: build a PromptedClipStep from a golden fixture, reference-name mapping,
and a shape-checking fake of the kernel layer for CPU-only host-logic tests."""
from __future__ import annotations

import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def load_golden(name):
    return dict(np.load(os.path.join(GOLD, name + ".npz"), allow_pickle=False))


def ref_name(key: str) -> str:
    """step parameter key -> reference state-dict / named_parameters key."""
    if key == "ctx":
        return "prompt_learner.ctx"
    if key == "vpt":
        return "image_encoder.VPT"
    side, i = key.split(".")
    enc = "text_encoder" if side == "text" else "image_encoder"
    return f"{enc}.transformer.resblocks.{i}.VPT_shallow"


def cos(a, b):
    a, b = torch.as_tensor(a).flatten().double().cpu(), torch.as_tensor(b).flatten().double().cpu()
    return float(a @ b / (a.norm() * b.norm() + 1e-300))


def promptsrc_setup(arch_name, depth, tokens, device, prec, batch, sd=None, fixed=None, world=1, rank=0, pg=None,
                    focal_alpha=None, truncate=True):
    """The PromptSRC configuration of oracle/make_golden.py: weights seed 0, deep prompts seed 100,
    ctx = token embedding of the prompt words (promptsrc.py:90-98)."""
    from psrc_b200 import synthetic as S
    from psrc_b200.step import PromptedClipStep, StepConfig, encode_text_frozen
    arch = S.ARCHS[arch_name]
    sd = sd if sd is not None else S.make_clip_state_dict(arch, seed=0)
    tokens = torch.as_tensor(np.asarray(tokens)).long()
    n_ctx = 4
    cfg = StepConfig(arch=arch, n_cls=tokens.shape[0], batch=batch, prec=prec, n_ctx_text=n_ctx, n_ctx_vision=n_ctx,
                     depth_text=depth, depth_vision=depth, zero_shot_branch=True, world_size=world, rank=rank,
                     process_group=pg, focal_alpha=focal_alpha, truncate_text=truncate)
    if fixed is None:
        fixed = encode_text_frozen(sd, tokens, device)
    step = PromptedClipStep(cfg, sd, tokens, device, fixed_embeddings=fixed)
    deep = S.make_deep_prompts(arch, n_ctx, n_ctx, depth, depth, seed=100)
    with torch.no_grad():
        step.params["ctx"].copy_(sd["token_embedding.weight"][tokens[0, 1:1 + n_ctx]])
        step.params["vpt"].copy_(deep["visual.VPT"])
        for i in range(1, depth):
            step.params[f"vision.{i}"].copy_(deep[f"visual.transformer.resblocks.{i}.VPT_shallow"])
            step.params[f"text.{i}"].copy_(deep[f"transformer.resblocks.{i}.VPT_shallow"])
    return step, sd, fixed


def maple_ref_name(key: str) -> str:
    """MaPLe step key -> reference named_parameters key (maple.py:145-156)."""
    if key == "ctx":
        return "prompt_learner.ctx"
    parts = key.split(".")
    if parts[0] == "text":
        return f"prompt_learner.compound_prompts_text.{int(parts[1]) - 1}"
    kind = "weight" if parts[1] == "w" else "bias"
    i = int(parts[2])
    return f"prompt_learner.proj.{kind}" if i == 0 else f"prompt_learner.compound_prompt_projections.{i - 1}.{kind}"


def maple_setup(arch_name, depth, tokens, device, prec, batch, n_ctx=2, sd=None, world=1, rank=0, pg=None,
                focal_alpha=None, ctx_init=None):
    """The MaPLe configuration of oracle/make_golden.py:run_maple: weights seed 0, coupling parameters seed 200,
    ctx = embedding of the first n_ctx prompt words (maple.py:125-131)."""
    from psrc_b200 import synthetic as S
    from psrc_b200.step import PromptedClipStep, StepConfig
    arch = S.ARCHS[arch_name]
    sd = sd if sd is not None else S.make_clip_state_dict(arch, seed=0)
    tokens = torch.as_tensor(np.asarray(tokens)).long()
    cfg = StepConfig(arch=arch, n_cls=tokens.shape[0], batch=batch, prec=prec, n_ctx_text=n_ctx, n_ctx_vision=n_ctx,
                     depth_text=depth, depth_vision=depth, coupling=True, world_size=world, rank=rank,
                     process_group=pg, focal_alpha=focal_alpha)
    step = PromptedClipStep(cfg, sd, tokens, device)
    mp_ = S.make_maple_params(arch, n_ctx, depth, seed=200)
    with torch.no_grad():
        step.params["ctx"].copy_(ctx_init if ctx_init is not None
                                 else sd["token_embedding.weight"][tokens[0, 1:1 + n_ctx]])
        for k in step.params:
            if k != "ctx":
                step.params[k].copy_(mp_[maple_ref_name(k)[len("prompt_learner."):]])
    return step, sd, mp_


def coop_setup(arch_name, tokens, device, prec, batch, ctx0, n_ctx=16, sd=None, focal_alpha=None, position="end",
               csc=False):
    """The CoOp configuration of oracle/make_golden.py:run_coop (16 generic context tokens, seeded ctx from the fixture)."""
    from psrc_b200 import synthetic as S
    from psrc_b200.step import PromptedClipStep, StepConfig
    arch = S.ARCHS[arch_name]
    sd = sd if sd is not None else S.make_clip_state_dict(arch, seed=0)
    tokens = torch.as_tensor(np.asarray(tokens)).long()
    cfg = StepConfig(arch=arch, n_cls=tokens.shape[0], batch=batch, prec=prec, n_ctx_text=n_ctx, n_ctx_vision=0,
                     depth_text=1, depth_vision=0, focal_alpha=focal_alpha, class_token_position=position,
                     ctx_per_class=csc)
    step = PromptedClipStep(cfg, sd, tokens, device)
    with torch.no_grad():
        step.params["ctx"].copy_(torch.as_tensor(ctx0))
    return step, sd


COCOOP_REF = {"ctx": "prompt_learner.ctx", "meta.w1": "prompt_learner.meta_net.linear1.weight",
              "meta.b1": "prompt_learner.meta_net.linear1.bias", "meta.w2": "prompt_learner.meta_net.linear2.weight",
              "meta.b2": "prompt_learner.meta_net.linear2.bias"}


def cocoop_setup(arch_name, tokens, device, prec, batch, n_ctx=4, sd=None, world=1, rank=0, pg=None,
                 focal_alpha=None):
    """The CoCoOp configuration of oracle/make_golden.py:run_cocoop: weights seed 0, meta-net seed 300,
    ctx = embedding of "a photo of a" (cocoop.py:124-131)."""
    from psrc_b200 import synthetic as S
    from psrc_b200.step import PromptedClipStep, StepConfig
    arch = S.ARCHS[arch_name]
    sd = sd if sd is not None else S.make_clip_state_dict(arch, seed=0)
    tokens = torch.as_tensor(np.asarray(tokens)).long()
    cfg = StepConfig(arch=arch, n_cls=tokens.shape[0], batch=batch, prec=prec, n_ctx_text=n_ctx, n_ctx_vision=0,
                     depth_text=1, depth_vision=0, per_image_text=True, meta_hidden=arch.embed_dim // 16,
                     world_size=world, rank=rank, process_group=pg, focal_alpha=focal_alpha)
    step = PromptedClipStep(cfg, sd, tokens, device)
    mp_ = S.make_cocoop_params(arch, seed=300)
    with torch.no_grad():
        step.params["ctx"].copy_(sd["token_embedding.weight"][tokens[0, 1:1 + n_ctx]])
        for k in step.params:
            if k != "ctx":
                step.params[k].copy_(mp_[COCOOP_REF[k][len("prompt_learner."):]])
    return step, sd, mp_


# ---------------------------------------------------------------------------------------------
class FakeKernels:
    """Replaces psrc_b200.ops.* with shape/dtype-checking no-ops that work on CPU tensors, so the host
    orchestration (buffer shapes, launch order, sharding arithmetic) can be exercised without a GPU.
    Never used on a GPU box and never part of the product."""

    def __init__(self):
        self.calls = []

    def install(self, monkeypatch):
        from psrc_b200 import ops
        monkeypatch.setattr(ops, "_FAKE_KERNELS", True, raising=False)
        for name in ("gemm", "gemm_f32_strided", "layernorm_fwd", "layernorm_bwd", "attn_fwd", "attn_bwd",
                     "attn_row_fwd", "attn_row_bwd", "im2col_patches", "assemble_vision_tokens", "assemble_text_prompts", "insert_prompt",
                     "prompt_grad", "gather_ln_fwd", "gather_ln_bwd", "l2norm_fwd", "l2norm_bwd", "loss_fwd_bwd",
                     "sgd_step", "cast", "axpy", "small_linear_fwd", "small_linear_bwd", "group_logits_fwd",
                     "group_logits_bwd", "ctx_bias_grad", "permute_rows"):
            monkeypatch.setattr(ops, name, getattr(self, name))

    def permute_rows(self, src, perm, dst, n_seq, L, D, add=None, inverse=False):
        self._rec("permute_rows")
        assert src.shape == dst.shape == (n_seq * L, D) and perm.shape == (n_seq, L)
        return dst

    def small_linear_fwd(self, x, w, b, y, relu=False):
        self._rec("small_linear_fwd")
        G, R, K = x.shape
        assert w.shape[0] == G and w.shape[2] == K and y.shape == (G, R, w.shape[1])
        return y

    def small_linear_bwd(self, x, w, dy, dx=None, dw=None, db=None, y_relu=None, accumulate_dx=False):
        self._rec("small_linear_bwd")
        assert dy.shape == (x.shape[0], x.shape[1], w.shape[1])
        assert dx is None or dx.shape == x.shape
        assert dw is None or dw.shape == w.shape

    def group_logits_fwd(self, f_hat, t_hat, logits, scale):
        self._rec("group_logits_fwd")
        assert t_hat.shape == (logits.numel(), f_hat.shape[1])
        return logits

    def group_logits_bwd(self, f_hat, d_logits, d_that, scale):
        self._rec("group_logits_bwd")
        assert d_that.shape == (d_logits.numel(), f_hat.shape[1])
        return d_that

    def ctx_bias_grad(self, gg, dctx, dbias):
        self._rec("ctx_bias_grad")
        assert dctx.shape == gg.shape[1:] and dbias.shape == (gg.shape[0], gg.shape[2])

    def _rec(self, name):
        self.calls.append(name)

    def gemm(self, a, w, out, bias=None, act=0, aux=None, residual=None, out2=None, alpha=1.0, backend=0):
        self._rec("gemm")
        M, K = a.shape
        N = w.shape[0]
        assert w.shape == (N, K) and out.shape == (M, N) and a.dtype == w.dtype, (a.shape, w.shape, out.shape)
        assert a.stride(1) == 1 and w.stride(1) == 1 and out.stride(1) == 1
        if bias is not None:
            assert bias.shape == (N,) and bias.dtype == torch.float32
        if residual is not None:
            assert residual.shape == (M, N) and residual.dtype == torch.float32 and out.dtype == torch.float32
        if aux is not None:
            assert aux.shape == (M, N)
        if out2 is not None:
            assert out2.shape == (M, N)
        return out

    def gemm_f32_strided(self, M, N, K, a, a_sm, a_sk, b, b_sn, b_sk, c, alpha=1.0, beta=0.0):
        self._rec("gemm_f32_strided")
        assert c.shape == (M, N)
        assert (M - 1) * a_sm + (K - 1) * a_sk < a.numel() and (N - 1) * b_sn + (K - 1) * b_sk < b.numel()
        return c

    def layernorm_fwd(self, x, gamma, beta, y, mean=None, rstd=None, eps=1e-5):
        self._rec("layernorm_fwd")
        assert x.shape == y.shape and gamma.shape == (x.shape[1],) and x.dtype == torch.float32
        assert mean is None or mean.shape == (x.shape[0],)
        return y

    def layernorm_bwd(self, dy, x, gamma, mean, rstd, dx_out, dx_in=None, dx_act=None):
        self._rec("layernorm_bwd")
        assert dy.shape == x.shape == dx_out.shape and mean.shape == (x.shape[0],)
        assert dx_act is None or dx_act.shape == x.shape
        return dx_out

    def attn_fwd(self, qkv, out, lse, n_seq, L, H, causal, backend=0):
        self._rec("attn_fwd")
        assert qkv.shape == (n_seq * L, 3 * H * 64) and out.shape == (n_seq * L, H * 64) and lse.shape == (n_seq, H, L)
        return out

    def attn_bwd(self, qkv, out, dout, lse, dqkv, n_seq, L, H, causal, backend=0):
        self._rec("attn_bwd")
        assert qkv.shape == dqkv.shape == (n_seq * L, 3 * H * 64) and out.shape == dout.shape == (n_seq * L, H * 64)
        return dqkv

    def attn_row_fwd(self, qkv, out, lse, n_seq, L, H, causal, idx=None):
        self._rec("attn_row_fwd")
        assert qkv.shape == (n_seq * L, 3 * H * 64) and out.shape == (n_seq * L, H * 64) and lse.shape == (n_seq, H, L)
        return out

    def attn_row_bwd(self, qkv, dout, lse, dqkv, n_seq, L, H, causal, idx=None):
        self._rec("attn_row_bwd")
        assert qkv.shape == dqkv.shape == (n_seq * L, 3 * H * 64) and dout.shape == (n_seq * L, H * 64)
        return dqkv

    def im2col_patches(self, image, patches, R, P):
        self._rec("im2col_patches")
        B = image.shape[0]
        assert image.shape == (B, 3, R, R) and patches.shape == (B * (R // P) ** 2, 3 * P * P)
        return patches

    def assemble_vision_tokens(self, pe, cls, pos, vpt, ln_g, ln_b, x, B, n_patches, n_ctx, D, eps=1e-5):
        self._rec("assemble_vision_tokens")
        assert pe.shape == (B * n_patches, D) and x.shape == (B * (1 + n_patches + n_ctx), D)
        assert pos.shape == (n_patches + 1, D) and (n_ctx == 0 or vpt.shape == (n_ctx, D))
        return x

    def assemble_text_prompts(self, prefix, suffix, ctx, pos, x, C_, L, n_ctx, D, ctx_per_class=False, ctx_bias=None,
                              n_groups=1):
        self._rec("assemble_text_prompts")
        assert prefix.shape == (C_, 1, D) and suffix.shape[0] == C_ and suffix.shape[2] == D
        assert L - 1 - n_ctx <= suffix.shape[1] and x.shape == (n_groups * C_ * L, D)
        if n_ctx:
            assert ctx.shape == ((C_, n_ctx, D) if ctx_per_class else (n_ctx, D))
        return x

    def insert_prompt(self, x, prompt, n_seq, L, start, n_ctx, D, round_fp16=True):
        self._rec("insert_prompt")
        assert x.shape == (n_seq * L, D) and prompt.shape == (n_ctx, D) and start + n_ctx <= L
        return x

    def prompt_grad(self, dx, grad, n_seq, L, start, n_ctx, D, zero_rows=True, accumulate=False, dx_act=None,
                    round_fp16=False):
        self._rec("prompt_grad")
        assert dx.shape == (n_seq * L, D) and grad.shape == (n_ctx, D) and start + n_ctx <= L
        return grad

    def gather_ln_fwd(self, x, idx, gamma, beta, y, mean, rstd, n_seq, L, D, eps=1e-5):
        self._rec("gather_ln_fwd")
        assert x.shape == (n_seq * L, D) and y.shape == (n_seq, D)
        assert idx is None or (idx.shape == (n_seq,) and idx.dtype == torch.int32 and int(idx.max()) < L)
        return y

    def gather_ln_bwd(self, dy, x, idx, gamma, mean, rstd, dx, n_seq, L, D, dx_act=None, dx_in=None):
        self._rec("gather_ln_bwd")
        assert dy.shape == (n_seq, D) and x.shape == dx.shape == (n_seq * L, D) and mean.shape == (n_seq,)
        return dx

    def l2norm_fwd(self, x, y, inv_norm):
        self._rec("l2norm_fwd")
        assert x.shape == y.shape and (inv_norm is None or inv_norm.shape == (x.shape[0],))
        return y

    def l2norm_bwd(self, dy, y, inv_norm, dx):
        self._rec("l2norm_bwd")
        assert dy.shape == y.shape == dx.shape
        return dx

    def loss_fwd_bwd(self, logits, labels, d_logits, losses, scratch, zs_logits=None, focal_alpha=None,
                     focal_gamma=2.0, t_hat=None, t_zs=None, d_that=None, f_hat=None, f_zs=None, d_fhat=None,
                     w_kl=1.0, w_text=25.0, w_image=10.0, B_global=None, Ct_global=None):
        self._rec("loss_fwd_bwd")
        B, C = logits.shape
        assert labels.shape == (B,) and labels.dtype == torch.int32 and d_logits.shape == (B, C)
        assert zs_logits is None or zs_logits.shape == (B, C)
        if t_hat is not None:
            assert t_hat.shape == t_zs.shape == d_that.shape
        if f_hat is not None:
            assert f_hat.shape == f_zs.shape == d_fhat.shape
        assert focal_alpha is None or focal_alpha.shape == (C,)
        return losses

    def sgd_step(self, p, g, buf, lr_dev, *a):
        self._rec("sgd_step")
        assert p.shape == g.shape
        return p

    def cast(self, src, dst):
        self._rec("cast")
        assert src.numel() == dst.numel()
        return dst

    def axpy(self, y, x, a=1.0):
        self._rec("axpy")
        assert y.shape == x.shape
        return y


# ---------------------------------------------------------------------------------------------
class Cfg(dict):
    """Attribute-access dict standing in for yacs.CfgNode in tests (the trainers only read keys)."""

    def __init__(self, d=None):
        super().__init__()
        for k, v in (d or {}).items():
            self[k] = Cfg(v) if isinstance(v, dict) else v

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


def trainer_cfg(trainer="PromptSRC", batch=3, test_batch=5, max_epoch=4, depth=3, prec="fp32", **over):
    """The keys train.py:extend_cfg + the PromptSRC/CoOp yamls define, at test sizes."""
    c = Cfg({
        "DATALOADER": {"TRAIN_X": {"BATCH_SIZE": batch}, "TEST": {"BATCH_SIZE": test_batch}},
        "DATASET": {"PER_CLASS_SHOTS": []},
        "INPUT": {"SIZE": (64, 64)},
        "OPTIM": {"NAME": "sgd", "LR": 0.0025, "MAX_EPOCH": max_epoch, "LR_SCHEDULER": "cosine", "WARMUP_EPOCH": 1,
                  "WARMUP_TYPE": "constant", "WARMUP_CONS_LR": 1e-5, "WARMUP_RECOUNT": True, "MOMENTUM": 0.9,
                  "WEIGHT_DECAY": 5e-4, "SGD_DAMPNING": 0.0, "SGD_NESTEROV": False, "STEPSIZE": (-1,), "GAMMA": 0.1},
        "TRAINER": {
            "NAME": trainer,
            "PROMPTSRC": {"N_CTX_VISION": 4, "N_CTX_TEXT": 4, "CTX_INIT": "a photo of a", "PREC": prec,
                          "PROMPT_DEPTH_VISION": depth, "PROMPT_DEPTH_TEXT": depth, "TEXT_LOSS_WEIGHT": 25,
                          "IMAGE_LOSS_WEIGHT": 10, "GPA_MEAN": 2, "GPA_STD": 1, "LOSS_TYPE": "ce"},
            "COOP": {"N_CTX": 4, "CSC": False, "CTX_INIT": "", "PREC": prec, "CLASS_TOKEN_POSITION": "end",
                     "LOSS_TYPE": "ce"},
            "MAPLE": {"N_CTX": 2, "CTX_INIT": "a photo of a", "PREC": prec, "PROMPT_DEPTH": depth,
                      "USE_FOCAL_LOSS": False},
            "COCOOP": {"N_CTX": 4, "CTX_INIT": "a photo of a", "PREC": prec, "USE_FOCAL_LOSS": False},
            "B200": {"CUDA_GRAPH": True},
        },
    })
    for k, v in over.items():
        node = c
        parts = k.split(".")
        for p in parts[:-1]:
            node = node[p]
        node[parts[-1]] = v
    return c


class FakeDM:
    def __init__(self, classnames):
        self.dataset = Cfg({})
        self.dataset.classnames = list(classnames)
