"""The prompted-CLIP training step (forward, losses, prompt-only backward) over the kernels.

`PromptedClipStep` is the engine behind the Dassl-facing `CustomCLIP` modules in
`trainers.py`.  It covers the PromptSRC step (CustomCLIP.forward, trainers/promptsrc.py:183-211,
plus the loss of forward_backward :299-324) and, with the frozen-teacher branch and the vision
prompts switched off, the CoOp step (trainers/coop.py:351-390).

Multi-GPU: one process per GPU; the image batch is data-parallel, the class prompts
are sharded by class.  Forward: all-gather of the normalised text features.  Backward:
reduce-scatter of dL/dt_hat (each rank holds a partial sum over its images), then an all-reduce
of the flat prompt-gradient buffer.  Loss terms are pre-scaled by the GLOBAL batch / class count
so the summed result equals the single-device step.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass, field

import torch
import torch.distributed as dist

from . import ops
from .engine import ACT_DTYPES, Tower, TowerWeights, _dev, tower_dtypes
from .synthetic import ClipArch


@dataclass
class StepConfig:
    arch: ClipArch
    n_cls: int
    batch: int                      # LOCAL batch (per rank)
    prec: str = "bf16"              # "bf16" fast mode | "fp32" validation mode
    n_ctx_text: int = 4
    n_ctx_vision: int = 0           # 0: no vision prompts (CoOp) -> no vision backward
    depth_text: int = 1             # layers 1..depth-1 carry deep prompts (model.py:343-349)
    depth_vision: int = 0
    ctx_per_class: bool = False     # CoOp CSC (coop.py:222-230)
    zero_shot_branch: bool = False  # PromptSRC frozen teacher: ZS image tower + fixed text embeddings
    w_text: float = 25.0
    w_image: float = 10.0
    w_logits: float = 1.0
    focal_alpha: list | None = None  # imbalance-aware CE (coop.py:326-346); None -> plain CE
    focal_gamma: float = 2.0
    fwd16: str = field(default_factory=lambda: os.environ.get("PSRC_FWD16", "fp16"))
    # ^ fast mode: 16-bit format of the FORWARD operands ("fp16": IEEE half, CLIP's native format, 3 more mantissa bits
    #   than bf16 at the same tensor-core rate; "bf16").  Gradients are always bf16.  See engine.TowerDtypes.
    truncate_text: bool = True      # run text positions 0..max(EOT) only (exact)
    class_token_position: str = "end"   # CoOp: "end" | "middle" | "front" (coop.py:267-293)
    coupling: bool = False          # MaPLe: vision prompts = Linear_i(text prompts) (maple.py:145, :155-156, :199-204)
    per_image_text: bool = False    # CoCoOp: one text-encoder pass per image, ctx + meta_net(f_hat) (cocoop.py:173-198)
    meta_hidden: int = 0            # CoCoOp meta-net hidden width (vis_dim // 16, cocoop.py:139-143)
    world_size: int = 1
    rank: int = 0
    process_group: object = field(default=None, repr=False)


def shard_range(n: int, world: int, rank: int):
    """Contiguous class shard of rank `rank` (first n % world ranks get one extra)."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


@torch.no_grad()
def encode_text_frozen(clip_sd: dict, tokens: torch.Tensor, device, prec: str = "fp32", chunk: int = 2048):
    """Frozen CLIP.encode_text (clip/model.py:605-618) through the kernels: the PromptSRC teacher
    `fixed_embeddings` (trainers/promptsrc.py:122-132).  Init-time, so it runs in fp32 mode by default.
    tokens: int64 [N, 77].  Returns un-normalised features fp32 [N, E] on `device`."""
    from .synthetic import arch_from_state_dict
    a = arch_from_state_dict(clip_sd)
    dev = torch.device(device)
    tokens = tokens.to(torch.int64).cpu()
    dts = tower_dtypes(prec, int(tokens.argmax(-1).max()) + 1)
    act = dts.fwd
    tw = TowerWeights(clip_sd, "transformer.", a.transformer_layers, dev, dts, False)
    tproj_t = _dev(clip_sd["text_projection"].t(), dev, act)
    ln_g, ln_b = _dev(clip_sd["ln_final.weight"], dev), _dev(clip_sd["ln_final.bias"], dev)
    pos = _dev(clip_sd["positional_embedding"], dev)
    Dt, E = a.transformer_width, a.embed_dim
    out = torch.empty(tokens.shape[0], E, device=dev)
    for lo in range(0, tokens.shape[0], chunk):
        tk = tokens[lo:lo + chunk]
        n = tk.shape[0]
        eot = tk.argmax(-1)
        L = int(eot.max()) + 1
        emb = _dev(clip_sd["token_embedding.weight"][tk], dev)               # model.py:606
        tower = Tower(tw, n, L, Dt, a.transformer_heads, True, False, dev)
        prefix, suffix = emb[:, :1].contiguous(), emb[:, 1:].contiguous()
        ops.assemble_text_prompts(prefix, suffix, prefix, pos, tower.x0, n, L, 0, Dt)   # + positional (model.py:608)
        tower.forward()
        h = torch.empty(n, Dt, device=dev, dtype=act)
        ops.gather_ln_fwd(tower.x_final, eot.to(torch.int32).to(dev), ln_g, ln_b, h, None, None, n, L, Dt)
        ops.gemm(h, tproj_t, out[lo:lo + n])
    return out


class PromptedClipStep:
    def __init__(self, cfg: StepConfig, clip_sd: dict, tokens: torch.Tensor, device, fixed_embeddings=None):
        """clip_sd: reference-keyed CLIP state-dict (fp32).  tokens: int64 [C, 77] tokenised prompts."""
        self.cfg, self.device = cfg, torch.device(device)
        a = cfg.arch
        dev = self.device
        C, B = cfg.n_cls, cfg.batch
        # ---- sequence lengths first: they decide the storage formats of each tower (engine.TowerDtypes)
        tokens = tokens.to(torch.int64).cpu()
        assert tokens.shape[0] == C
        eot_all = tokens.argmax(-1)                                          # promptsrc.py:68
        self.Lt = int(eot_all.max()) + 1 if cfg.truncate_text else a.context_length
        self.Lv = 1 + a.n_patches + cfg.n_ctx_vision
        self.vdt = tower_dtypes(cfg.prec, self.Lv, cfg.fwd16)
        self.tdt = tower_dtypes(cfg.prec, self.Lt, cfg.fwd16)
        self.act = self.vdt.grad          # gradient operands (and the fp32 validation mode's only format)
        self.fwd = self.vdt.fwd           # forward operands outside the towers (patches, class-token / EOT rows)
        self.per_image = bool(cfg.per_image_text)
        if self.per_image:                        # CoCoOp: pure batch sharding, every rank runs all classes
            assert cfg.n_ctx_vision == 0 and cfg.depth_text == 1 and not cfg.zero_shot_branch and not cfg.ctx_per_class
            self.c0, self.c1 = 0, C
        else:
            # equal shards of ceil(C / R) classes (the collectives need equal chunks); when C is not a multiple of R the
            # tail of the last shard(s) is padding: dummy sequences (a copy of the last class) that run through the text
            # tower but never reach the logits, receive a zero gradient and are excluded from the L1-text term
            per = -(-C // cfg.world_size)
            self.c0 = cfg.rank * per
            self.c1 = self.c0 + per
            if cfg.world_size > 1 and per * cfg.world_size != C:
                assert not cfg.ctx_per_class, "class-specific contexts need a class count divisible by the world size"
        if cfg.coupling:
            assert cfg.n_ctx_vision == cfg.n_ctx_text and cfg.depth_vision == cfg.depth_text and not cfg.ctx_per_class, \
                "MaPLe couples one vision prompt to each text prompt (maple.py:150-156)"
        self.Cl = self.c1 - self.c0
        self.Cv = max(0, min(C, self.c1) - self.c0)            # real (non-padding) classes of this shard
        self.C_pad = self.Cl * cfg.world_size if not self.per_image else C
        cls_idx = torch.arange(self.c0, self.c1).clamp(max=C - 1)
        self.share_text = cfg.world_size > 1 and not self.per_image   # all-gather t_hat / reduce-scatter dt_hat
        self.B_global = B * cfg.world_size
        self.Dv, self.Dt, self.E = a.vision_width, a.transformer_width, a.embed_dim
        self.scale = float(math.exp(float(clip_sd["logit_scale"])))

        # ---- frozen weights in kernel layout
        need_vb = cfg.n_ctx_vision > 0
        self.vw = TowerWeights(clip_sd, "visual.transformer.", a.vision_layers, dev, self.vdt, need_vb)
        self.tw = TowerWeights(clip_sd, "transformer.", a.transformer_layers, dev, self.tdt, True)
        f = lambda k: _dev(clip_sd[k], dev)  # noqa: E731
        self.conv_w = _dev(clip_sd["visual.conv1.weight"].reshape(self.Dv, -1), dev, self.fwd)
        self.cls, self.vpos = f("visual.class_embedding"), f("visual.positional_embedding")
        self.ln_pre = (f("visual.ln_pre.weight"), f("visual.ln_pre.bias"))
        self.ln_post = (f("visual.ln_post.weight"), f("visual.ln_post.bias"))
        self.ln_final = (f("ln_final.weight"), f("ln_final.bias"))
        self.vproj = _dev(clip_sd["visual.proj"], dev, self.act)            # [Dv, E]  (dgrad operand)
        self.vproj_t = _dev(clip_sd["visual.proj"].t(), dev, self.fwd)      # [E, Dv]  (forward operand)
        self.tproj = _dev(clip_sd["text_projection"], dev, self.act)        # [Dt, E]
        self.tproj_t = _dev(clip_sd["text_projection"].t(), dev, self.fwd)  # [E, Dt]  (forward operand)
        self.tpos = f("positional_embedding")

        # ---- class tokens (this rank's shard): prefix / suffix embeddings, EOT rows
        emb = clip_sd["token_embedding.weight"][tokens[cls_idx]]             # promptsrc.py:121
        self.token_prefix = _dev(emb[:, :1], dev)                           # promptsrc.py:136
        self.token_suffix = _dev(emb[:, 1 + cfg.n_ctx_text:], dev)          # promptsrc.py:137
        self.eot = eot_all[cls_idx].to(torch.int32).to(dev)
        # CoOp class-token positions other than "end": the prompt is the "end" layout with its rows permuted per class
        self.text_perm = None
        if cfg.class_token_position != "end":
            assert cfg.class_token_position in ("middle", "front"), cfg.class_token_position
            assert cfg.depth_text == 1 and not self.per_image, "class-token positions are a CoOp option (coop.py:267)"
            perm = torch.arange(self.Lt, dtype=torch.int32).repeat(self.Cl, 1)
            n_c = cfg.n_ctx_text
            for i, e_ in enumerate(eot_all[cls_idx].tolist()):
                name_len = e_ - n_c - 2                                      # SOT, ctx, name tokens, ".", EOT
                assert name_len >= 0
                ctx_pos = list(range(1, 1 + n_c))
                cls_pos = list(range(1 + n_c, 1 + n_c + name_len))
                half = n_c // 2
                order = (ctx_pos[:half] + cls_pos + ctx_pos[half:]) if cfg.class_token_position == "middle" \
                    else (cls_pos + ctx_pos)                                 # coop.py:268-281 / :282-291
                perm[i, 1:1 + len(order)] = torch.tensor(order, dtype=torch.int32)
            self.text_perm = perm.to(dev)
        self.n_text_seq = self.Cl * (B if self.per_image else 1)
        if self.per_image:
            self.eot = self.eot.repeat(B).contiguous()                       # sequence s = image * C + class

        # ---- towers
        P = a.n_patches
        v_prompts = {i: (1 + P, cfg.n_ctx_vision) for i in range(1, cfg.depth_vision)} if need_vb else {}
        t_prompts = {i: (1, cfg.n_ctx_text) for i in range(1, cfg.depth_text)}
        self.vt = Tower(self.vw, B, self.Lv, self.Dv, a.vision_heads, False, need_vb, dev, v_prompts)
        self.tt = Tower(self.tw, self.n_text_seq, self.Lt, self.Dt, a.transformer_heads, True, True, dev,
                        t_prompts)
        self.zt = Tower(self.vw, B, 1 + P, self.Dv, a.vision_heads, False, False, dev) \
            if cfg.zero_shot_branch else None
        if self.zt is not None:
            self.zt.cls_only_last_mlp = True    # only ln_post(x[:, 0]) of the frozen tower is ever read (:301)
        # only ln_post(x[:, 0]) is read (model.py:426), with or without a backward (PSRC_CLS_ONLY=0: A/B switch)
        self.vt.cls_only_last_mlp = need_vb is False or os.environ.get("PSRC_CLS_ONLY", "1") != "0"

        e = lambda *s, dt=torch.float32: torch.empty(*s, device=dev, dtype=dt)  # noqa: E731
        self.patches = e(B * P, 3 * a.vision_patch_size ** 2, dt=self.fwd)
        self.pe = e(B * P, self.Dv, dt=self.fwd)
        Ts = self.n_text_seq
        self.hv, self.ht = e(B, self.Dv, dt=self.fwd), e(Ts, self.Dt, dt=self.fwd)
        self.hv_stats, self.ht_stats = (e(B), e(B)), (e(Ts), e(Ts))
        self.f, self.f_hat, self.f_inv = e(B, self.E), e(B, self.E), e(B)
        self.t, self.t_hat_l, self.t_inv = e(Ts, self.E), e(Ts, self.E), e(Ts)
        self.t_hat = e(self.C_pad, self.E) if self.share_text else self.t_hat_l
        self.logits, self.d_logits = e(B, C), e(B, C)
        self.d_fhat, self.d_that_l = e(B, self.E), e(Ts, self.E)
        # zero-initialised: the rows of padding classes are never written and must stay zero
        self.d_that_full = torch.zeros(self.C_pad, self.E, device=dev) if self.share_text else None
        self.d_that_l1 = torch.zeros(self.Cl, self.E, device=dev) if self.share_text else None
        self.df, self.dt_ = e(B, self.E), e(Ts, self.E)
        self.df_a = e(B, self.E, dt=self.act) if self.act != torch.float32 else self.df
        self.dt_a = e(Ts, self.E, dt=self.act) if self.act != torch.float32 else self.dt_
        self.dhv, self.dht = e(B, self.Dv, dt=self.act), e(Ts, self.Dt, dt=self.act)
        if self.text_perm is not None:
            self.x0_end, self.dx0_end = e(self.Cl * self.Lt, self.Dt), e(self.Cl * self.Lt, self.Dt)
            self.zero_pos = torch.zeros_like(self.tpos)
        self.losses_host = torch.zeros(4, dtype=torch.float32).pin_memory() if dev.type == "cuda" else torch.zeros(4)
        self._text_ahead = False     # True: the text-tower prefix of the NEXT step is already enqueued (graphs only)
        self.loss_scratch = e(4 * B + 2 * 296)
        self.labels = torch.zeros(B, device=dev, dtype=torch.int32)
        self.image = e(B, 3, a.image_resolution, a.image_resolution)
        self.focal_alpha = torch.tensor(cfg.focal_alpha, device=dev, dtype=torch.float32) \
            if cfg.focal_alpha is not None else None
        if need_vb:
            nr = B * cfg.n_ctx_vision           # the shallow prompt rows of every image (layer-0 VPT gradient)
            self.vpt_rounded = e(nr, self.Dv)
            self.vpt_scratch = e(nr, self.Dv, dt=self.fwd)
            self.vpt_stats = (e(nr), e(nr))
            self.vpt_dy, self.vpt_dx = e(nr, self.Dv), e(nr, self.Dv)

        # ---- frozen teacher (PromptSRC): ZS image features per step, fixed text embeddings once
        if cfg.zero_shot_branch:
            self.zhv, self.zf = e(B, self.Dv, dt=self.fwd), e(B, self.E)
            self.zf_hat, self.zs_logits = e(B, self.E), e(B, C)
            assert fixed_embeddings is not None, "PromptSRC needs the frozen-CLIP text embeddings"
            fe = _dev(fixed_embeddings, dev)                                   # [C, E] un-normalised
            self.fixed_hat = e(C, self.E)
            ops.l2norm_fwd(fe, self.fixed_hat, None)                          # promptsrc.py:196-197
            self.fixed_hat_h = self.fixed_hat.to(torch.float16).to(torch.float32)  # `.half()` at :201 (init-time)
            self.fixed_hat_l = torch.zeros(self.Cl, self.E, device=dev)
            self.fixed_hat_l[:self.Cv].copy_(self.fixed_hat[self.c0:self.c0 + self.Cv])

        # ---- trainable prompt parameters live in ONE flat fp32 buffer (+ flat grad buffer)
        self.param_shapes = {}
        if cfg.ctx_per_class:
            self.param_shapes["ctx"] = (C, cfg.n_ctx_text, self.Dt)
        else:
            self.param_shapes["ctx"] = (cfg.n_ctx_text, self.Dt)
        for i in range(1, cfg.depth_text):
            self.param_shapes[f"text.{i}"] = (cfg.n_ctx_text, self.Dt)
        nd = cfg.depth_vision
        if cfg.coupling:
            # MaPLe: the vision prompts are DERIVED (maple.py:199-204); trainable are the coupling Linears.
            # Index 0 = `proj` (ctx -> first-layer prompt), i >= 1 = compound_prompt_projections[i-1].
            for i in range(nd):
                self.param_shapes[f"couple.w.{i}"] = (self.Dv, self.Dt)
            for i in range(nd):
                self.param_shapes[f"couple.b.{i}"] = (self.Dv,)
        elif need_vb:
            self.param_shapes["vpt"] = (cfg.n_ctx_vision, self.Dv)
            for i in range(1, cfg.depth_vision):
                self.param_shapes[f"vision.{i}"] = (cfg.n_ctx_vision, self.Dv)
        if self.per_image:
            Hm = cfg.meta_hidden
            assert Hm > 0, "CoCoOp needs the meta-net hidden width"
            self.param_shapes.update({"meta.w1": (Hm, self.E), "meta.b1": (Hm,), "meta.w2": (self.Dt, Hm),
                                      "meta.b2": (self.Dt,)})
        n = sum(math.prod(s) for s in self.param_shapes.values())
        # gradients and the 4 loss terms live in ONE buffer: multi-GPU runs sum both with a single all-reduce
        self.flat_grad_ext = torch.zeros(n + 4, device=dev)
        self.flat_param, self.flat_grad = torch.zeros(n, device=dev), self.flat_grad_ext[:n]
        self.losses = self.flat_grad_ext[n:]
        self.params, self.grads, self.param_offsets, off = {}, {}, {}, 0
        for k, s in self.param_shapes.items():
            m = math.prod(s)
            self.params[k] = self.flat_param[off:off + m].view(*s)
            self.grads[k] = self.flat_grad[off:off + m].view(*s)
            self.param_offsets[k] = off
            off += m
        # vision prompt values / gradients by layer (0 = the shallow prompt appended before ln_pre)
        if cfg.coupling:
            n_c = cfg.n_ctx_text

            def span(buf, first, count, *shape):
                o = self.param_offsets[first]
                return buf[o:o + count].view(*shape)
            nt = nd * n_c * self.Dt              # ctx, text.1 .. text.{nd-1} are contiguous: X of the batched Linear
            self.couple_x, self.couple_dx = (span(b_, "ctx", nt, nd, n_c, self.Dt) for b_ in (self.flat_param, self.flat_grad))
            nw = nd * self.Dv * self.Dt
            self.couple_w, self.couple_dw = (span(b_, "couple.w.0", nw, nd, self.Dv, self.Dt)
                                             for b_ in (self.flat_param, self.flat_grad))
            self.couple_b, self.couple_db = (span(b_, "couple.b.0", nd * self.Dv, nd, self.Dv)
                                             for b_ in (self.flat_param, self.flat_grad))
            self.vis_prompt, self.vis_prompt_grad = e(nd, n_c, self.Dv), e(nd, n_c, self.Dv)
            self.vprompts = {i: self.vis_prompt[i] for i in range(nd)}
            self.vgrads = {i: self.vis_prompt_grad[i] for i in range(nd)}
        elif need_vb:
            self.vprompts = {0: self.params["vpt"], **{i: self.params[f"vision.{i}"] for i in range(1, nd)}}
            self.vgrads = {0: self.grads["vpt"], **{i: self.grads[f"vision.{i}"] for i in range(1, nd)}}
        else:
            self.vprompts, self.vgrads = {}, {}
        if self.per_image:
            Hm = cfg.meta_hidden
            self.meta_h, self.meta_dh = e(1, B, Hm), e(1, B, Hm)
            self.ctx_bias, self.d_ctx_bias = e(1, B, self.Dt), e(1, B, self.Dt)
            self.ctx_gg = e(B, cfg.n_ctx_text, self.Dt)

    # ------------------------------------------------------------------ pieces
    def _vision_embed(self):
        a = self.cfg.arch
        ops.im2col_patches(self.image, self.patches, a.image_resolution, a.vision_patch_size)
        ops.gemm(self.patches, self.conv_w, self.pe)                         # conv1 (model.py:402)

    def _vision_head(self, tower, hv, stats, f_out):
        ops.gather_ln_fwd(tower.x_final, None, *self.ln_post, hv, *stats, tower.n_seq, tower.L, self.Dv)  # :426
        ops.gemm(hv, self.vproj_t, f_out)                                    # :428-429

    def text_features(self, ctx=None, gather=True):
        """Normalised text features of this rank's class shard -> self.t_hat_l (and, gathered, self.t_hat)."""
        cfg = self.cfg
        ctx = self.params["ctx"] if ctx is None else ctx
        if cfg.ctx_per_class:
            ctx = ctx[self.c0:self.c1]
        if self.per_image:
            # meta-net on the normalised image features, one context shift per image (cocoop.py:139-143, :185-188)
            p = self.params
            ops.small_linear_fwd(self.f_hat.view(1, cfg.batch, self.E), p["meta.w1"].unsqueeze(0),
                                 p["meta.b1"].unsqueeze(0), self.meta_h, relu=True)
            ops.small_linear_fwd(self.meta_h, p["meta.w2"].unsqueeze(0), p["meta.b2"].unsqueeze(0), self.ctx_bias)
            ops.assemble_text_prompts(self.token_prefix, self.token_suffix, ctx, self.tpos, self.tt.x0, self.Cl,
                                      self.Lt, cfg.n_ctx_text, self.Dt, ctx_bias=self.ctx_bias.view(cfg.batch, self.Dt),
                                      n_groups=cfg.batch)
        elif self.text_perm is not None:
            ops.assemble_text_prompts(self.token_prefix, self.token_suffix, ctx, self.zero_pos, self.x0_end, self.Cl,
                                      self.Lt, cfg.n_ctx_text, self.Dt, ctx_per_class=cfg.ctx_per_class)
            ops.permute_rows(self.x0_end, self.text_perm, self.tt.x0, self.Cl, self.Lt, self.Dt, add=self.tpos)
        else:
            ops.assemble_text_prompts(self.token_prefix, self.token_suffix, ctx, self.tpos, self.tt.x0, self.Cl,
                                      self.Lt, cfg.n_ctx_text, self.Dt, ctx_per_class=cfg.ctx_per_class)
        self.tt.forward({i: self.params[f"text.{i}"] for i in range(1, cfg.depth_text)})
        ops.gather_ln_fwd(self.tt.x_final, self.eot, *self.ln_final, self.ht, *self.ht_stats, self.n_text_seq,
                          self.Lt, self.Dt)                                  # promptsrc.py:64-68
        ops.gemm(self.ht, self.tproj_t, self.t)
        ops.l2norm_fwd(self.t, self.t_hat_l, self.t_inv)                     # promptsrc.py:191
        if self.share_text and gather:
            dist.all_gather_into_tensor(self.t_hat, self.t_hat_l, group=cfg.process_group)
        return self.t_hat

    def couple_prompts(self):
        """MaPLe: vision prompt of layer i = Linear_i(text prompt of layer i) (maple.py:199-204), one launch."""
        if self.cfg.coupling:
            ops.small_linear_fwd(self.couple_x, self.couple_w, self.couple_b, self.vis_prompt)

    def image_features(self, embed=True):
        cfg = self.cfg
        if embed:
            self._vision_embed()
        self.couple_prompts()
        vpt = self.vprompts[0] if cfg.n_ctx_vision > 0 else None
        ops.assemble_vision_tokens(self.pe, self.cls, self.vpos, vpt, *self.ln_pre, self.vt.x0, cfg.batch,
                                   cfg.arch.n_patches, cfg.n_ctx_vision, self.Dv)
        self.vt.forward(self.vprompts if cfg.n_ctx_vision > 0 else None)
        self._vision_head(self.vt, self.hv, self.hv_stats, self.f)
        ops.l2norm_fwd(self.f, self.f_hat, self.f_inv)                       # promptsrc.py:190
        return self.f_hat

    def zero_shot_features(self):
        """Frozen CLIP image tower on the same patch embeddings (promptsrc.py:198-200)."""
        cfg = self.cfg
        ops.assemble_vision_tokens(self.pe, self.cls, self.vpos, None, *self.ln_pre, self.zt.x0, cfg.batch,
                                   cfg.arch.n_patches, 0, self.Dv)
        self.zt.forward()
        self._vision_head(self.zt, self.zhv, (None, None), self.zf)
        ops.l2norm_fwd(self.zf, self.zf_hat, None)
        return self.zf_hat

    def compute_logits(self):
        B, C, E = self.cfg.batch, self.cfg.n_cls, self.E
        if self.per_image:
            ops.group_logits_fwd(self.f_hat, self.t_hat_l, self.logits, self.scale)   # cocoop.py:249-252
        else:
            ops.gemm_f32_strided(B, C, E, self.f_hat, E, 1, self.t_hat, E, 1, self.logits, alpha=self.scale)  # :193
        return self.logits

    # ------------------------------------------------------------------ full step
    def forward_text(self):
        """The image-independent part of the forward: the text tower over this rank's class prompts (+ all-gather).
        Runs while the batch is still on its way to the device (`train_step_host`).  CoCoOp's text pass depends on
        the image features and therefore belongs to `forward_rest`."""
        self._text_ahead = False     # an eager pass: whatever was enqueued ahead no longer counts
        if not self.per_image:
            self.text_features()

    def forward_rest(self):
        self._forward_vision()
        return self._forward_logits()

    def _forward_logits(self):
        if self.per_image:                 # the context shift needs the image features first (cocoop.py:241-246)
            self.text_features()
        return self.compute_logits()

    def _forward_vision(self):
        """Everything of the forward that needs the images but not the (gathered) text features: both vision towers,
        image features and -- PromptSRC -- the teacher logits."""
        cfg = self.cfg
        B, C, E = cfg.batch, cfg.n_cls, self.E
        side = self._side_stream() if (cfg.zero_shot_branch and self.overlap_zero_shot) else None
        if side is None:
            self.image_features()
        else:
            # the frozen (zero-shot) tower only shares the patch embeddings with the prompted tower: run it on a second
            # stream so its LayerNorm / attention kernels fill the gaps of the other tower's GEMMs and vice versa
            # (fork / join through events, so the pair is capturable into the step graph)
            main = torch.cuda.current_stream()
            self._vision_embed()
            side.wait_stream(main)
            with torch.cuda.stream(side):
                self.zero_shot_features()
                ops.gemm_f32_strided(B, C, E, self.zf_hat, E, 1, self.fixed_hat_h, E, 1, self.zs_logits,
                                     alpha=self.scale)                      # promptsrc.py:201
            self.image_features(embed=False)
        if side is not None:
            torch.cuda.current_stream().wait_stream(side)
        elif cfg.zero_shot_branch:
            self.zero_shot_features()
            ops.gemm_f32_strided(B, C, E, self.zf_hat, E, 1, self.fixed_hat_h, E, 1, self.zs_logits,
                                 alpha=self.scale)                          # promptsrc.py:201

    overlap_zero_shot = True     # measured on B200 at cfg 3: 24.03 -> 23.47 ms/step

    def _side_stream(self):
        if self.device.type != "cuda":
            return None
        if getattr(self, "_side", None) is None:
            self._side = torch.cuda.Stream(device=self.device)
        return self._side

    def forward_train(self):
        """Forward of CustomCLIP.forward in training mode (promptsrc.py:183-201; coop.py:351-365): text features,
        image features, logits and -- PromptSRC -- the frozen-teacher features / logits.  No loss, no backward."""
        self.forward_text()
        return self.forward_rest()

    def forward_backward(self):
        """Inputs: self.image / self.labels (device).  Fills self.losses, self.logits and self.flat_grad
        (all-reduced across ranks).  The optimizer step is the caller's (trainer / bench)."""
        self.forward_text()
        return self.backward_after_text()

    def backward_after_text(self):
        """Everything of the step that needs the image batch: vision towers, logits, losses, prompt-only backward."""
        self.forward_rest()
        self._loss()
        self.backward_from_buffers(feature_grads=self.cfg.zero_shot_branch, reduce=False)
        if self.cfg.world_size > 1:        # prompt gradients and the 4 loss terms in one message
            dist.all_reduce(self.flat_grad_ext, group=self.cfg.process_group)
        return self.losses

    def _loss(self):
        cfg = self.cfg
        zs = cfg.zero_shot_branch
        d_that_l1 = self.d_that_l1 if self.share_text else self.d_that_l
        ops.loss_fwd_bwd(self.logits, self.labels, self.d_logits, self.losses, self.loss_scratch,
                         zs_logits=self.zs_logits if zs else None, focal_alpha=self.focal_alpha,
                         focal_gamma=cfg.focal_gamma,
                         t_hat=self.t_hat_l[:self.Cv] if (zs and self.Cv) else None,
                         t_zs=self.fixed_hat_l[:self.Cv] if (zs and self.Cv) else None,
                         d_that=d_that_l1[:self.Cv] if (zs and self.Cv) else None,
                         f_hat=self.f_hat if zs else None, f_zs=self.zf_hat if zs else None,
                         d_fhat=self.d_fhat if zs else None,
                         w_kl=cfg.w_logits, w_text=cfg.w_text, w_image=cfg.w_image,
                         B_global=self.B_global, Ct_global=cfg.n_cls)

    # ------------------------------------------------------------------ multi-GPU: segments between the collectives
    # world_size > 1: the step is cut where a collective sits, each segment is a CUDA graph (no NCCL call
    # inside a graph), and the collectives are issued asynchronously so that they run under independent compute:
    #   seg 0  text tower forward (own class shard)            -> all-gather t_hat        | runs under seg 1
    #   seg 1  vision towers forward (+ frozen tower)            (needs the images, not t_hat)
    #   seg 2  logits, fused loss, dL/dt_hat partial sums      -> reduce-scatter          | runs under seg 3
    #   seg 3  vision backward                                   (needs d_logits, t_hat -- not the scattered sum)
    #   seg 4  text backward (+ MaPLe coupling backward)       -> ONE all-reduce: flat prompt gradient + 4 loss terms
    #   seg 5  SGD
    # CoCoOp (pure batch sharding): seg 0 = the whole forward + backward, all-reduce, seg 1 = SGD.
    def _segments(self):
        zs = self.cfg.zero_shot_branch
        if not self.share_text:
            def whole():
                self.forward_text()
                self.forward_rest()
                self._loss()
                self.backward_from_buffers(feature_grads=zs, reduce=False)
            return [whole, self.sgd]

        def seg_text_fwd():
            self._text_ahead = False
            self.text_features(gather=False)

        def seg_logits_loss():
            self._forward_logits()
            self._loss()
            self._backward_text_pre()

        def seg_text_bwd():
            self._backward_text_post(zs)
            self._backward_coupling()
        return [seg_text_fwd, self._forward_vision, seg_logits_loss, lambda: self._backward_vision(zs), seg_text_bwd,
                self.sgd]

    def run_segments(self, graphs=None, wait_event=None):
        """One train step at world_size > 1.  graphs: the captured segments (capture()), None = eager launches."""
        pg = self.cfg.process_group
        segs = self._segments()
        run = (lambda i: graphs[i].replay()) if graphs is not None else (lambda i: segs[i]())
        main = torch.cuda.current_stream() if self.device.type == "cuda" else None
        if not self.share_text:
            if wait_event is not None:
                main.wait_event(wait_event)
            run(0)
            dist.all_reduce(self.flat_grad_ext, group=pg)
            run(1)
            return self.losses
        run(0)
        w_ag = dist.all_gather_into_tensor(self.t_hat, self.t_hat_l, group=pg, async_op=True)
        if wait_event is not None:
            main.wait_event(wait_event)         # the batch's host-to-device copy (run_host)
        run(1)
        w_ag.wait()
        run(2)
        w_rs = dist.reduce_scatter_tensor(self.d_that_l, self.d_that_full, group=pg, async_op=True)
        run(3)
        w_rs.wait()
        run(4)
        dist.all_reduce(self.flat_grad_ext, group=pg)
        run(5)
        return self.losses

    def backward_from_buffers(self, feature_grads: bool, reduce: bool = True):
        """Prompt-only backward from d(loss)/d(logits) in self.d_logits.  feature_grads: the direct gradients
        w.r.t. the normalised features (PromptSRC's L1 terms, or an autograd caller's) are already in
        self.d_fhat and self.d_that_l (single GPU) / self.d_that_l1 (sharded) and are added to the logits part.
        Fills self.flat_grad (all-reduced across ranks)."""
        cfg = self.cfg
        multi = cfg.world_size > 1
        # the two towers' backward chains are independent: the text chain (small-M GEMMs, 58 % wave occupancy) runs
        # on a second stream when the vision tower has a backward of its own, joined before the coupling / all-reduce
        side = self._side_stream() if (self.overlap_backward and cfg.n_ctx_vision > 0 and not multi) else None
        if side is None:
            self._backward_text(feature_grads)
            self._backward_vision(feature_grads)
        else:
            main = torch.cuda.current_stream()
            side.wait_stream(main)
            with torch.cuda.stream(side):
                self._backward_text(feature_grads)
            self._backward_vision(feature_grads)
            main.wait_stream(side)
        self._backward_coupling()
        if multi and reduce:
            dist.all_reduce(self.flat_grad, group=cfg.process_group)
        return self.flat_grad

    def _backward_coupling(self):
        if self.cfg.coupling:
            # d(text prompts) += d(vision prompts) . W ; dW, db of the coupling Linears (autograd of maple.py:199-204)
            ops.small_linear_bwd(self.couple_x, self.couple_w, self.vis_prompt_grad, dx=self.couple_dx,
                                 dw=self.couple_dw, db=self.couple_db, accumulate_dx=True)

    overlap_backward = False     # measured: no gain (22.98 vs 23.09 ms), kept as a switch

    def _backward_text(self, zs: bool):
        self._backward_text_pre(zs)
        if self.share_text:
            dist.reduce_scatter_tensor(self.d_that_l, self.d_that_full, group=self.cfg.process_group)
        self._backward_text_post(zs)

    def _backward_text_pre(self, zs: bool = False):
        """dL/dt_hat = s * dlogits^T f_hat (+ direct part): with a sharded text tower the partial sum over this
        rank's images for ALL classes (input of the reduce-scatter), else the complete gradient."""
        cfg = self.cfg
        B, C, E = cfg.batch, cfg.n_cls, self.E
        if self.per_image:
            assert not zs
            ops.group_logits_bwd(self.f_hat, self.d_logits, self.d_that_l, self.scale)
        elif self.share_text:
            ops.gemm_f32_strided(C, E, B, self.d_logits, 1, C, self.f_hat, 1, E, self.d_that_full[:C], alpha=self.scale)
        else:
            ops.gemm_f32_strided(C, E, B, self.d_logits, 1, C, self.f_hat, 1, E, self.d_that_l, alpha=self.scale,
                                 beta=1.0 if zs else 0.0)

    def _backward_text_post(self, zs: bool):
        cfg = self.cfg
        B = cfg.batch
        if self.share_text and zs:
            ops.axpy(self.d_that_l, self.d_that_l1, 1.0)       # + the L1-text gradient of the own shard
        ops.l2norm_bwd(self.d_that_l, self.t_hat_l, self.t_inv, self.dt_)
        if self.dt_a is not self.dt_:
            ops.cast(self.dt_, self.dt_a)
        ops.gemm(self.dt_a, self.tproj, self.dht)                            # d(ln_final row) = dt . proj^T
        ops.gather_ln_bwd(self.dht, self.tt.x_final, self.eot, self.ln_final[0], *self.ht_stats, self.tt.dx,
                          self.n_text_seq, self.Lt, self.Dt, dx_act=self.tt.dxa)
        self.tt.backward({i: self.grads[f"text.{i}"] for i in range(1, cfg.depth_text)})
        if self.per_image:
            # context rows summed over the classes of each image, then split into d ctx / d bias and pushed
            # through the meta-net (autograd of cocoop.py:185-188, :139-143); the image tower is frozen
            n_c, rows = cfg.n_ctx_text, self.Cl * self.Lt
            for g in range(B):
                ops.prompt_grad(self.tt.dx[g * rows:(g + 1) * rows], self.ctx_gg[g], self.Cl, self.Lt, 1, n_c,
                                self.Dt, zero_rows=False, round_fp16=False)
            ops.ctx_bias_grad(self.ctx_gg, self.grads["ctx"], self.d_ctx_bias.view(B, self.Dt))
            p, g_ = self.params, self.grads
            ops.small_linear_bwd(self.meta_h, p["meta.w2"].unsqueeze(0), self.d_ctx_bias, dx=self.meta_dh,
                                 dw=g_["meta.w2"].unsqueeze(0), db=g_["meta.b2"].unsqueeze(0))
            ops.small_linear_bwd(self.f_hat.view(1, B, self.E), p["meta.w1"].unsqueeze(0), self.meta_dh,
                                 dw=g_["meta.w1"].unsqueeze(0), db=g_["meta.b1"].unsqueeze(0), y_relu=self.meta_h)
        dx0 = self.tt.dx
        if self.text_perm is not None:          # route d(x0) back to the "end" layout the context rows live in
            dx0 = ops.permute_rows(self.tt.dx, self.text_perm, self.dx0_end, self.Cl, self.Lt, self.Dt, inverse=True)
        if self.per_image:
            pass
        elif cfg.ctx_per_class:
            g = self.grads["ctx"]
            g.zero_()
            g[self.c0:self.c0 + self.Cv].copy_(dx0.view(self.Cl, self.Lt, self.Dt)[:self.Cv, 1:1 + cfg.n_ctx_text])
        else:
            ops.prompt_grad(dx0, self.grads["ctx"], self.Cl, self.Lt, 1, cfg.n_ctx_text, self.Dt,
                            zero_rows=False, round_fp16=False)               # ctx has no .half() (promptsrc.py:160)

    def _backward_vision(self, zs: bool):
        cfg = self.cfg
        B, C, E = cfg.batch, cfg.n_cls, self.E
        # ---- vision side (only when vision prompts exist)
        if cfg.n_ctx_vision > 0:
            ops.gemm_f32_strided(B, E, C, self.d_logits, C, 1, self.t_hat, 1, E, self.d_fhat, alpha=self.scale,
                                 beta=1.0 if zs else 0.0)
            ops.l2norm_bwd(self.d_fhat, self.f_hat, self.f_inv, self.df)
            if self.df_a is not self.df:
                ops.cast(self.df, self.df_a)
            ops.gemm(self.df_a, self.vproj, self.dhv)
            ops.gather_ln_bwd(self.dhv, self.vt.x_final, None, self.ln_post[0], *self.hv_stats, self.vt.dx, B,
                              self.Lv, self.Dv, dx_act=self.vt.dxa)
            self.vt.backward(self.vgrads)
            # VPT enters before ln_pre as `VPT.expand(B, -1, -1).half()` (model.py:414-420): autograd runs the ln_pre
            # backward on every image's prompt rows, casts each image's gradient through fp16 (the backward of the
            # type promotion / `.half()`), and only then sums over the images.  Same order here: the prompt rows of
            # d(ln_pre output) are gathered, one LayerNorm backward over the B * n_ctx rows (the rounded prompt and its
            # statistics are the same for every image), then the per-sequence fp16 rounding + sum of psrc_prompt_grad.
            n, P = cfg.n_ctx_vision, cfg.arch.n_patches
            ops.insert_prompt(self.vpt_rounded, self.vprompts[0], B, n, 0, n, self.Dv, round_fp16=True)
            ops.layernorm_fwd(self.vpt_rounded, *self.ln_pre, self.vpt_scratch, *self.vpt_stats)
            self.vpt_dy.view(B, n, self.Dv).copy_(self.vt.dx.view(B, self.Lv, self.Dv)[:, 1 + P:])
            ops.layernorm_bwd(self.vpt_dy, self.vpt_rounded, self.ln_pre[0], *self.vpt_stats, self.vpt_dx)
            ops.prompt_grad(self.vpt_dx, self.vgrads[0], B, n, 0, n, self.Dv, zero_rows=False, round_fp16=True)

    def backward_from(self, g_logits, g_that=None, g_fhat=None):
        """Autograd bridge (trainers.CustomCLIP): upstream gradients of an arbitrary loss w.r.t. logits [B, C],
        normalised text features [C, E] and normalised image features [B, E] (fp32 device tensors or None)
        -> prompt gradients in self.flat_grad.  Single-process form (the sharded step uses the fused loss)."""
        assert self.cfg.world_size == 1, "the autograd bridge is single-process; sharded runs use forward_backward()"
        if g_logits is None:
            self.d_logits.zero_()
        else:
            self.d_logits.copy_(g_logits)
        direct = g_that is not None or g_fhat is not None
        if direct:
            self.d_that_l.zero_() if g_that is None else self.d_that_l.copy_(g_that)
            self.d_fhat.zero_() if g_fhat is None else self.d_fhat.copy_(g_fhat)
        return self.backward_from_buffers(feature_grads=direct)

    @torch.no_grad()
    def inference_logits(self, recompute_text=True):
        """Eval branch of CustomCLIP.forward (promptsrc.py:187-193, 212-213): logits [B, C]."""
        if self.per_image:
            self.image_features()
            self.text_features()
            return self.compute_logits()
        if recompute_text:
            self.text_features()
        self.image_features()
        return self.compute_logits()

    @torch.no_grad()
    def per_image_logits_for(self, image: torch.Tensor) -> torch.Tensor:
        """CoCoOp evaluation (cocoop.py:235-260, eval branch): the text features depend on the image, so nothing
        can be cached; test images are run in chunks of the configured batch through the step's own buffers."""
        assert self.per_image
        Bc, outs = self.cfg.batch, []
        for lo in range(0, image.shape[0], Bc):
            chunk = image[lo:lo + Bc]
            b = chunk.shape[0]
            self.image[:b].copy_(chunk)
            if b < Bc:
                self.image[b:].zero_()
            outs.append(self.inference_logits()[:b].clone())
        return outs[0] if len(outs) == 1 else torch.cat(outs, 0)

    def total_loss(self):
        return self.losses.sum()

    # ------------------------------------------------------------------ optimizer + host-facing step
    def init_optimizer(self, lr=0.0025, momentum=0.9, weight_decay=5e-4):
        """torch.optim.SGD as Dassl builds it (dassl/optim/optimizer.py:105-113; PromptSRC yaml:
        lr 0.0025, momentum 0.9, wd 5e-4) over the flat prompt buffer.  lr lives in device memory so a
        captured graph follows the scheduler (`set_lr`)."""
        self.lr_dev = torch.tensor([lr], device=self.device, dtype=torch.float32)
        self.mom_buf = torch.zeros_like(self.flat_param)   # zero buffer == torch's first-step `buf = g` (dampening 0)
        self.sgd_hp = (float(momentum), 0.0, float(weight_decay))
        self.sgd_nesterov = False
        self.drop_graphs()

    def drop_graphs(self):
        """Captured graphs bake the SGD hyper-parameters in as kernel arguments: drop them when those change."""
        self._graph = None
        self._seg_graphs = None
        self._text_ahead = False

    def set_lr(self, lr: float):
        self.lr_dev.fill_(float(lr))

    def sgd(self):
        m, d, wd = self.sgd_hp
        ops.sgd_step(self.flat_param, self.flat_grad, self.mom_buf, self.lr_dev, m, d, wd, self.sgd_nesterov, False)

    def train_step(self):
        """forward + losses + prompt-only backward + SGD on the resident inputs (self.image / self.labels).
        world_size > 1: the overlapped segment order of `run_segments`, launched eagerly."""
        if self.cfg.world_size > 1:
            return self.run_segments(None)
        self.forward_backward()
        self.sgd()
        return self.losses

    def _step_part(self, part):
        """part 0: image-independent prefix (text tower forward); part 1: the rest of the train step."""
        if part == 0:
            self.forward_text()
        else:
            self.backward_after_text()
            self.sgd()

    def capture(self, warmup: int = 2):
        """Capture train_step into CUDA graphs (shapes are static; every kernel is stateless).  With world_size > 1
        every segment between two collectives becomes one graph and the NCCL calls stay outside (`run_segments`);
        every rank must call capture() at the same point."""
        if self.cfg.world_size > 1:
            return self._capture_segments(warmup)
        keep = (self.flat_param.clone(), self.mom_buf.clone())
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(warmup):
                self.train_step()
        torch.cuda.current_stream().wait_stream(s)
        # two graphs: the text-tower prefix does not read the image batch, so `train_step_host` replays it while the
        # host->device copy of the batch is still in flight on the copy stream
        g0, g1 = (None if self.per_image else torch.cuda.CUDAGraph()), torch.cuda.CUDAGraph()
        if g0 is not None:
            with torch.cuda.graph(g0):
                self._step_part(0)
        with torch.cuda.graph(g1):
            self._step_part(1)
        self.flat_param.copy_(keep[0])
        self.mom_buf.copy_(keep[1])
        self._graph = (g0, g1)
        self._text_ahead = False
        return self._graph

    def _capture_segments(self, warmup: int = 1):
        keep = (self.flat_param.clone(), self.mom_buf.clone())
        for _ in range(max(1, warmup)):                 # eager: initialises NCCL, fills every buffer with valid data
            self.run_segments(None)
        torch.cuda.synchronize()
        graphs, pool = [], None
        for fn in self._segments():                     # each segment replays the SAME launches on the same buffers
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, pool=pool):
                fn()
            pool = pool or g.pool()
            graphs.append(g)
        self.flat_param.copy_(keep[0])
        self.mom_buf.copy_(keep[1])
        self._seg_graphs = graphs
        self._graph = None
        self._text_ahead = False
        return graphs

    # The text-tower prefix reads only the prompt parameters, so as soon as one step's SGD is enqueued the NEXT step's
    # prefix can follow it on the stream: it then runs while the host is still waiting for this step's loss, logging
    # and fetching the next batch (Dassl's loop calls `.item()` every iteration, promptsrc.py:330), instead of after.
    # The loss terms are copied to pinned memory BEFORE that prefix is enqueued, so reading them does not wait for it.
    text_ahead = os.environ.get("PSRC_TEXT_AHEAD", "1") != "0"

    def invalidate_text(self):
        """Call after changing prompt parameters from outside the step (load_state_dict, a foreign optimizer)."""
        self._text_ahead = False

    def _replay(self, wait_event=None):
        g0, g1 = self._graph
        main = torch.cuda.current_stream()
        if g0 is not None and not self._text_ahead:
            g0.replay()
        if wait_event is not None:
            main.wait_event(wait_event)
        g1.replay()
        self.losses_host.copy_(self.losses, non_blocking=True)
        if getattr(self, "_loss_ready", None) is None:
            self._loss_ready, self._step_done = torch.cuda.Event(), torch.cuda.Event()
        self._loss_ready.record(main)
        self._step_done.record(main)
        self._text_ahead = False
        if g0 is not None and self.text_ahead and self.cfg.world_size == 1:
            g0.replay()
            self._text_ahead = True

    def loss_value(self) -> float:
        """Sum of the step's loss terms on the host (the trainers' `loss.item()`), without waiting for work that was
        enqueued after the step (the next step's text prefix)."""
        if getattr(self, "_graph", None) is not None and getattr(self, "_loss_ready", None) is not None:
            self._loss_ready.synchronize()
            return float(self.losses_host.sum())
        return float(self.losses.sum().item())

    def run(self):
        """One train step on the resident inputs: graph replay when captured, eager otherwise."""
        if self.cfg.world_size > 1:
            return self.run_segments(getattr(self, "_seg_graphs", None))
        if getattr(self, "_graph", None) is not None:
            self._replay()
        else:
            self.train_step()
        return self.losses

    def train_step_host(self, image_host: torch.Tensor, label_host: torch.Tensor, next_host=None) -> float:
        """The call the Dassl trainer makes per batch (PromptSRC.forward_backward, promptsrc.py:285-330):
        host batch in (pinned fp32 [B,3,R,R], int64 [B]) -> H2D -> step -> SGD -> `loss.item()` (D2H sync).
        `next_host = (image, label)` of the FOLLOWING batch, when the loader can look one batch ahead
        (trainers.LookaheadLoader): its H2D copy is started behind this step's launch and runs under its compute."""
        if getattr(self, "_staged_key", None) == id(image_host):
            self.run_staged()
        else:
            self.run_host(image_host, label_host)
        if next_host is not None:
            self.stage_host(*next_host)
        return self.loss_value()                                            # :330 (16-byte D2H)

    def train_step_host_u8(self, item: dict, label_host: torch.Tensor, next_host=None, canvas=None) -> float:
        """`train_step_host` fed by the on-device input pipeline (input_pipeline.DeviceTransform /
        DevicePreprocessor): `item` = collated {"u8" [B,Hc,Wc,3] uint8, "geom_f" [B,4], "geom_i" [B,5]} from pinned host
        memory (the crop boxes / flips were drawn on the host with the reference's RNG calls), label_host int64 [B].
        H2D of the uint8 windows (a third of the fp32 batch at 256 x 256 canvases) -> crop / bicubic resample / flip /
        normalise kernel into self.image -> step -> SGD -> loss.  `next_host = (item, labels)` starts the following
        batch's copies under this step's compute."""
        self.run_host_u8(item, label_host, next_host=next_host, canvas=canvas)
        return self.loss_value()

    def run_host_u8(self, item: dict, label_host: torch.Tensor, next_host=None, canvas=None):
        """`train_step_host_u8` without the final loss read (the trainers read it through `loss_value()` later).
        `self.input_kwargs` (mean / std / interpolation of cfg.INPUT) configures the preprocessor on first use."""
        pre = getattr(self, "_pre", None)
        if pre is None:
            from .input_pipeline import DevicePreprocessor
            pre = self._pre = DevicePreprocessor(self.device, self.cfg.batch, self.cfg.arch.image_resolution,
                                                 canvas=canvas or tuple(item["u8"].shape[1:3]),
                                                 **getattr(self, "input_kwargs", {}))
        pre.run(item, self.image, labels_out=self.labels, label_host=label_host)
        self.run()
        if next_host is not None:
            pre.stage(*next_host)
        return self.losses

    def stage_host(self, image_host: torch.Tensor, label_host: torch.Tensor):
        """Start the H2D copy of the NEXT batch into staging buffers on the copy stream and return at once."""
        cs = getattr(self, "_copy_stream", None)
        if cs is None:
            cs = self._copy_stream = torch.cuda.Stream(device=self.device)
            self._copy_done = torch.cuda.Event()
        if getattr(self, "image_next", None) is None:
            self.image_next, self.labels_next = torch.empty_like(self.image), torch.empty_like(self.labels)
            self._stage_done, self._stage_free = torch.cuda.Event(), None
        if self._stage_free is not None:
            cs.wait_event(self._stage_free)         # the previous batch has left the staging buffers
        with torch.cuda.stream(cs):
            self.image_next.copy_(image_host, non_blocking=True)
            self.labels_next.copy_(label_host.to(torch.int32), non_blocking=True)
            self._stage_done.record(cs)
        self._staged_key = id(image_host)

    def run_staged(self):
        """One train step on the batch `stage_host` brought over: 77 MB device-to-device (24 us at cfg 3), then the step."""
        main = torch.cuda.current_stream()
        main.wait_event(self._stage_done)
        self.image.copy_(self.image_next, non_blocking=True)
        self.labels.copy_(self.labels_next, non_blocking=True)
        if self._stage_free is None:
            self._stage_free = torch.cuda.Event()
        self._stage_free.record(main)
        self._staged_key = None
        return self.run()

    def run_host(self, image_host: torch.Tensor, label_host: torch.Tensor):
        """Batch in from the host (pinned: truly asynchronous) on a copy stream, overlapped with the image-independent
        text-tower forward on the compute stream; the rest of the step waits for the copy.  Losses stay on the device."""
        main = torch.cuda.current_stream()
        cs = getattr(self, "_copy_stream", None)
        if cs is None:
            cs = self._copy_stream = torch.cuda.Stream(device=self.device)
            self._copy_done = torch.cuda.Event()
        if self._text_ahead:
            cs.wait_event(self._step_done)          # not the whole stream: the prefix enqueued ahead may still be running
        else:
            cs.wait_stream(main)                    # the previous step has finished reading self.image / self.labels
        with torch.cuda.stream(cs):
            self.image.copy_(image_host, non_blocking=True)                 # parse_batch_train :384-387
            self.labels.copy_(label_host.to(torch.int32), non_blocking=True)
            self._copy_done.record(cs)
        if self.cfg.world_size > 1:
            self.run_segments(getattr(self, "_seg_graphs", None), wait_event=self._copy_done)
        elif getattr(self, "_graph", None) is not None:
            self._replay(wait_event=self._copy_done)
        else:
            self._step_part(0)
            main.wait_event(self._copy_done)
            self._step_part(1)
        return self.losses


class VisionInference:
    """Evaluation path with cached text features: `SimpleTrainer.test` (dassl/engine/trainer.py:
    445-486) calls `model_inference` once per test batch and the reference re-runs the text tower over every class
    each time (promptsrc.py:187-188).  Here the normalised text features are computed once per
    evaluation (`refresh_text`) and each batch costs one forward of the prompted vision tower plus the [B, C]
    cosine-logit GEMM.  Shares the frozen weights and the live prompt parameters of a `PromptedClipStep`; owns
    only its own activation buffers (the test batch size differs from the training one)."""

    def __init__(self, step: PromptedClipStep, batch: int):
        self.step, self.batch = step, batch
        cfg, a, dev, act = step.cfg, step.cfg.arch, step.device, step.fwd
        P = a.n_patches
        v_prompts = {i: (1 + P, cfg.n_ctx_vision) for i in range(1, cfg.depth_vision)} if cfg.n_ctx_vision > 0 else {}
        self.vt = Tower(step.vw, batch, step.Lv, step.Dv, a.vision_heads, False, False, dev, v_prompts)
        self.vt.cls_only_last_mlp = True        # evaluation reads ln_post(x[:, 0]) only
        e = lambda *s, dt=torch.float32: torch.empty(*s, device=dev, dtype=dt)  # noqa: E731
        self.image = e(batch, 3, a.image_resolution, a.image_resolution)
        self.patches = e(batch * P, 3 * a.vision_patch_size ** 2, dt=act)
        self.pe = e(batch * P, step.Dv, dt=act)
        self.hv = e(batch, step.Dv, dt=act)
        self.f, self.f_hat = e(batch, step.E), e(batch, step.E)
        self.logits = e(batch, cfg.n_cls)
        self.text_valid = False

    @torch.no_grad()
    def refresh_text(self):
        self.step.text_features()
        self.step.couple_prompts()          # MaPLe: the vision prompts follow the text prompts
        self.text_valid = True

    @torch.no_grad()
    def logits_for(self, image: torch.Tensor) -> torch.Tensor:
        """image: fp32 [b, 3, R, R] on the device, any b (chunked / zero-padded to the buffer batch).
        Returns a fresh [b, C] fp32 tensor of logits (promptsrc.py:193)."""
        st, cfg, a = self.step, self.step.cfg, self.step.cfg.arch
        if not self.text_valid:
            self.refresh_text()
        outs = []
        for lo in range(0, image.shape[0], self.batch):
            chunk = image[lo:lo + self.batch]
            b = chunk.shape[0]
            self.image[:b].copy_(chunk)
            if b < self.batch:
                self.image[b:].zero_()
            ops.im2col_patches(self.image, self.patches, a.image_resolution, a.vision_patch_size)
            ops.gemm(self.patches, st.conv_w, self.pe)
            vpt = st.vprompts[0] if cfg.n_ctx_vision > 0 else None
            ops.assemble_vision_tokens(self.pe, st.cls, st.vpos, vpt, *st.ln_pre, self.vt.x0, self.batch,
                                       a.n_patches, cfg.n_ctx_vision, st.Dv)
            self.vt.forward(st.vprompts if cfg.n_ctx_vision > 0 else None)
            ops.gather_ln_fwd(self.vt.x_final, None, *st.ln_post, self.hv, None, None, self.batch, self.vt.L, st.Dv)
            ops.gemm(self.hv, st.vproj_t, self.f)
            ops.l2norm_fwd(self.f, self.f_hat, None)
            ops.gemm_f32_strided(self.batch, cfg.n_cls, st.E, self.f_hat, st.E, 1, st.t_hat, st.E, 1, self.logits,
                                 alpha=st.scale)
            outs.append(self.logits[:b].clone())
        return outs[0] if len(outs) == 1 else torch.cat(outs, 0)
