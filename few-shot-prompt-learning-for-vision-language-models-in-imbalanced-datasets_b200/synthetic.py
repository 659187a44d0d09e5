"""Deterministic synthetic CLIP weights, prompts, class tokens and inputs.

There is no network for checkpoints or datasets (BASELINE.json: "random-init CLIP",
"synthetic ...-shaped data"), so every config is built from seeds.  The state-dict uses
the reference's key names (PromptSRC/clip/model.py:488-561, :662-705) so the same tensors
load into the reference model, the oracle and this package.  Init scales follow
CLIP.initialize_parameters (model.py:562-590) and the nn defaults of the vision tower;
LayerNorm affine parameters and biases are perturbed so that no term is trivially 0/1.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from pathlib import Path

import numpy as np
import torch

SOT, EOT = 49406, 49407  # clip/simple_tokenizer.py: <|startoftext|>, <|endoftext|>
DATA_DIR = Path(__file__).resolve().parent / "data"


@dataclass(frozen=True)
class ClipArch:
    embed_dim: int
    image_resolution: int
    vision_layers: int
    vision_width: int
    vision_patch_size: int
    context_length: int
    vocab_size: int
    transformer_width: int
    transformer_layers: int

    @property
    def vision_heads(self):  # model.py:519
        return self.vision_width // 64

    @property
    def transformer_heads(self):  # model.py:686
        return self.transformer_width // 64

    @property
    def n_patches(self):
        return (self.image_resolution // self.vision_patch_size) ** 2


ARCHS = {
    "ViT-B/16": ClipArch(512, 224, 12, 768, 16, 77, 49408, 512, 12),
    "ViT-L/14": ClipArch(768, 224, 24, 1024, 14, 77, 49408, 768, 12),
    # small shapes for fast CPU oracle tests (same code paths: 2 heads, 3 layers, 16 patches)
    "tiny": ClipArch(64, 64, 3, 128, 16, 77, 49408, 128, 3),
}


def arch_from_state_dict(sd) -> ClipArch:
    """Same inference as clip.model.build_model (model.py:662-687)."""
    vw = sd["visual.conv1.weight"].shape[0]
    vl = len([k for k in sd if k.startswith("visual.") and k.endswith(".attn.in_proj_weight")])
    ps = sd["visual.conv1.weight"].shape[-1]
    grid = round((sd["visual.positional_embedding"].shape[0] - 1) ** 0.5)
    tw = sd["ln_final.weight"].shape[0]
    tl = len(set(k.split(".")[2] for k in sd if k.startswith("transformer.resblocks")))
    return ClipArch(sd["text_projection"].shape[1], ps * grid, vl, vw, ps, sd["positional_embedding"].shape[0],
                    sd["token_embedding.weight"].shape[0], tw, tl)


def make_clip_state_dict(arch: ClipArch | str, seed: int = 0) -> dict:
    """Random-init CLIP state-dict (fp32, CPU) with the reference's key names."""
    if isinstance(arch, str):
        arch = ARCHS[arch]
    g = torch.Generator().manual_seed(seed)

    def n(*shape, std=1.0):
        return torch.randn(*shape, generator=g) * std

    def u(*shape, bound=1.0):
        return (torch.rand(*shape, generator=g) * 2 - 1) * bound

    sd = {}
    vw, tw = arch.vision_width, arch.transformer_width
    fan_conv = 3 * arch.vision_patch_size ** 2
    sd["visual.conv1.weight"] = u(vw, 3, arch.vision_patch_size, arch.vision_patch_size, bound=1 / math.sqrt(fan_conv))
    sd["visual.class_embedding"] = n(vw, std=vw ** -0.5)
    sd["visual.positional_embedding"] = n(arch.n_patches + 1, vw, std=vw ** -0.5)
    for nm in ("visual.ln_pre", "visual.ln_post", "ln_final"):
        d = tw if nm == "ln_final" else vw
        sd[nm + ".weight"] = 1 + n(d, std=0.05)
        sd[nm + ".bias"] = n(d, std=0.02)
    sd["visual.proj"] = n(vw, arch.embed_dim, std=vw ** -0.5)

    def blocks(prefix, width, layers, clip_text_init):
        proj_std = (width ** -0.5) * ((2 * layers) ** -0.5)
        attn_std = width ** -0.5
        fc_std = (2 * width) ** -0.5
        for i in range(layers):
            p = f"{prefix}.resblocks.{i}."
            if clip_text_init:  # model.py:580-587
                sd[p + "attn.in_proj_weight"] = n(3 * width, width, std=attn_std)
                sd[p + "attn.out_proj.weight"] = n(width, width, std=proj_std)
                sd[p + "mlp.c_fc.weight"] = n(4 * width, width, std=fc_std)
                sd[p + "mlp.c_proj.weight"] = n(width, 4 * width, std=proj_std)
            else:  # nn.MultiheadAttention xavier_uniform / nn.Linear kaiming_uniform(a=sqrt(5)) defaults
                sd[p + "attn.in_proj_weight"] = u(3 * width, width, bound=math.sqrt(6 / (4 * width)))
                sd[p + "attn.out_proj.weight"] = u(width, width, bound=1 / math.sqrt(width))
                sd[p + "mlp.c_fc.weight"] = u(4 * width, width, bound=1 / math.sqrt(width))
                sd[p + "mlp.c_proj.weight"] = u(width, 4 * width, bound=1 / math.sqrt(4 * width))
            sd[p + "attn.in_proj_bias"] = n(3 * width, std=0.02)
            sd[p + "attn.out_proj.bias"] = n(width, std=0.02)
            sd[p + "mlp.c_fc.bias"] = u(4 * width, bound=1 / math.sqrt(width))
            sd[p + "mlp.c_proj.bias"] = u(width, bound=1 / math.sqrt(4 * width))
            for ln in ("ln_1", "ln_2"):
                sd[p + ln + ".weight"] = 1 + n(width, std=0.05)
                sd[p + ln + ".bias"] = n(width, std=0.02)

    blocks("visual.transformer", vw, arch.vision_layers, False)
    blocks("transformer", tw, arch.transformer_layers, True)
    sd["token_embedding.weight"] = n(arch.vocab_size, tw, std=0.02)
    sd["positional_embedding"] = n(arch.context_length, tw, std=0.01)
    sd["text_projection"] = n(tw, arch.embed_dim, std=tw ** -0.5)
    sd["logit_scale"] = torch.tensor(math.log(1 / 0.07))
    return sd


def make_deep_prompts(arch: ClipArch, n_ctx_vision: int, n_ctx_text: int, depth_vision: int, depth_text: int,
                      seed: int = 100) -> dict:
    """IVLP/PromptSRC deep prompts, std 0.02 (model.py:215-217, :388-390).  Keys relative to CLIP."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    if depth_vision > 0:
        out["visual.VPT"] = torch.randn(n_ctx_vision, arch.vision_width, generator=g) * 0.02
        for i in range(1, depth_vision):
            out[f"visual.transformer.resblocks.{i}.VPT_shallow"] = torch.randn(n_ctx_vision, arch.vision_width, generator=g) * 0.02
    for i in range(1, depth_text):
        out[f"transformer.resblocks.{i}.VPT_shallow"] = torch.randn(n_ctx_text, arch.transformer_width, generator=g) * 0.02
    return out


def make_maple_params(arch: ClipArch, n_ctx: int, depth: int, seed: int = 200) -> dict:
    """MaPLe's trainable coupling (trainers/maple.py:144-156): `proj` and one Linear per deeper layer
    (nn.Linear default init scale), compound text prompts std 0.02.  Keys relative to `prompt_learner.`."""
    g = torch.Generator().manual_seed(seed)
    Dt, Dv = arch.transformer_width, arch.vision_width
    bound = 1 / math.sqrt(Dt)
    u = lambda *s: (torch.rand(*s, generator=g) * 2 - 1) * bound  # noqa: E731
    out = {"proj.weight": u(Dv, Dt), "proj.bias": u(Dv)}
    for i in range(depth - 1):
        out[f"compound_prompts_text.{i}"] = torch.randn(n_ctx, Dt, generator=g) * 0.02
        out[f"compound_prompt_projections.{i}.weight"] = u(Dv, Dt)
        out[f"compound_prompt_projections.{i}.bias"] = u(Dv)
    return out


def make_cocoop_params(arch: ClipArch, seed: int = 300) -> dict:
    """CoCoOp's meta-net (trainers/cocoop.py:139-143): Linear(vis_dim, vis_dim // 16) -> ReLU -> Linear(.., ctx_dim),
    nn.Linear default init scale.  Keys relative to `prompt_learner.`."""
    g = torch.Generator().manual_seed(seed)
    E, Dt = arch.embed_dim, arch.transformer_width
    Hm = E // 16
    u = lambda bound, *s: (torch.rand(*s, generator=g) * 2 - 1) * bound  # noqa: E731
    return {"meta_net.linear1.weight": u(1 / math.sqrt(E), Hm, E), "meta_net.linear1.bias": u(1 / math.sqrt(E), Hm),
            "meta_net.linear2.weight": u(1 / math.sqrt(Hm), Dt, Hm), "meta_net.linear2.bias": u(1 / math.sqrt(Hm), Dt)}


# ------------------------------------------------------------------ tokens
def load_token_fixture(name: str = "tokens_object1000") -> np.ndarray:
    """int32 [C,77] CLIP token ids of "a photo of a object {i}." produced by the reference tokenizer
    (clip/clip.py:185-221); generated by oracle/make_golden.py, committed under data/."""
    return np.load(DATA_DIR / f"{name}.npz")["tokens"].astype(np.int64)


def synthetic_tokens(n_cls: int, n_ctx_words: int = 4, name_lens=None, context_length: int = 77, seed: int = 7) -> np.ndarray:
    """Tokenizer-free stand-in with the same structure: SOT, prompt words, name tokens, '.', EOT, zeros.
    Only for shapes the fixture does not cover; ids are arbitrary but deterministic."""
    rng = np.random.RandomState(seed)
    toks = np.zeros((n_cls, context_length), dtype=np.int64)
    for c in range(n_cls):
        nl = 2 if name_lens is None else int(name_lens[c])
        body = [320, 1125, 539, 320][:n_ctx_words] + list(rng.randint(1000, 40000, size=nl)) + [269]
        seq = [SOT] + body + [EOT]
        toks[c, :len(seq)] = seq
    return toks


# ------------------------------------------------------------------ imbalanced class counts
def lt_counts_cifar100(n_cls: int = 100, n_max: int = 500, ratio: float = 0.01):
    """CIFAR-100-LT-shaped exponential profile, imbalance ratio 100 (SURVEY 8d cfg 2)."""
    return [int(round(n_max * ratio ** (c / max(1, n_cls - 1)))) for c in range(n_cls)]


def lt_counts_pareto(n_cls: int = 1000, alpha: float = 6.0, lo: int = 5, hi: int = 1280, seed: int = 3):
    """ImageNet-LT-shaped Pareto profile clipped to [5, 1280] (SURVEY 8d cfg 3)."""
    rng = np.random.RandomState(seed)
    raw = (rng.pareto(alpha, size=n_cls) + 1) * lo * 4
    return sorted((int(min(hi, max(lo, round(v)))) for v in raw), reverse=True)


def focal_alpha(per_class_shots, n_cls: int):
    """alpha_c = sum(n) / (C * n_c), 0 when n_c == 0 (trainers/coop.py:326-346)."""
    total = float(sum(per_class_shots))
    return [total / (n_cls * c) if c > 0 else 0.0 for c in per_class_shots]


def synthetic_batch(batch: int, n_cls: int, resolution: int = 224, seed: int = 1):
    g = torch.Generator().manual_seed(seed)
    img = torch.randn(batch, 3, resolution, resolution, generator=g)
    lab = torch.randint(0, n_cls, (batch,), generator=g)
    return img, lab
