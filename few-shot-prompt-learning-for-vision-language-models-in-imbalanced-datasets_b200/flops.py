"""Algorithmic FLOP accounting of the prompted-CLIP step (SURVEY.md 8d formulas).

Per block forward, N sequences of length L, width D:  F = N*L*(24*D^2 + 4*L*D)
dgrad-only backward (weights frozen):                  B = N*L*(24*D^2 + 8*L*D)
Host-side arithmetic only -- used by bench.py and DESIGN.md, never on the compute path.
"""
from __future__ import annotations


def block_fwd(n, L, D):
    return n * L * (24 * D * D + 4 * L * D)


def block_bwd(n, L, D):
    return n * L * (24 * D * D + 8 * L * D)


def last_block_saving(n, L, D, attn_coeff):
    """FLOPs NOT executed in a tower's last block when only row 0 of every sequence is consumed."""
    return n * (L - 1) * (18 * D * D + attn_coeff * L * D)


def step_flops(arch, B, C, Lt, n_ctx_vision, zero_shot, vision_backward, text_backward=True):
    """Executed FLOPs of one train step on one rank (B local images, C local classes, Lt text positions)."""
    Dv, Dt, E, P = arch.vision_width, arch.transformer_width, arch.embed_dim, arch.n_patches
    Lv = 1 + P + n_ctx_vision
    out = {}
    out["patch_embed"] = 2 * B * P * (3 * arch.vision_patch_size ** 2) * Dv
    # The LAST block of every vision tower runs LN1 + the q/k/v projection on all rows, the attention core for ONE
    # query per (sequence, head) and the out-projection + MLP on the class-token rows only (engine.Tower
    # .cls_only_last_mlp): per skipped row 18 D^2 of GEMM and 4 L D (forward) / 8 L D (backward) of attention less.
    out["vision_fwd"] = arch.vision_layers * block_fwd(B, Lv, Dv) - last_block_saving(B, Lv, Dv, 4)
    out["vision_bwd"] = (arch.vision_layers * block_bwd(B, Lv, Dv) - last_block_saving(B, Lv, Dv, 8)) \
        if vision_backward else 0
    out["frozen_vision_fwd"] = (arch.vision_layers * block_fwd(B, 1 + P, Dv) - last_block_saving(B, 1 + P, Dv, 4)) \
        if zero_shot else 0
    out["text_fwd"] = arch.transformer_layers * block_fwd(C, Lt, Dt)
    out["text_bwd"] = arch.transformer_layers * block_bwd(C, Lt, Dt) if text_backward else 0
    heads = 2 * B * Dv * E * (2 if zero_shot else 1) + 2 * C * Dt * E
    out["heads"] = heads * (2 if text_backward else 1)
    out["logits"] = 2 * B * C * E * ((2 if zero_shot else 1) + 2)   # fwd (+zs) and the two gradient GEMMs
    out["total"] = sum(out.values())
    return out


def gemm_shapes(arch, B, C, Lt, n_ctx_vision, zero_shot, vision_backward):
    """(M, N, K, count) of every tcgen05 GEMM launch in one step (forward + dgrad)."""
    Dv, Dt, P = arch.vision_width, arch.transformer_width, arch.n_patches
    shapes = []

    def tower(M, D, layers, bwd):
        shapes.extend([(M, 3 * D, D, layers), (M, D, D, layers), (M, 4 * D, D, layers), (M, D, 4 * D, layers)])
        if bwd:
            shapes.extend([(M, 4 * D, D, layers), (M, D, 4 * D, layers), (M, D, D, layers), (M, D, 3 * D, layers)])

    shapes.append((B * P, Dv, 3 * arch.vision_patch_size ** 2, 1))
    tower(B * (1 + P + n_ctx_vision), Dv, arch.vision_layers, vision_backward)
    if zero_shot:
        tower(B * (1 + P), Dv, arch.vision_layers, False)
    tower(C * Lt, Dt, arch.transformer_layers, True)
    return shapes


def gemm_launches(arch, B, C, Lt, n_ctx_vision, zero_shot, vision_backward, vision_cls_rows=0):
    """(M, N, K, count, epilogue) of every tcgen05 GEMM launch in one step, with the fused epilogue each one runs:
    "none" (bias, 16-bit out), "res" (fp32 residual in, fp32 out), "res_inplace" (forward-only towers: the TMA store
    adds), "gelu" / "gelu_save" (QuickGELU, + saved pre-activation), "gelugrad" (QuickGELU' * aux)."""
    Dv, Dt, P = arch.vision_width, arch.transformer_width, arch.n_patches
    out = [(B * P, Dv, 3 * arch.vision_patch_size ** 2, 1, "none")]

    def tower(M, D, layers, bwd, cls_rows=0):
        res, gelu = ("res", "gelu_save") if bwd else ("res_inplace", "gelu")
        full = layers - 1 if cls_rows else layers      # vision towers: the last block's out-projection / MLP (and
        out.extend([(M, 3 * D, D, layers, "none"), (M, D, D, full, res), (M, 4 * D, D, full, gelu),   # their dgrads)
                    (M, D, 4 * D, full, res)])                                          # run on class-token rows
        if cls_rows:
            out.extend([(cls_rows, D, D, 1, res), (cls_rows, 4 * D, D, 1, gelu), (cls_rows, D, 4 * D, 1, res)])
        if bwd:
            out.extend([(M, 4 * D, D, full, "gelugrad"), (M, D, 4 * D, full, "none"), (M, D, D, full, "none"),
                        (M, D, 3 * D, layers, "none")])
            if cls_rows:
                out.extend([(cls_rows, 4 * D, D, 1, "gelugrad"), (cls_rows, D, 4 * D, 1, "none"),
                            (cls_rows, D, D, 1, "none")])

    tower(B * (1 + P + n_ctx_vision), Dv, arch.vision_layers, vision_backward, cls_rows=vision_cls_rows)
    if zero_shot:
        tower(B * (1 + P), Dv, arch.vision_layers, False, cls_rows=B)
    tower(C * Lt, Dt, arch.transformer_layers, True)
    return out
