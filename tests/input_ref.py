"""TEST INFRASTRUCTURE -- CPU restatement of `psrc_crop_resize_normalize` (include/psrc_b200.h) in matrix form, plus the
seeded synthetic uint8 images the input-pipeline tests use.  out[c] = Wy . window[:, :, c] . Wx^T with the per-axis
weight matrices of PIL's `precompute_coeffs` geometry (src/libImaging/Resample.c of Pillow 12.2, the pinned dependency
behind torchvision / Dassl's Resize and RandomResizedCrop)."""
from __future__ import annotations

import numpy as np


def _filter(x, bicubic):
    x = np.abs(x)
    if not bicubic:
        return np.where(x < 1, 1 - x, 0.0)
    a = -0.5
    return np.where(x < 1, ((a + 2) * x - (a + 3)) * x * x + 1, np.where(x < 2, (((x - 5) * x + 8) * x - 4) * a, 0.0))


def axis_weights(n_out, scale, off, size, bicubic):
    """[n_out, size] float64 weight matrix of one axis."""
    fs = max(scale, 1.0)
    support = (2.0 if bicubic else 1.0) * fs
    Wm = np.zeros((n_out, size))
    for o in range(n_out):
        c = np.float32((np.float32(o) + np.float32(off) + np.float32(0.5)) * np.float32(scale))   # kernel computes in fp32
        lo = max(int(np.float32(c - np.float32(support) + np.float32(0.5))), 0)
        hi = min(int(np.float32(c + np.float32(support) + np.float32(0.5))), size)
        k = np.arange(lo, hi)
        w = _filter((k - float(c) + 0.5) / fs, bicubic)
        if w.sum() != 0:
            Wm[o, lo:hi] = w / w.sum()
    return Wm


def crop_resize_normalize(src_u8, geom_f, geom_i, R, interpolation, quantize, mean, std):
    """src_u8: uint8 [B, Hc, Wc, 3] numpy; returns fp32 [B, 3, R, R] numpy."""
    B = src_u8.shape[0]
    out = np.zeros((B, 3, R, R), dtype=np.float32)
    for b in range(B):
        sx, sy, ox, oy = [float(v) for v in geom_f[b]]
        left, top, w, h, flip = [int(v) for v in geom_i[b]]
        win = src_u8[b, top:top + h, left:left + w].astype(np.float64)
        Wy, Wx = axis_weights(R, sy, oy, h, interpolation == 1), axis_weights(R, sx, ox, w, interpolation == 1)
        for c in range(3):
            hp = win[:, :, c] @ Wx.T                     # horizontal pass [h, R]
            if quantize == 2:
                hp = np.clip(np.rint(hp), 0, 255)
            v = Wy @ hp
            if quantize:
                v = np.clip(np.rint(v), 0, 255)
            if flip:
                v = v[:, ::-1]
            out[b, c] = ((v / 255.0 - mean[c]) / std[c]).astype(np.float32)
    return out


def synthetic_image(h, w, seed):
    """Smooth low-frequency colour field + texture + noise, uint8 HWC (deterministic)."""
    rng = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    img = np.zeros((h, w, 3))
    for c in range(3):
        for _ in range(4):
            fy, fx, ph = rng.uniform(0.005, 0.08, 2).tolist() + [rng.uniform(0, 6.28)]
            img[:, :, c] += rng.uniform(20, 60) * np.sin(fy * yy + fx * xx + ph)
    img += 128 + rng.normal(0, 12, size=(h, w, 3))
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


IMAGE_SIZES = [(375, 500), (500, 333), (224, 224), (180, 240), (64, 96), (768, 1024)]
