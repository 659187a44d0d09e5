"""Host side of the on-device input pipeline.

The reference feeds the step through Dassl's per-image PIL chain in DataLoader worker processes
(Dassl.pytorch/dassl/data/transforms/transforms.py:176-326, `data_manager.py:13-52`): at the step rates of this package
(thousands of images per second per GPU) eight PIL workers cannot keep up and every batch crosses PCIe as fp32
(602 KB per image).  Here the workers only DRAW the augmentation parameters and hand over the uint8 pixels of the
sampled window; resampling, flip, /255 and normalisation run in one CUDA launch (`csrc/preprocess.cu`).

Index / label / parameter flow is bit-identical to the reference: `rrc_params` and `flip_param` make exactly the torch
RNG calls torchvision's `RandomResizedCrop.get_params` and `RandomHorizontalFlip.forward` make, in the same order
(checked against torchvision in tests/test_input_pipeline_cpu.py).  Pixels are not bit-identical to PIL (float weights
instead of PIL's 8-bit fixed point): <= 1 grey level on nearly every pixel, gated in the tests -- AS LONG AS the sampled
window fits the canvas.  A window larger than the canvas is first shrunk by an integer box filter (`_reduce`), which
the reference's PIL `resize` does not do; measured against the reference's PIL output on noisy synthetic images
(tests/test_input_pipeline_cpu.py::test_canvas_reduction_deviation_from_pil_is_bounded):
  - the shrunk window is still larger than the output (768 x 1024 image on the default 512 canvas): mean 0.8, max 5 levels;
  - the shrunk window falls below the output size and is scaled up again (a 330 x 370 window on a 256 canvas ->
    165 x 185 -> 224): mean 3, max 24 grey levels.  A 256 canvas does this to ordinary ImageNet-sized images, so the
    default canvas is 512 x 512 (786 KB per image over PCIe, more than the 602 KB of the fp32 tensor: the gain left is
    the resampling work taken off the loader processes).  Choose a smaller canvas only for datasets whose images fit
    it (CIFAR-sized sets, pre-shrunk caches) or when that pixel error is acceptable.

    tfm = DeviceTransform(size=224, is_train=True)                 # use as Dassl `custom_tfm_train`
    item = tfm(pil_image)          -> {"u8": uint8 [Hc, Wc, 3], "geom_f": fp32 [4], "geom_i": int32 [5]}
    (default collate stacks them)  -> batch["img"] = {"u8": [B, Hc, Wc, 3], ...}
    pre = DevicePreprocessor(device, batch, size=224)
    pre.run(batch["img"], out=step.image)                          # H2D of the uint8 canvases + one kernel
"""
from __future__ import annotations

import math

import numpy as np
import torch

from . import ops

CLIP_MEAN = (0.48145466, 0.4578275, 0.40821073)     # PromptSRC yaml INPUT.PIXEL_MEAN / PIXEL_STD
CLIP_STD = (0.26862954, 0.26130258, 0.27577711)
INTERPOLATION = {"bilinear": 0, "bicubic": 1}


def rrc_params(height: int, width: int, scale=(0.08, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0)):
    """Crop box (top, left, h, w) of a random-resized crop, drawn with the same torch RNG calls, in the same order, as
    torchvision's RandomResizedCrop.get_params (the transform behind Dassl's "random_resized_crop",
    transforms.py:237-243): up to ten (area, log-uniform aspect) tries, each accepted try followed by two randint
    draws for the position; centre-crop fallback."""
    area = height * width
    log_ratio = torch.log(torch.tensor(ratio))
    for _ in range(10):
        target_area = area * torch.empty(1).uniform_(scale[0], scale[1]).item()
        aspect = torch.exp(torch.empty(1).uniform_(log_ratio[0], log_ratio[1])).item()
        w = int(round(math.sqrt(target_area * aspect)))
        h = int(round(math.sqrt(target_area / aspect)))
        if 0 < w <= width and 0 < h <= height:
            top = torch.randint(0, height - h + 1, size=(1,)).item()
            left = torch.randint(0, width - w + 1, size=(1,)).item()
            return top, left, h, w
    in_ratio = float(width) / float(height)
    if in_ratio < min(ratio):
        w, h = width, int(round(width / min(ratio)))
    elif in_ratio > max(ratio):
        h, w = height, int(round(height * max(ratio)))
    else:
        w, h = width, height
    return (height - h) // 2, (width - w) // 2, h, w


def flip_param(p: float = 0.5) -> bool:
    """RandomHorizontalFlip.forward's draw (`torch.rand(1) < p`), Dassl "random_flip" (transforms.py:245-247)."""
    return bool(torch.rand(1) < p)


def eval_geometry(height: int, width: int, size: int):
    """Resize(size) on the smaller edge followed by CenterCrop(size) (transforms.py:337-341) as resampling geometry:
    returns (new_h, new_w, crop_top, crop_left) in the resized image."""
    if height <= width:
        new_h, new_w = size, int(size * width / height)
    else:
        new_w, new_h = size, int(size * height / width)
    return new_h, new_w, int(round((new_h - size) / 2.0)), int(round((new_w - size) / 2.0))


def max_taps(geom_f, interpolation: int) -> int:
    """Largest filter window of a batch: 2 * support + 2 with support = base * max(scale, 1)."""
    base = 2.0 if interpolation == 1 else 1.0
    s = float(max(1.0, float(np.asarray(geom_f)[..., :2].max())))
    return int(2 * math.ceil(base * s) + 2)


def _reduce(arr: np.ndarray, f: int) -> np.ndarray:
    """Integer box reduction (what PIL's `Image.reduce` / `resize(reducing_gap=...)` does) for windows larger than the
    canvas; rounds half up like PIL."""
    h, w = arr.shape[:2]
    hh, ww = -(-h // f) * f, -(-w // f) * f
    if (hh, ww) != (h, w):
        arr = np.pad(arr, ((0, hh - h), (0, ww - w), (0, 0)), mode="edge")
    v = arr.reshape(hh // f, f, ww // f, f, 3).astype(np.uint32).sum(axis=(1, 3))
    return ((v + (f * f) // 2) // (f * f)).astype(np.uint8)


class DeviceTransform:
    """Drop-in for the transform Dassl's `build_transform(cfg, is_train)` returns (`DataManager(custom_tfm_train=...,
    custom_tfm_test=...)`, data_manager.py:64-85): called on one PIL image (or HWC uint8 array) inside a DataLoader
    worker, it draws the augmentation parameters and returns the uint8 window + geometry instead of a float tensor."""

    def __init__(self, size=224, is_train=True, scale=(0.08, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0), flip_p=0.5,
                 canvas=(512, 512), random_flip=True):
        self.size, self.is_train, self.scale, self.ratio = int(size), bool(is_train), tuple(scale), tuple(ratio)
        self.flip_p, self.canvas, self.random_flip = flip_p, tuple(canvas), random_flip

    @classmethod
    def from_cfg(cls, cfg, is_train=True, canvas=(512, 512)):
        """Same cfg keys as transforms.py:176-243; only the PromptSRC-family choices are expressible on the device."""
        choices = list(cfg.INPUT.TRANSFORMS)
        known = {"random_resized_crop", "random_flip", "normalize"}
        if is_train and (set(choices) - known or "random_resized_crop" not in choices):
            raise NotImplementedError(f"device input pipeline covers {sorted(known)}; cfg.INPUT.TRANSFORMS = {choices}")
        size = cfg.INPUT.SIZE
        assert size[0] == size[1], "square inputs (CLIP)"
        return cls(size=size[0], is_train=is_train, scale=tuple(cfg.INPUT.RRCROP_SCALE), canvas=canvas,
                   random_flip="random_flip" in choices)

    def __call__(self, img):
        arr = np.asarray(img)
        if arr.ndim == 2:
            arr = np.repeat(arr[:, :, None], 3, axis=2)
        arr = arr[:, :, :3]
        assert arr.dtype == np.uint8
        H, W = arr.shape[:2]
        R = self.size
        if self.is_train:
            top, left, h, w = rrc_params(H, W, self.scale, self.ratio)
            flip = flip_param(self.flip_p) if self.random_flip else False
            win = arr[top:top + h, left:left + w]
            sx, sy, ox, oy = w / R, h / R, 0.0, 0.0
        else:
            new_h, new_w, ct, cl = eval_geometry(H, W, R)
            sx, sy = W / new_w, H / new_h
            mx, my = 2.0 * max(sx, 1.0) + 1.0, 2.0 * max(sy, 1.0) + 1.0      # bicubic support margin
            x0, x1 = max(0, int(cl * sx - mx)), min(W, int(math.ceil((cl + R) * sx + mx)))
            y0, y1 = max(0, int(ct * sy - my)), min(H, int(math.ceil((ct + R) * sy + my)))
            win = arr[y0:y1, x0:x1]
            ox, oy, flip = cl - x0 / sx, ct - y0 / sy, False
        Hc, Wc = self.canvas
        f = max(1, -(-win.shape[0] // Hc), -(-win.shape[1] // Wc))
        if f > 1:
            win = _reduce(win, f)
            sx, sy = sx / f, sy / f
        h, w = win.shape[:2]
        canvas = np.zeros((Hc, Wc, 3), dtype=np.uint8)
        canvas[:h, :w] = win
        return {"u8": torch.from_numpy(canvas),
                "geom_f": torch.tensor([sx, sy, ox, oy], dtype=torch.float32),
                "geom_i": torch.tensor([0, 0, w, h, int(flip)], dtype=torch.int32)}


class DevicePreprocessor:
    """Device half: pinned-host uint8 canvases -> H2D on a copy stream -> `psrc_crop_resize_normalize` into the step's
    fp32 image buffer.  Buffers are allocated once; `stage()` may be called for batch i+1 as soon as batch i's kernel
    has been launched (it waits for that kernel only, not for the step that follows it) -- the same look-ahead contract
    as `PromptedClipStep.stage_host`."""

    def __init__(self, device, batch, size=224, canvas=(512, 512), interpolation="bicubic", quantize=2,
                 mean=CLIP_MEAN, std=CLIP_STD):
        self.device, self.batch, self.size = torch.device(device), batch, size
        self.interp, self.quantize, self.mean, self.std = INTERPOLATION[interpolation], int(quantize), mean, std
        self.u8 = torch.empty(batch, canvas[0], canvas[1], 3, dtype=torch.uint8, device=self.device)
        self.geom_f = torch.empty(batch, 4, dtype=torch.float32, device=self.device)
        self.geom_i = torch.empty(batch, 5, dtype=torch.int32, device=self.device)
        self.labels = torch.empty(batch, dtype=torch.int32, device=self.device)
        self.taps = 0
        self._copy, self._ready, self._consumed = None, None, None
        self.staged_key = None

    def bytes_per_batch(self) -> int:
        return self.u8.numel() + self.geom_f.numel() * 4 + self.geom_i.numel() * 4 + self.labels.numel() * 8

    def stage(self, item: dict, label_host=None):
        """Start the H2D copies of one collated batch ({"u8", "geom_f", "geom_i"}, pinned for true overlap)."""
        if self._copy is None:
            self._copy, self._ready = torch.cuda.Stream(device=self.device), torch.cuda.Event()
        self.taps = max_taps(item["geom_f"].numpy(), self.interp)
        if self._consumed is not None:
            self._copy.wait_event(self._consumed)                # the previous batch has left the staging buffers
        with torch.cuda.stream(self._copy):
            self.u8.copy_(item["u8"], non_blocking=True)
            self.geom_f.copy_(item["geom_f"], non_blocking=True)
            self.geom_i.copy_(item["geom_i"], non_blocking=True)
            if label_host is not None:
                self.labels.copy_(label_host.to(torch.int32), non_blocking=True)
            self._ready.record(self._copy)
        self.staged_key = id(item["u8"])

    def run(self, item: dict | None, out: torch.Tensor, labels_out: torch.Tensor | None = None, label_host=None):
        """Resample the staged (or given) batch into `out` [B, 3, R, R] fp32 on the current stream."""
        if item is not None and self.staged_key != id(item["u8"]):
            self.stage(item, label_host)
        main = torch.cuda.current_stream()
        main.wait_event(self._ready)
        ops.crop_resize_normalize(self.u8, self.geom_f, self.geom_i, out, self.interp, self.quantize, self.taps,
                                  self.mean, self.std)
        if labels_out is not None:
            labels_out.copy_(self.labels, non_blocking=True)
        if self._consumed is None:
            self._consumed = torch.cuda.Event()
        self._consumed.record(main)
        self.staged_key = None
        return out
