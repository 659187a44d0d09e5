"""GPU parity of the on-device input pipeline: `psrc_crop_resize_normalize` through the C ABI against the
matrix-form restatement (tests/input_ref.py) and against the output of the reference's own PIL transform chain
(tests/golden/input_pipeline.npz, oracle/make_golden_input.py); and the uint8-fed train step against the fp32-fed one."""
import os

import numpy as np
import pytest
import torch

import psrc_b200  # noqa: F401
from psrc_b200 import input_pipeline as IP, ops
from input_ref import IMAGE_SIZES, crop_resize_normalize, synthetic_image

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "input_pipeline.npz")


def _batch(is_train, R, canvas, seed):
    imgs = [synthetic_image(h, w, 100 + k) for k, (h, w) in enumerate(IMAGE_SIZES)]
    torch.manual_seed(seed)
    tf = IP.DeviceTransform(size=R, is_train=is_train, canvas=canvas)
    items = [tf(im) for im in imgs]
    return {k: torch.stack([it[k] for it in items]) for k in ("u8", "geom_f", "geom_i")}


def _gpu(item, R, interp, quantize, mean, std):
    dev = "cuda"
    out = torch.empty(item["u8"].shape[0], 3, R, R, device=dev)
    taps = IP.max_taps(item["geom_f"].numpy(), interp)
    ops.crop_resize_normalize(item["u8"].to(dev), item["geom_f"].to(dev), item["geom_i"].to(dev), out, interp, quantize,
                              taps, mean, std)
    torch.cuda.synchronize()
    return out.cpu().numpy()


@pytest.mark.parametrize("is_train", [True, False])
@pytest.mark.parametrize("interp", [1, 0])
@pytest.mark.parametrize("quantize", [0, 1, 2])
def test_kernel_matches_restatement(is_train, interp, quantize):
    R = 224
    item = _batch(is_train, R, (1024, 1024), 4321)
    ours = _gpu(item, R, interp, quantize, (0, 0, 0), (1, 1, 1)) * 255.0
    ref = crop_resize_normalize(item["u8"].numpy(), item["geom_f"].numpy(), item["geom_i"].numpy(), R, interp, quantize,
                                (0, 0, 0), (1, 1, 1)) * 255.0
    d = np.abs(ours - ref)
    if quantize == 0:
        assert d.max() < 5e-3, d.max()                      # fp32 accumulation order, in grey levels
    else:                                                   # rounding ties: a handful of pixels, one grey level
        assert d.max() <= 1.0 + 1e-3 and (d > 0.5).mean() < (1e-2 if quantize == 2 else 2e-3), (d.max(), (d > 0.5).mean())


def test_kernel_normalises_like_the_reference():
    R = 224
    item = _batch(True, R, (1024, 1024), 4321)
    ours = _gpu(item, R, 1, 2, IP.CLIP_MEAN, IP.CLIP_STD)
    ref = crop_resize_normalize(item["u8"].numpy(), item["geom_f"].numpy(), item["geom_i"].numpy(), R, 1, 2,
                                IP.CLIP_MEAN, IP.CLIP_STD)
    d = np.abs(ours - ref)
    assert (d > 1e-5).mean() < 2e-3 and d.max() < 0.016     # one grey level = 1 / (255 * 0.26) = 0.015


@pytest.mark.parametrize("is_train", [True, False])
def test_kernel_matches_reference_pil_pipeline(is_train):
    """Against the pixels Dassl's build_transform produced on PIL images (parameters drawn from the same seed)."""
    fx = np.load(GOLD)
    R = int(fx["size"])
    item = _batch(is_train, R, (1024, 1024), int(fx["seed"]))
    if is_train:
        assert item["geom_i"][:, 2:4].tolist() == fx["train_box"][:, [3, 2]].tolist()      # (w, h) of the drawn boxes
        assert item["geom_i"][:, 4].tolist() == fx["train_flip"].tolist()
    ours = _gpu(item, R, 1, 2, (0, 0, 0), (1, 1, 1)) * 255.0
    d = np.abs(ours - fx["train_out" if is_train else "test_out"].astype(np.float32))
    assert d.max() <= 2.0 + 1e-3 and (d > 0.5).mean() < 0.01 and (d > 1.5).mean() < 1e-4, (d.max(), (d > 0.5).mean())


def test_canvas_reduction_and_u8_fed_step_equals_fp32_fed_step():
    """train_step_host_u8 (uint8 windows in, device resample) == train_step_host fed with the same pixels as fp32."""
    from helpers import promptsrc_setup
    from psrc_b200 import synthetic as S
    R, B, dev = 64, 3, "cuda"
    tokens = S.synthetic_tokens(10, name_lens=[1, 2, 3, 8, 1, 1, 2, 2, 1, 2])
    imgs = [synthetic_image(h, w, 300 + k) for k, (h, w) in enumerate([(90, 120), (300, 200), (64, 64)])]
    torch.manual_seed(9)
    tf = IP.DeviceTransform(size=R, is_train=True, canvas=(128, 128))
    items = [tf(im) for im in imgs]
    item = {k: torch.stack([it[k] for it in items]).pin_memory() for k in ("u8", "geom_f", "geom_i")}
    label = torch.tensor([1, 7, 3])
    ref_img = torch.from_numpy(crop_resize_normalize(item["u8"].numpy(), item["geom_f"].numpy(), item["geom_i"].numpy(),
                                                     R, 1, 2, IP.CLIP_MEAN, IP.CLIP_STD))
    losses, params = [], []
    for mode in ("u8", "fp32"):
        step, _, _ = promptsrc_setup("tiny", 3, tokens, dev, "fp32", batch=B)
        step.init_optimizer()
        if mode == "u8":
            loss = step.train_step_host_u8(item, label)
            img_dev = step.image.cpu()
        else:
            loss = step.train_step_host(img_dev.pin_memory(), label)
        losses.append(loss)
        params.append(step.flat_param.cpu().clone())
    assert (img_dev - ref_img).abs().max() < 0.016 and ((img_dev - ref_img).abs() > 1e-5).float().mean() < 5e-3
    assert losses[0] == losses[1] and torch.equal(params[0], params[1])
