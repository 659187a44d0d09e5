"""CPU checks of the on-device input pipeline's host side: the augmentation-parameter stream is
bit-identical to torchvision's (the library behind Dassl's transforms), and the matrix-form restatement the GPU kernel
is held against agrees with torch's antialiased resize and with the reference's own PIL pipeline (golden fixture)."""
import os

import numpy as np
import pytest
import torch

import psrc_b200  # noqa: F401
from psrc_b200 import input_pipeline as IP
from input_ref import IMAGE_SIZES, crop_resize_normalize, synthetic_image

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "input_pipeline.npz")


def test_parameter_stream_matches_torchvision():
    tv = pytest.importorskip("torchvision.transforms")
    rrc, flip = tv.RandomResizedCrop(224, scale=(0.08, 1.0)), tv.RandomHorizontalFlip()
    sizes = [(375, 500), (500, 333), (224, 224), (32, 900), (900, 20), (1, 1), (1200, 1600)]
    torch.manual_seed(1234)
    ref = []
    for (h, w) in sizes * 40:
        img = torch.zeros(3, h, w, dtype=torch.uint8)
        ref.append((tuple(rrc.get_params(img, rrc.scale, rrc.ratio)), bool(torch.rand(1) < flip.p)))
    torch.manual_seed(1234)
    ours = [(IP.rrc_params(h, w), IP.flip_param()) for (h, w) in sizes * 40]
    assert ours == ref
    assert torch.rand(1).item() == (torch.manual_seed(1234), [(IP.rrc_params(h, w), IP.flip_param()) for (h, w) in sizes * 40],
                                    torch.rand(1).item())[2]     # the generator is left in the same state


def test_parameter_stream_matches_reference_fixture():
    """Same check without torchvision: the parameters the reference's own Dassl transform drew (oracle/make_golden_input.py)."""
    fx = np.load(GOLD)
    torch.manual_seed(int(fx["seed"]))
    for k, (h, w) in enumerate(IMAGE_SIZES):
        box, fl = IP.rrc_params(h, w), IP.flip_param()
        assert list(box) == fx["train_box"][k].tolist() and int(fl) == int(fx["train_flip"][k])


def _run_restatement(tf, imgs, R):
    items = [tf(im) for im in imgs]
    u8 = np.stack([it["u8"].numpy() for it in items])
    gf = np.stack([it["geom_f"].numpy() for it in items])
    gi = np.stack([it["geom_i"].numpy() for it in items])
    return u8, gf, gi


@pytest.mark.parametrize("is_train", [True, False])
def test_restatement_matches_reference_pil_pipeline(is_train):
    """DeviceTransform geometry + the restatement (two-pass uint8 rounding like PIL) against the output of Dassl's
    build_transform(cfg, is_train) run on PIL images by the unmodified reference (stored un-normalised as uint8)."""
    fx = np.load(GOLD)
    R = int(fx["size"])
    imgs = [synthetic_image(h, w, 100 + k) for k, (h, w) in enumerate(IMAGE_SIZES)]
    torch.manual_seed(int(fx["seed"]))
    tf = IP.DeviceTransform(size=R, is_train=is_train, canvas=(1024, 1024))
    u8, gf, gi = _run_restatement(tf, imgs, R)
    out = crop_resize_normalize(u8, gf, gi, R, 1, 2, (0, 0, 0), (1, 1, 1)) * 255.0
    ref = fx["train_out" if is_train else "test_out"].astype(np.float32)
    d = np.abs(out - ref)
    # PIL uses 8-bit fixed-point coefficients: float weights land on the other side of a rounding boundary on < 1 % of the
    # pixels (measured 0.04 % when down-scaling, 0.8 % on the 3.5x up-scaled noisy image), never by more than 2 grey levels
    assert d.max() <= 2.0 + 1e-3, d.max()
    assert (d > 0.5).mean() < 0.01, (d > 0.5).mean()
    assert (d > 1.5).mean() < 1e-4, (d > 1.5).mean()


@pytest.mark.parametrize("is_train", [True, False])
@pytest.mark.parametrize("canvas", [(512, 512), (256, 256)])
def test_canvas_reduction_deviation_from_pil_is_bounded(is_train, canvas):
    """windows larger than the canvas go through an integer box reduction that PIL's resize does not do, so
    those images are NOT within the 1-2 grey levels of the other tests.  Same reference fixture (Dassl's PIL chain), same
    seed, the default 512 canvas and the 256 one: images whose window fits stay within 2 levels; reduced ones are held to
    the deviation measured when this test was written (the numbers input_pipeline.py's docstring quotes)."""
    fx = np.load(GOLD)
    R = int(fx["size"])
    imgs = [synthetic_image(h, w, 100 + k) for k, (h, w) in enumerate(IMAGE_SIZES)]
    torch.manual_seed(int(fx["seed"]))
    big = IP.DeviceTransform(size=R, is_train=is_train, canvas=(1024, 1024))
    _, _, gi_full = _run_restatement(big, imgs, R)            # the windows before any reduction
    torch.manual_seed(int(fx["seed"]))
    tf = IP.DeviceTransform(size=R, is_train=is_train, canvas=canvas)
    u8, gf, gi = _run_restatement(tf, imgs, R)
    assert u8.shape[1:3] == canvas
    out = crop_resize_normalize(u8, gf, gi, R, 1, 2, (0, 0, 0), (1, 1, 1)) * 255.0
    ref = fx["train_out" if is_train else "test_out"].astype(np.float32)
    reduced = 0
    for k in range(len(imgs)):
        d = np.abs(out[k] - ref[k])
        w0, h0 = int(gi_full[k][2]), int(gi_full[k][3])
        if w0 <= canvas[1] and h0 <= canvas[0]:
            assert d.max() <= 2.0 + 1e-3, (k, d.max())           # untouched by the canvas size
            continue
        reduced += 1
        upscaled_again = int(gi[k][2]) < R or int(gi[k][3]) < R
        lim_mean, lim_max = (4.0, 32.0) if upscaled_again else (1.0, 6.0)
        assert d.mean() <= lim_mean and d.max() <= lim_max, (k, (w0, h0), tuple(gi[k][2:4]), d.mean(), d.max())
    assert reduced >= (1 if canvas == (512, 512) else 2)       # the fixture does exercise the reduction


def test_restatement_matches_torch_antialias_resize():
    img = synthetic_image(300, 420, 7)
    top, left, h, w, R = 17, 40, 233, 301, 224
    u8 = img[None]
    gf = np.array([[w / R, h / R, 0, 0]], dtype=np.float32)
    gi = np.array([[left, top, w, h, 1]], dtype=np.int32)
    for interp, mode in ((1, "bicubic"), (0, "bilinear")):
        ours = crop_resize_normalize(u8, gf, gi, R, interp, 0, (0, 0, 0), (1, 1, 1)) * 255.0
        crop = torch.from_numpy(img[top:top + h, left:left + w].copy()).permute(2, 0, 1).float()[None]
        ref = torch.nn.functional.interpolate(crop, size=(R, R), mode=mode, antialias=True, align_corners=False)
        ref = ref.flip(-1).numpy()
        assert np.abs(ours - ref).max() < 2e-3, (mode, np.abs(ours - ref).max())


def test_large_window_is_reduced_to_the_canvas():
    img = synthetic_image(768, 1024, 3)
    torch.manual_seed(5)
    tf = IP.DeviceTransform(size=224, is_train=False, canvas=(256, 256))
    it = tf(img)
    assert it["u8"].shape == (256, 256, 3) and int(it["geom_i"][2]) <= 256 and int(it["geom_i"][3]) <= 256
    assert IP.max_taps(it["geom_f"].numpy()[None], 1) <= 16
