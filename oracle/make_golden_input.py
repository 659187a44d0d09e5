"""TEST INFRASTRUCTURE -- golden outputs of the reference's own input pipeline (build container only).

Runs Dassl's `build_transform(cfg, is_train)` (Dassl.pytorch/dassl/data/transforms/transforms.py:176-357) with the
PromptSRC yaml's INPUT block (224 x 224, bicubic, random_resized_crop + random_flip + normalize) on seeded synthetic PIL
images and stores (a) the crop boxes / flip bits it drew, recovered by replaying torchvision under the same seed, and (b)
its output un-normalised back to the uint8 grey levels PIL produced.  tests/test_input_pipeline_cpu.py and the GPU test
hold `psrc_b200.input_pipeline` + `psrc_crop_resize_normalize` against it."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import ref_harness as H  # noqa: E402
from input_ref import IMAGE_SIZES, synthetic_image  # noqa: E402

SEED, R = 4321, 224
MEAN, STD = [0.48145466, 0.4578275, 0.40821073], [0.26862954, 0.26130258, 0.27577711]

if __name__ == "__main__":
    from PIL import Image
    import torchvision.transforms as T
    cfg = H.make_cfg("PromptSRC", **{"INPUT.SIZE": (R, R), "INPUT.INTERPOLATION": "bicubic", "INPUT.PIXEL_MEAN": MEAN,
                                     "INPUT.PIXEL_STD": STD,
                                     "INPUT.TRANSFORMS": ["random_resized_crop", "random_flip", "normalize"]})
    from dassl.data.transforms import build_transform
    tfm_train, tfm_test = build_transform(cfg, is_train=True), build_transform(cfg, is_train=False)
    imgs = [Image.fromarray(synthetic_image(h, w, 100 + k)) for k, (h, w) in enumerate(IMAGE_SIZES)]
    m, s = torch.tensor(MEAN)[:, None, None], torch.tensor(STD)[:, None, None]

    def run(tfm):
        outs = []
        for im in imgs:
            v = (tfm(im) * s + m) * 255.0
            assert float((v - v.round()).abs().max()) < 1e-3       # ToTensor of a uint8 PIL image
            outs.append(v.round().clamp(0, 255).to(torch.uint8).numpy())
        return np.stack(outs)

    torch.manual_seed(SEED)
    train_out = run(tfm_train)
    torch.manual_seed(SEED)
    boxes, flips = [], []
    rrc = T.RandomResizedCrop(R, scale=tuple(cfg.INPUT.RRCROP_SCALE))
    for im in imgs:                                                 # replay: the draws Compose([RRC, Flip]) made
        boxes.append(list(rrc.get_params(im, rrc.scale, rrc.ratio)))
        flips.append(int(torch.rand(1) < 0.5))
    test_out = run(tfm_test)
    path = os.path.join(ROOT, "tests", "golden", "input_pipeline.npz")
    np.savez_compressed(path, seed=np.int32(SEED), size=np.int32(R), train_box=np.array(boxes, dtype=np.int32),
                        train_flip=np.array(flips, dtype=np.int32), train_out=train_out, test_out=test_out)
    print("wrote", path, os.path.getsize(path) / 1e6, "MB; boxes", boxes, "flips", flips)
