"""TEST INFRASTRUCTURE -- the drop-in under the REAL Dassl (build container only: needs /root/reference).

Runs Dassl's own `build_trainer(cfg)` -> `SimpleTrainer.__init__` -> `trainer.train()` (run_epoch / after_epoch /
save_model / test with the Classification evaluator; Dassl.pytorch/dassl/engine/build.py:3-11, trainer.py:118-201,
:296-486, :596-639) twice on the same stub DataManager and the same synthetic CLIP weights:
  (1) with this package's trainer, selected by name from Dassl's TRAINER_REGISTRY exactly as `train.py --trainer X`
      does, its kernels replaced by the plain-torch restatements (tests/torch_kernels.py) so it runs on this CPU box;
  (2) with the unmodified reference trainer class.
and compares per-batch losses, the final prompt parameters, the test accuracy and a checkpoint round trip in both
directions (reference-written file -> our model, our file -> reference model).  Prints one JSON line.
Run in a fresh interpreter (tests/test_dassl_dropin_cpu.py): Dassl must be importable BEFORE psrc_b200.trainers is.
"""
from __future__ import annotations

import contextlib
import importlib
import io
import json
import os
import sys
import tempfile
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

import ref_harness as R  # noqa: E402

CLASSNAMES = ["cat", "dog", "big red apple", "aeroplane", "fire truck", "owl", "sea lion", "tree", "cup", "violin"]
TRAINER_KEYS = {"PromptSRC": "PROMPTSRC", "CoOp": "COOP", "MaPLe": "MAPLE", "CoCoOp": "COCOOP"}


def make_cfg(trainer, outdir, epochs=2):
    over = {"OUTPUT_DIR": outdir, "USE_CUDA": False, "INPUT.SIZE": (64, 64), "OPTIM.NAME": "sgd", "OPTIM.LR": 0.0025,
            "OPTIM.MAX_EPOCH": epochs, "OPTIM.LR_SCHEDULER": "cosine", "OPTIM.WARMUP_EPOCH": 1,
            "OPTIM.WARMUP_TYPE": "constant", "OPTIM.WARMUP_CONS_LR": 1e-5, "DATALOADER.TRAIN_X.BATCH_SIZE": 3,
            "DATALOADER.TEST.BATCH_SIZE": 5, "TRAIN.PRINT_FREQ": 1, "TRAIN.CHECKPOINT_FREQ": 0,
            "TEST.FINAL_MODEL": "last_step", "TEST.NO_TEST": False, "MODEL.BACKBONE.NAME": "ViT-B/16"}
    if trainer == "PromptSRC":
        over.update({"TRAINER.PROMPTSRC.PROMPT_DEPTH_VISION": 3, "TRAINER.PROMPTSRC.PROMPT_DEPTH_TEXT": 3,
                     "TRAINER.PROMPTSRC.N_CTX_VISION": 4, "TRAINER.PROMPTSRC.N_CTX_TEXT": 4,
                     "TRAINER.PROMPTSRC.CTX_INIT": "a photo of a", "TRAINER.PROMPTSRC.LOSS_TYPE": "ce"})
    elif trainer == "CoOp":
        over.update({"TRAINER.COOP.N_CTX": 4, "TRAINER.COOP.CTX_INIT": "a photo of a", "TRAINER.COOP.LOSS_TYPE": "ce"})
    return R.make_cfg(trainer, **over)


def batches(n, bs, seed):
    from psrc_b200 import synthetic as S
    out = []
    for i in range(n):
        img, lab = S.synthetic_batch(bs, len(CLASSNAMES), 64, seed=seed + i)
        out.append({"img": img, "label": lab, "impath": [f"synthetic/{seed + i}_{k}.jpg" for k in range(bs)]})
    return out


class StubDataManager:
    """What SimpleTrainer.build_data_loader reads from dassl.data.DataManager (trainer.py:341-358)."""

    last_kwargs = None      # what the trainer's build_data_loader passed (custom_tfm_train / custom_tfm_test)

    def __init__(self, cfg, *a, **k):
        from psrc_b200 import synthetic as S  # noqa: F401
        StubDataManager.last_kwargs = dict(k)
        self.train_loader_x = batches(2, 3, seed=1)
        self.train_loader_u, self.val_loader = None, None
        test = batches(1, 5, seed=50) + batches(1, 2, seed=60)           # a ragged last test batch
        self.test_loader = test
        self.num_classes, self.num_source_domains = len(CLASSNAMES), 1
        self.lab2cname = dict(enumerate(CLASSNAMES))
        self.dataset = types.SimpleNamespace(classnames=list(CLASSNAMES), lab2cname=self.lab2cname,
                                             num_classes=len(CLASSNAMES))


class _MP:
    def setattr(self, obj, name, value, raising=True):
        setattr(obj, name, value)


def run(trainer_name: str) -> dict:
    R.setup()
    importlib.import_module("train")               # registers the REFERENCE trainers in Dassl's registry (train.py:16-27)
    import dassl.engine.trainer as dtr
    from dassl.engine import TRAINER_REGISTRY, build_trainer
    ref_cls = TRAINER_REGISTRY.get(trainer_name)
    import psrc_b200  # noqa: F401  (imports psrc_b200.trainers: registers the drop-ins with force=True)
    from psrc_b200 import synthetic as S, trainers as T
    from torch_kernels import TorchKernels
    assert T.HAVE_DASSL and T.TrainerX is dtr.TrainerX, "the drop-in must subclass Dassl's own TrainerX"
    ours_cls = TRAINER_REGISTRY.get(trainer_name)
    assert ours_cls is getattr(T, trainer_name) and ours_cls is not ref_cls
    dtr.DataManager = StubDataManager
    # torch >= 2.7 compatibility of the reference's pinned-torch-1.9 code, harness side only: Dassl's warm-up schedulers
    # pass `verbose` positionally to _LRScheduler.__init__ (dassl/optim/lr_scheduler.py:22), which no longer takes it
    import dassl.optim.lr_scheduler as LS

    def _warmup_init(self, optimizer, successor, warmup_epoch, last_epoch=-1, verbose=False):
        self.successor, self.warmup_epoch = successor, warmup_epoch
        torch.optim.lr_scheduler._LRScheduler.__init__(self, optimizer, last_epoch)
    LS._BaseWarmupScheduler.__init__ = _warmup_init
    arch = S.ARCHS["tiny"]
    sd = S.make_clip_state_dict(arch, seed=0)
    tm = importlib.import_module(f"trainers.{trainer_name.lower()}")
    R._patched_loader(tm, sd, trainer_name)         # load_clip_to_cpu -> synthetic weights (no download)
    TorchKernels().install(_MP())
    m = {"trainer": trainer_name, "registry_class": f"{ours_cls.__module__}.{ours_cls.__name__}"}
    quiet = io.StringIO()

    def losses_of(tr):
        rec = []
        fb = tr.forward_backward

        def wrapped(batch):
            out = fb(batch)
            rec.append(float(out["loss"]))
            return out
        tr.forward_backward = wrapped
        return rec

    with tempfile.TemporaryDirectory() as d_ours, tempfile.TemporaryDirectory() as d_ref:
        torch.manual_seed(0)
        with R.cpu_patches(half_roundtrip=(trainer_name == "PromptSRC")), contextlib.redirect_stdout(quiet):
            ours = build_trainer(make_cfg(trainer_name, d_ours))           # Dassl's own factory (engine/build.py:3-11)
            assert type(ours) is ours_cls
            lo = losses_of(ours)
            ours.train()
            acc_ours = ours.test()
            torch.manual_seed(0)
            ref = ref_cls(make_cfg(trainer_name, d_ref))
            R.freeze_like_trainer(trainer_name, ref.model)
            lr_ = losses_of(ref)
            ref.train()
            acc_ref = ref.test()
        m["n_batches"] = len(lo)
        m["loss_maxabs"] = max(abs(a - b) for a, b in zip(lo, lr_))
        m["acc_ours"], m["acc_ref"] = float(acc_ours), float(acc_ref)
        with torch.no_grad(), R.cpu_patches(half_roundtrip=(trainer_name == "PromptSRC")):
            ours.set_model_mode("eval"), ref.set_model_mode("eval")
            tb = ours.test_loader[0]["img"]
            lo_, lr__ = ours.model_inference(tb), ref.model_inference(tb)
        m["test_logits_maxabs"] = float((lo_ - lr__).abs().max())
        m["test_argmax_equal"] = bool((lo_.argmax(1) == lr__.argmax(1)).all())
        ours_named = {k: v.detach() for k, v in ours.model.named_parameters() if v.requires_grad}
        ref_named = {k: v.detach() for k, v in ref.model.named_parameters() if v.requires_grad}
        m["trainable_names_equal"] = sorted(ours_named) == sorted(ref_named)
        m["param_maxabs"] = max(float((ours_named[k] - ref_named[k]).abs().max()) for k in ref_named)
        # checkpoints written by Dassl's save_model (after_epoch of the last epoch): same files, same keys
        name = ours.get_model_names()[0]
        m["model_name"], m["ref_model_name"] = name, ref.get_model_names()[0]
        f_ours = os.path.join(d_ours, name, "model.pth.tar-2")
        f_ref = os.path.join(d_ref, name, "model.pth.tar-2")
        m["checkpoint_files"] = [os.path.exists(f_ours), os.path.exists(f_ref)]
        ck_o = torch.load(f_ours, map_location="cpu", weights_only=False)
        ck_r = torch.load(f_ref, map_location="cpu", weights_only=False)
        m["ckpt_keys_missing_in_ours"] = sorted(set(ck_r["state_dict"]) - set(ck_o["state_dict"]))[:8]
        # cross-load: perturb, then load the OTHER implementation's file through Dassl's load_model
        with contextlib.redirect_stdout(quiet), R.cpu_patches(half_roundtrip=(trainer_name == "PromptSRC")):
            with torch.no_grad():
                for v in ours.model.parameters():
                    if v.requires_grad:
                        v.add_(1.0)
            ours.load_model(d_ref, epoch=2)
            ours_after = {k: v.detach() for k, v in ours.model.named_parameters() if v.requires_grad}
            m["ours_loaded_ref_ckpt_maxabs"] = max(float((ours_after[k] - ref_named[k]).abs().max()) for k in ref_named)
            acc2 = ours.test()
            m["acc_after_cross_load"] = float(acc2)
            _load = torch.load     # torch >= 2.6 defaults to weights_only=True; Dassl's load_checkpoint predates it
            torch.load = lambda *a, **k: _load(*a, **{"weights_only": False, **k})
            try:
                ref.load_model(d_ours, epoch=2)
            finally:
                torch.load = _load
            ref_after = {k: v.detach() for k, v in ref.model.named_parameters() if v.requires_grad}
            m["ref_loaded_our_ckpt_maxabs"] = max(float((ref_after[k] - ours_named[k]).abs().max()) for k in ref_named)
        # TRAINER.B200.DEVICE_INPUT: the same Dassl DataManager, handed the on-device input pipeline's transforms
        # (data_manager.py:64-85) -- and nothing else changes (without the flag no custom transform is passed)
        m["default_datamanager_kwargs"] = sorted(StubDataManager.last_kwargs)
        from yacs.config import CfgNode
        from psrc_b200.input_pipeline import DeviceTransform
        cfg_dev = make_cfg(trainer_name, d_ours)
        cfg_dev.defrost()
        cfg_dev.TRAINER.B200 = CfgNode({"DEVICE_INPUT": True, "DEVICE_INPUT_CANVAS": 96})
        cfg_dev.INPUT.TRANSFORMS = ("random_resized_crop", "random_flip", "normalize")     # the PromptSRC yamls' chain
        with R.cpu_patches(half_roundtrip=(trainer_name == "PromptSRC")), contextlib.redirect_stdout(quiet):
            dev_tr = build_trainer(cfg_dev)
        k = StubDataManager.last_kwargs
        tr_, te_ = k.get("custom_tfm_train"), k.get("custom_tfm_test")
        m["device_input_transforms"] = bool(isinstance(tr_, DeviceTransform) and isinstance(te_, DeviceTransform)
                                            and tr_.is_train and not te_.is_train and tr_.canvas == (96, 96)
                                            and tr_.size == 64 and te_.size == 64)
        m["device_input_loader_attrs"] = all(hasattr(dev_tr, a) for a in ("train_loader_x", "test_loader", "num_classes",
                                                                         "lab2cname", "dm"))
        m["device_input_norm"] = list(dev_tr.model.input_kwargs.get("mean", ()))
    return m


if __name__ == "__main__":
    name = sys.argv[1] if len(sys.argv) > 1 else "PromptSRC"
    torch.set_num_threads(min(8, os.cpu_count() or 1))
    print("DROPIN " + json.dumps(run(name)))
