"""The drop-in under the REAL Dassl (reference checkout at /root/reference; skipped on the GPU box where it is absent).

`tests/dassl_dropin_check.py` is run in a fresh interpreter per trainer: Dassl's `build_trainer(cfg)` picks this
package's trainer from its own registry, `trainer.train()` runs Dassl's epoch loop, checkpointing and test(), and the
unmodified reference trainer runs beside it on the same stub DataManager and synthetic CLIP weights."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("PSRC_REFERENCE_ROOT", "/root/reference")


@pytest.mark.parametrize("trainer", ["PromptSRC", "CoOp", "CoCoOp"])   # MaPLe: the reference hard-codes ViT-B widths (F7)
def test_dropin_trains_and_tests_under_real_dassl_like_the_reference(trainer):
    if not os.path.isdir(os.path.join(REF, "Dassl.pytorch", "dassl")):
        pytest.skip("reference checkout not present (GPU box)")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "dassl_dropin_check.py"), trainer],
                       capture_output=True, text=True, timeout=900)
    lines = [l for l in p.stdout.splitlines() if l.startswith("DROPIN ")]
    assert p.returncode == 0 and lines, p.stdout[-2000:] + p.stderr[-4000:]
    m = json.loads(lines[-1][len("DROPIN "):])
    assert m["registry_class"].endswith(f"_b200.trainers.{trainer}"), m      # what `train.py --trainer X` would build
    assert m["model_name"] == m["ref_model_name"] and m["trainable_names_equal"], m
    assert m["n_batches"] == 4 and m["loss_maxabs"] < 2e-5 and m["param_maxabs"] < 1e-5, m
    assert m["test_logits_maxabs"] < 2e-4 and m["test_argmax_equal"] and m["acc_ours"] == m["acc_ref"], m
    assert all(m["checkpoint_files"]), m
    # frozen duplicate the reference stores inside prompt_learner (a second copy of the vision tower) is the only gap
    assert all(k.startswith("prompt_learner.ZS_image_encoder.") for k in m["ckpt_keys_missing_in_ours"]), m
    assert m["ours_loaded_ref_ckpt_maxabs"] == 0.0 and m["ref_loaded_our_ckpt_maxabs"] < 1e-5, m
    assert m["acc_after_cross_load"] == m["acc_ref"], m
    # the on-device input pipeline is opt-in: by default Dassl's DataManager is built exactly as the reference builds it
    assert m["default_datamanager_kwargs"] == [] and m["device_input_transforms"] and m["device_input_loader_attrs"], m
    assert len(m["device_input_norm"]) == 3, m
