"""GPU parity of the Dassl-facing drop-in (trainers.PromptSRC / CoOp, CustomCLIP) against the oracle."""
import pytest
import torch

import trainer_checks as TC

pytestmark = pytest.mark.gpu


def _need_gpu():
    if not torch.cuda.is_available():
        pytest.fail("CUDA device required: this package has no CPU fallback")


def test_promptsrc_trainer_fused_steps_match_oracle():
    _need_gpu()
    m = TC.check_promptsrc_trainer("fp32")
    assert m["graph_captured"], m
    assert max(m["loss_abs_0"], m["loss_abs_1"], m["loss_abs_2"]) < 5e-5, m
    assert m["param_after3_maxabs"] < 1e-6 and m["update_cos"] > 0.9999, m
    assert abs(m["lr_after_epoch"] - 0.0025) < 1e-12 and abs(m["lr_device"] - 0.0025) < 1e-9, m


def test_customclip_forward_autograd_bridge():
    _need_gpu()
    m = TC.check_bridge("fp32")
    assert m["loss_abs"] < 5e-5 and m["logits_maxabs"] < 1e-4, m
    assert m["gradcos_worst"] > 0.9999 and m["param_after_maxabs"] < 1e-6, m
    assert m["n_trainable"] == 2 * TC.DEPTH, m      # ctx + VPT + (depth-1) deep prompts per tower


def test_eval_logits_cached_text_and_checkpoint_roundtrip():
    _need_gpu()
    m = TC.check_eval_and_checkpoint("fp32")
    assert m["ckpt_restore_maxabs"] == 0.0 and not m["missing_state_keys"], m
    assert m["eval_logits_maxabs"] < 1e-4 and m["eval_argmax_equal"] and m["eval_shape_ok"], m
    assert len(m["slim_keys"]) == 2 * TC.DEPTH and m["slim_restore_maxabs"] == 0.0 and m["state_dict_method_restored"], m
    assert m["slim_bytes"] < 4 * m["full_state_numel"] / 20, m        # a small fraction of the full state


def test_promptsrc_gpa_three_epochs_matches_reference_aggregation():
    _need_gpu()
    m = TC.check_gpa("fp32")
    assert m["step_counter"] == 4 and m["loss_abs_max"] < 5e-5, m
    assert m["gpa_param_maxabs"] < 1e-5 and m["last_vs_gpa_maxabs"] > 1e-3, m    # GPA != the last iterate
    assert m["eval_logits_maxabs"] < 1e-4, m


def test_coop_eval_text_cache_follows_prompt_learner_load_state_dict():
    _need_gpu()
    m = TC.check_coop_eval_follows_prompt_learner_load()
    assert m["changed_maxabs"] > 1e-3 and m["stale_maxabs"] == 0.0, m


@pytest.mark.parametrize("loss_type", ["ce", "focal"])
def test_coop_trainer_step(loss_type):
    _need_gpu()
    m = TC.check_coop("fp32", loss_type)
    assert m["loss_abs"] < 5e-5 and m["gradcos"] > 0.9999 and m["ctx_after_maxabs"] < 1e-6, m
    assert m["registered"] == ["prompt_learner"], m
    if loss_type == "ce":
        assert m["acc_abs"] < 1e-4, m


def test_trainer_bf16_mode_runs():
    _need_gpu()
    m = TC.check_promptsrc_trainer("fp16")          # cfg PREC fp16 -> 16-bit tensor-core mode
    assert max(m["loss_abs_0"], m["loss_abs_1"], m["loss_abs_2"]) < 5e-2, m
    assert m["update_cos"] > 0.99, m


@pytest.mark.parametrize("focal", [False, True])
def test_maple_trainer_steps_and_eval(focal):
    _need_gpu()
    m = TC.check_maple("fp32", focal)
    assert max(m["loss_abs_0"], m["loss_abs_1"], m["loss_abs_2"]) < 5e-5, m
    assert m["param_after3_maxabs"] < 2e-6 and m["graph_captured"], m
    assert m["registered"] == ["MultiModalPromptLearner"] and not m["missing_state_keys"], m
    assert m["n_trainable"] == 1 + 2 + 3 * (TC.DEPTH - 1), m      # ctx, proj.{w,b}, per deeper layer: text + Linear
    assert m["eval_logits_maxabs"] < 1e-4 and m["eval_argmax_equal"], m


@pytest.mark.parametrize("focal", [False, True])
def test_cocoop_trainer_steps_and_eval(focal):
    _need_gpu()
    m = TC.check_cocoop("fp32", focal)
    assert max(m["loss_abs_0"], m["loss_abs_1"], m["loss_abs_2"]) < 5e-5, m
    assert m["param_after3_maxabs"] < 2e-6 and m["graph_captured"], m
    assert m["registered"] == ["prompt_learner"] and m["n_trainable"] == 5, m
    assert m["eval_logits_maxabs"] < 1e-4 and m["eval_argmax_equal"] and m["eval_shape_ok"], m


def test_maple_cocoop_bf16_mode_runs():
    _need_gpu()
    for fn in (TC.check_maple, TC.check_cocoop):
        m = fn("fp16")
        assert max(m["loss_abs_0"], m["loss_abs_1"], m["loss_abs_2"]) < 5e-2, m


def test_frozen_teacher_precompute_single_and_ensemble():
    _need_gpu()
    m = TC.check_teacher_ensemble()
    assert m["ensemble_rel"] < 2e-5 and m["single_rel"] < 2e-5, m


def test_text_prefix_enqueued_ahead_is_exact_and_invalidated_by_outside_edits():
    _need_gpu()
    m = TC.check_text_ahead()
    assert m["ahead_flag_True"] and not m["ahead_flag_False"], m
    assert m["param_maxabs"] == 0.0 and m["loss_maxabs"] == 0.0, m
    assert m["loss_jump_after_edit"] > 0.0, m
    assert m["staged_steps_used"] and m["lookahead_losses_distinct"], m
    assert m["lookahead_param_maxabs"] == 0.0 and m["lookahead_loss_maxabs"] == 0.0, m


@pytest.mark.parametrize("name", ["PromptSRC", "CoCoOp"])
def test_trainer_fed_by_the_device_input_pipeline_equals_the_fp32_fed_trainer(name):
    """SURVEY 8f-4 through the drop-in: batches collated from `input_pipeline.DeviceTransform` items (what Dassl's
    DataManager delivers under TRAINER.B200.DEVICE_INPUT) go through `forward_backward` -- uint8 windows over PCIe,
    `psrc_crop_resize_normalize`, graph replay, look-ahead staging of the next batch -- and through `model_inference`
    with a different (test) batch size.  The fp32-fed twin receives the very images that kernel produces, so losses,
    parameters and test logits must agree to the last bit; the kernel's parity with Dassl's PIL chain is
    test_input_pipeline_gpu.py's subject."""
    _need_gpu()
    from torch.utils.data import default_collate
    from input_ref import synthetic_image
    from psrc_b200 import input_pipeline as IP, ops
    from psrc_b200.trainers import LookaheadLoader
    R, canvas = 64, (96, 96)
    mean, std = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)

    def u8_batch(n, seed, is_train):
        torch.manual_seed(seed)
        tf = IP.DeviceTransform(size=R, is_train=is_train, canvas=canvas)
        items = [tf(synthetic_image(70 + 9 * k, 120 - 7 * k, seed + k)) for k in range(n)]       # some exceed the canvas
        lab = torch.randint(0, len(TC.NAME_LENS), (n,), generator=torch.Generator().manual_seed(seed))
        return {"img": default_collate(items), "label": lab}

    def as_fp32(b):
        it, out = b["img"], torch.empty(b["label"].shape[0], 3, R, R, device="cuda")
        ops.crop_resize_normalize(it["u8"].cuda(), it["geom_f"].cuda(), it["geom_i"].cuda(), out, 1, 2,
                                  IP.max_taps(it["geom_f"].numpy(), 1), mean, std)
        torch.cuda.synchronize()
        return {"img": out.cpu(), "label": b["label"]}

    train = [u8_batch(TC.B, 10 * s, True) for s in range(4)]
    test = u8_batch(5, 99, False)
    out = {}
    for feed in ("u8", "fp32"):
        torch.manual_seed(1234)                            # CoCoOp's meta-net is randomly initialised at build time
        tr, cfg, _ = TC.make_trainer(name, "fp32")
        tr.model.input_kwargs = {"interpolation": "bicubic", "mean": mean, "std": std}
        batches = train if feed == "u8" else [as_fp32(b) for b in train]
        losses = []
        for b in LookaheadLoader(batches):                 # step 2 onwards replays the captured graph
            tr.batch_idx = 0
            losses.append(tr.forward_backward(b)["loss"])
        tr.set_model_mode("eval")
        with torch.no_grad():
            tb = test if feed == "u8" else as_fp32(test)
            inp, _ = tr.parse_batch_test({"img": tb["img"], "label": tb["label"]})
            logits = tr.model_inference(inp).float().cpu()
        out[feed] = (losses, tr.model._step.flat_param.clone(), logits,
                     getattr(tr.model._step, "_graph", None) is not None)
    (l_u, p_u, g_u, cap_u), (l_f, p_f, g_f, _) = out["u8"], out["fp32"]
    assert cap_u, "the u8-fed trainer must reach the captured-graph path"
    assert l_u == l_f, (l_u, l_f)
    assert torch.equal(p_u, p_f) and torch.equal(g_u, g_f)
    assert g_u.shape == (5, len(TC.NAME_LENS)) and all(torch.isfinite(torch.tensor(l_u)))
