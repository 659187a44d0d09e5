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
