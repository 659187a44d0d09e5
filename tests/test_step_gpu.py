"""GPU parity of the full PromptSRC training step against the reference's committed outputs."""
import pytest
import torch

from step_checks import CASES, OTHER_CASES, check, check_generic, run_promptsrc_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("name", list(CASES))
def test_promptsrc_step_matches_reference(name, prec):
    if not torch.cuda.is_available():
        pytest.fail("CUDA device required: this package has no CPU fallback")
    m = run_promptsrc_case(name, prec)
    bad = check(m, prec, name)
    assert not bad, "; ".join(bad) + f"\nmetrics: {m}"


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("name", list(OTHER_CASES))
def test_maple_cocoop_step_matches_reference(name, prec):
    """MaPLe (coupled prompts, ViT-B/16) and CoCoOp (per-image text passes + meta-net) against the reference's outputs."""
    if not torch.cuda.is_available():
        pytest.fail("CUDA device required: this package has no CPU fallback")
    m = OTHER_CASES[name](name, prec)
    bad = check_generic(m, prec, name)
    assert not bad, "; ".join(bad) + f"\nmetrics: {m}"


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("name", ["tiny_promptsrc", "cfg1_promptsrc"])
def test_promptsrc_step_dense_text_positions_matches_reference(name, prec):
    """All 77 text positions instead of the truncation to max(EOT) + 1 (exact under the causal mask): the
    text tower then runs the tcgen05 attention kernels (L = 77, causal) and -- in the 16-bit mode -- the mixed layout
    of engine.TowerDtypes (fp16 GEMM operands, bf16 q/k/v/O).  Same fixture, same gates."""
    if not torch.cuda.is_available():
        pytest.fail("CUDA device required: this package has no CPU fallback")
    m = run_promptsrc_case(name, prec, truncate=False)
    bad = check(m, prec, name)
    assert not bad, "; ".join(bad) + f"\nmetrics: {m}"


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("B,name_lens", [(1, [1, 2]), (2, [30]), (1, [1, 1, 1, 1, 1, 1, 1])])
def test_promptsrc_step_edge_sizes_match_oracle(B, name_lens, prec):
    """One image; two classes / ONE class (softmax over a single logit: CE = 0, KL = 0) / a 30-token class name (text
    length 37: the tcgen05 causal attention instead of the short-sequence kernels); seven classes for one image."""
    from step_checks import TOL, run_tiny_oracle_case
    if not torch.cuda.is_available():
        pytest.fail("CUDA device required: this package has no CPU fallback")
    m = run_tiny_oracle_case(B, name_lens, prec)
    t = TOL[prec]
    assert m["logits_maxabs"] <= t["logits"] and m["loss_rel"] <= t["loss"], m
    assert m["gradcos_worst"] >= t["cos"], m
    if prec == "fp32":
        assert m["argmax_equal"], m
