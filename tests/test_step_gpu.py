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
