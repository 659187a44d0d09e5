"""GPU parity of the full PromptSRC training step against the reference's committed outputs."""
import pytest
import torch

from step_checks import CASES, check, run_promptsrc_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("name", list(CASES))
def test_promptsrc_step_matches_reference(name, prec):
    if not torch.cuda.is_available():
        pytest.fail("CUDA device required: this package has no CPU fallback")
    m = run_promptsrc_case(name, prec)
    bad = check(m, prec, name)
    assert not bad, "; ".join(bad) + f"\nmetrics: {m}"
