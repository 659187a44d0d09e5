"""GPU parity of every kernel family against plain PyTorch fp32 (run with -m gpu on the B200)."""
import pytest
import torch

from kernel_checks import GROUPS, run_group

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", list(GROUPS))
def test_kernel_group(name):
    if not torch.cuda.is_available():
        pytest.fail("CUDA device required: this package has no CPU fallback")
    res = run_group(name)
    bad = [r for r in res.rows if not r[3]]
    assert not bad, "\n".join(f"{n}: err={e:.3e} tol={t:.1e}" for n, e, t, _ in bad)
