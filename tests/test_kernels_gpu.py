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


def test_empty_and_single_row_inputs_through_the_c_abi():
    """Edge cases the host code can produce (an empty class shard on a padded rank, a batch of one): zero-sized inputs
    return without a launch, single rows / single sequences go through every kernel family."""
    if not torch.cuda.is_available():
        pytest.fail("CUDA device required: this package has no CPU fallback")
    from psrc_b200 import cabi, ops
    dev, bf = "cuda", torch.bfloat16
    n0 = cabi.launch_count()
    ops.gemm(torch.empty(0, 64, device=dev, dtype=bf), torch.randn(32, 64, device=dev).to(bf), torch.empty(0, 32, device=dev))
    ops.layernorm_fwd(torch.empty(0, 128, device=dev), torch.ones(128, device=dev), torch.zeros(128, device=dev),
                      torch.empty(0, 128, device=dev, dtype=bf))
    ops.attn_fwd(torch.empty(0, 3 * 64, device=dev, dtype=bf), torch.empty(0, 64, device=dev, dtype=bf),
                 torch.empty(0, 1, 7, device=dev), 0, 7, 1, True)
    ops.attn_row_fwd(torch.empty(0, 3 * 64, device=dev, dtype=bf), torch.empty(0, 64, device=dev, dtype=bf),
                     torch.empty(0, 1, 7, device=dev), 0, 7, 1, False)
    assert cabi.launch_count() == n0                       # nothing was launched for empty inputs
    # one row / one sequence of one token
    a, w = torch.randn(1, 64, device=dev).to(bf), torch.randn(8, 64, device=dev).to(bf)
    out = torch.empty(1, 8, device=dev)
    ops.gemm(a, w, out)
    assert torch.allclose(out, a.float() @ w.float().t(), atol=1e-2, rtol=1e-2)
    qkv = torch.randn(1, 3 * 64, device=dev).to(bf)
    o, lse = torch.empty(1, 64, device=dev, dtype=bf), torch.empty(1, 1, 1, device=dev)
    ops.attn_fwd(qkv, o, lse, 1, 1, 1, True)
    assert torch.allclose(o.float(), qkv[:, 128:].float(), atol=1e-2)      # softmax over one key = that key's value
    o2 = torch.zeros(1, 64, device=dev, dtype=bf)
    ops.attn_row_fwd(qkv, o2, lse, 1, 1, 1, False)
    assert torch.allclose(o2.float(), qkv[:, 128:].float(), atol=1e-2)
    torch.cuda.synchronize()
