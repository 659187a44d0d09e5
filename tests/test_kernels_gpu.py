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


@pytest.mark.parametrize("L", [280, 300])
def test_class_token_only_last_block_matches_full_block_also_beyond_the_one_query_kernel_limit(L):
    """A tower whose consumer reads the class-token row only runs its last block on that row (one-query attention,
    L <= 288).  Longer sequences (ViT-L/14@336px has 577 tokens, `clip._MODELS`) must fall back to the full-row block
    instead of failing in psrc_attn_row_*: both settings of `cls_only_last_mlp` give the same row 0 and the same
    input-gradient at L = 280 (one-query kernels) and at L = 300 (fallback)."""
    if not torch.cuda.is_available():
        pytest.fail("CUDA device required: this package has no CPU fallback")
    from psrc_b200.engine import Tower, TowerWeights
    dev, D, H, n_seq, nl = torch.device("cuda"), 64, 1, 3, 2
    g = torch.Generator().manual_seed(3)
    r = lambda *s, sc=0.05: torch.randn(*s, generator=g) * sc  # noqa: E731
    sd = {}
    for i in range(nl):
        p = f"t.resblocks.{i}."
        sd.update({p + "ln_1.weight": 1 + r(D), p + "ln_1.bias": r(D), p + "ln_2.weight": 1 + r(D), p + "ln_2.bias": r(D),
                   p + "attn.in_proj_weight": r(3 * D, D, sc=0.1), p + "attn.in_proj_bias": r(3 * D),
                   p + "attn.out_proj.weight": r(D, D, sc=0.1), p + "attn.out_proj.bias": r(D),
                   p + "mlp.c_fc.weight": r(4 * D, D, sc=0.1), p + "mlp.c_fc.bias": r(4 * D),
                   p + "mlp.c_proj.weight": r(D, 4 * D, sc=0.1), p + "mlp.c_proj.bias": r(D)})
    w = TowerWeights(sd, "t.", nl, dev, torch.bfloat16, need_backward=True)
    x0 = r(n_seq * L, D, sc=1.0).to(dev)
    d_cls = r(n_seq, D, sc=1.0).to(dev)
    outs = []
    for flag in (False, True):
        t = Tower(w, n_seq, L, D, H, False, True, dev)
        t.cls_only_last_mlp = flag
        t.x0.copy_(x0)
        y = t.forward().view(n_seq, L, D)[:, 0].clone()
        t.dx.zero_()
        t.dx.view(n_seq, L, D)[:, 0] = d_cls          # the gradient enters on the class-token rows only
        t.dxa.copy_(t.dx)
        dx = t.backward().clone()
        torch.cuda.synchronize()
        outs.append((y, dx))
    (y0, dx0), (y1, dx1) = outs
    assert torch.isfinite(y1).all() and torch.isfinite(dx1).all()
    if L > 288:     # same launches either way
        assert torch.equal(y0, y1) and torch.equal(dx0, dx1)
    else:           # one-query kernels vs the full attention kernels: bf16 rounding of different summation orders
        assert float((y0 - y1).abs().max()) <= 2e-2 * float(y0.abs().max())
        assert float(torch.nn.functional.cosine_similarity(dx0.flatten(), dx1.flatten(), dim=0)) >= 0.999
