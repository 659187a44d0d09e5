"""psrc_loss_fwd_bwd_noalloc: the loss entry point for embedders that cannot allocate device memory.

Some hosts of the C-ABI library only own the buffers they are handed: a plugin inside a runtime that has already
reserved the whole GPU, code running while another thread captures a CUDA graph (cudaMalloc is illegal then), a
latency-critical loop that must not touch the allocator's locks.  They cannot pass the `row_scratch` workspace that
psrc_loss_fwd_bwd needs, so the no-alloc variant keeps that workspace in a fixed `__device__` array inside the library
(at most 4096 logits rows per call; include/psrc_b200.h states the contract).

This file checks, through ctypes -> C ABI only, that the variant returns what the ordinary entry point returns, twice in
a row with different shapes and loss terms (the second call starts on the first call's leftovers in the shared array),
that it refuses a batch beyond its limit with a message naming the limit, and that torch's allocator saw no request
while it ran.  The file name sorts last on purpose: `pytest -x` stops at the first failure and this test must not hide
the parity tests; it passes both alone in a fresh process (its first call then does the per-device address lookup) and
after the rest of the suite.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ROWS = 4096     # include/psrc_b200.h: limit of psrc_loss_fwd_bwd_noalloc
TOL = 2e-5          # the loss group's tolerance in tests/kernel_checks.py:check_loss


def _rel(a, b):
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


def _case(B, Cn, E, focal, with_zs, with_l1, seed):
    """Seeded inputs of one loss call (device tensors)."""
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(seed)
    n = torch.nn.functional.normalize
    t = dict(
        logits=torch.randn(B, Cn, device=dev, generator=g) * 2,
        labels=torch.randint(0, Cn, (B,), device=dev, generator=g).int(),
        zs=(torch.randn(B, Cn, device=dev, generator=g) * 2) if with_zs else None,
        alpha=(torch.rand(Cn, device=dev, generator=g) * 3) if focal else None,
    )
    if with_l1:
        t.update(t_hat=n(torch.randn(Cn, E, device=dev, generator=g), dim=-1),
                 t_zs=n(torch.randn(Cn, E, device=dev, generator=g), dim=-1),
                 f_hat=n(torch.randn(B, E, device=dev, generator=g), dim=-1),
                 f_zs=n(torch.randn(B, E, device=dev, generator=g), dim=-1))
    return t


def _call(entry, t, scratch):
    """One call of `entry` (a bound C-ABI function) on the inputs `t`; returns (rc, outputs)."""
    from psrc_b200 import cabi
    B, Cn = t["logits"].shape
    with_l1 = "t_hat" in t
    E = t["t_hat"].shape[1] if with_l1 else 0
    dev = "cuda"
    # outputs start as NaN so that a slot the call forgot to write cannot pass the comparison
    out = dict(d_logits=torch.full((B, Cn), float("nan"), device=dev), losses=torch.full((4,), float("nan"), device=dev))
    if with_l1:
        out.update(d_that=torch.full((Cn, E), float("nan"), device=dev), d_fhat=torch.full((B, E), float("nan"), device=dev))
    a = cabi.LossArgs()
    a.B, a.C, a.E, a.Ct = B, Cn, E, (Cn if with_l1 else 0)
    a.B_global, a.Ct_global = B, a.Ct
    a.logits, a.zs_logits, a.labels = cabi.ptr(t["logits"]), cabi.ptr(t["zs"]), cabi.ptr(t["labels"])
    a.focal_alpha, a.focal_gamma = cabi.ptr(t["alpha"]), 2.0
    a.t_hat, a.t_zs = cabi.ptr(t.get("t_hat")), cabi.ptr(t.get("t_zs"))
    a.f_hat, a.f_zs = cabi.ptr(t.get("f_hat")), cabi.ptr(t.get("f_zs"))
    a.w_kl, a.w_text, a.w_image = 1.0, 25.0, 10.0
    a.d_logits, a.d_that, a.d_fhat = cabi.ptr(out["d_logits"]), cabi.ptr(out.get("d_that")), cabi.ptr(out.get("d_fhat"))
    a.losses = cabi.ptr(out["losses"])
    a.row_scratch = cabi.ptr(scratch)           # None -> NULL: the no-alloc variant gets no workspace at all
    rc = entry(C.byref(a), cabi.stream())
    return rc, out


def test_noalloc_loss_matches_the_allocating_entry_point_on_two_calls_in_a_row():
    if not torch.cuda.is_available():
        pytest.fail("CUDA device required: this package has no CPU fallback")
    from psrc_b200 import cabi
    lib = cabi.load()
    torch.cuda.set_device(0)
    # first call: every term, focal CE, the benchmark's batch; second call: a smaller, ragged batch with CE only --
    # it runs on top of the 128-row partial sums and both L1 partial blocks that the first call left in the array
    cases = [_case(128, 1000, 512, True, True, True, seed=11), _case(7, 37, 0, False, False, False, seed=12)]
    expected = []
    for t in cases:
        B = t["logits"].shape[0]
        rc, ref = _call(lib.psrc_loss_fwd_bwd, t, torch.empty(4 * B + 2 * 296, device="cuda"))
        cabi.check(rc, "psrc_loss_fwd_bwd")
        expected.append(ref)
    torch.cuda.synchronize()

    requests_before = torch.cuda.memory_stats()["allocation.all.allocated"]
    got = []
    for t in cases:                               # two calls in a row, same stream, nothing between them
        rc, out = _call(lib.psrc_loss_fwd_bwd_noalloc, t, None)
        cabi.check(rc, "psrc_loss_fwd_bwd_noalloc")
        got.append(out)
    torch.cuda.synchronize()
    # the only torch allocations in that loop are the output tensors made by _call (2 + 4 tensors for
    # these two cases, +0 for a workspace): the entry point itself asked torch's allocator for nothing
    assert torch.cuda.memory_stats()["allocation.all.allocated"] - requests_before == 4 + 2

    for i, (ref, out) in enumerate(zip(expected, got)):
        assert set(ref) == set(out)
        for name in ref:
            assert torch.isfinite(out[name]).all(), f"call {i}: {name} has unwritten / non-finite entries"
            err = _rel(out[name], ref[name])
            assert err <= TOL, f"call {i}: {name} differs from psrc_loss_fwd_bwd by {err:.3e} (tolerance {TOL:.0e})"
    # CE-only call: the KL and L1 slots are reported as exact zeros, not as leftovers of the first call
    assert got[1]["losses"][1:].tolist() == [0.0, 0.0, 0.0]


def test_noalloc_loss_refuses_a_batch_beyond_its_builtin_workspace():
    if not torch.cuda.is_available():
        pytest.fail("CUDA device required: this package has no CPU fallback")
    from psrc_b200 import cabi
    lib = cabi.load()
    torch.cuda.set_device(0)
    n0 = cabi.launch_count()
    rc, _ = _call(lib.psrc_loss_fwd_bwd_noalloc, _case(MAX_ROWS + 1, 8, 0, False, False, False, seed=13), None)
    assert rc == 1                                # PSRC_ERR_ARG
    msg = cabi.last_error()
    assert str(MAX_ROWS) in msg and "psrc_loss_fwd_bwd_noalloc" in msg, msg
    assert cabi.launch_count() == n0              # rejected before any kernel was launched
    # the limit itself is accepted and agrees with the ordinary entry point
    t = _case(MAX_ROWS, 8, 0, False, True, False, seed=14)
    rc, ref = _call(lib.psrc_loss_fwd_bwd, t, torch.empty(4 * MAX_ROWS + 2 * 296, device="cuda"))
    cabi.check(rc, "psrc_loss_fwd_bwd")
    rc, out = _call(lib.psrc_loss_fwd_bwd_noalloc, t, None)
    cabi.check(rc, "psrc_loss_fwd_bwd_noalloc")
    torch.cuda.synchronize()
    for name in ref:
        assert _rel(out[name], ref[name]) <= TOL, name
