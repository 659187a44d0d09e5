"""CPU tests of the host orchestration with the kernel layer replaced by shape-checking fakes
(no compute): buffer shapes, launch order, class sharding arithmetic, C-ABI symbol table."""
import ctypes
import os
import re

import pytest
import torch

from helpers import FakeKernels, load_golden, promptsrc_setup

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cabi_exports_every_declared_symbol():
    from psrc_b200 import cabi
    lib = cabi.load()  # raises if the .so is missing: the product has no fallback
    header = open(os.path.join(ROOT, "include", "psrc_b200.h")).read()
    declared = set(re.findall(r"\b(psrc_[a-z0-9_]+)\s*\(", header))
    declared -= {"psrc_stream_t"}
    assert declared, "no declarations parsed"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/psrc_b200.h but not exported"
    assert declared == set(cabi.SIGNATURES), declared ^ set(cabi.SIGNATURES)
    assert lib.psrc_version() >= 100
    assert isinstance(cabi.launch_count(), int)


def test_ops_refuse_cpu_tensors():
    from psrc_b200 import ops
    x = torch.zeros(4, 128)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.layernorm_fwd(x, torch.ones(128), torch.zeros(128), torch.empty_like(x))


@pytest.mark.parametrize("prec", ["bf16", "fp32"])
def test_promptsrc_step_launch_sequence(monkeypatch, prec):
    fk = FakeKernels()
    fk.install(monkeypatch)
    fx = load_golden("tiny_promptsrc")
    step, sd, fixed = promptsrc_setup("tiny", 3, fx["tokens"], "cpu", prec, batch=3)
    n_teacher = len(fk.calls)
    assert fk.calls.count("attn_fwd") == 3  # frozen teacher: 3 text layers
    step.forward_backward()
    calls = fk.calls[n_teacher:]
    # 3 towers forward (text, prompted vision, frozen vision) x 3 layers; backward on two of them
    assert calls.count("attn_fwd") == 9 and calls.count("attn_bwd") == 6
    assert calls.count("loss_fwd_bwd") == 1
    # prompt grads: 2 deep text + ctx + 2 deep vision + vpt
    assert calls.count("prompt_grad") == 6
    assert calls.count("insert_prompt") == 2 + 2 + 1
    assert step.Lt == int(torch.as_tensor(fx["tokens"]).argmax(-1).max()) + 1
    assert step.flat_grad.numel() == 4 * 128 * 6


def test_shard_range_partitions_classes():
    from psrc_b200.step import shard_range
    for n, w in ((1000, 8), (100, 3), (7, 8)):
        spans = [shard_range(n, w, r) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
