"""2-rank NCCL parity of the class-sharded / data-parallel step -- runs where >= 2 GPUs are visible
(`gpurun --gpus 2 -- python -m pytest tests/test_multigpu_gpu.py -m gpu`), skipped on a one-GPU box.

Body = tests/multigpu_check.py under torchrun: every rank runs its image / class shard through NCCL (all-gather of
t_hat, reduce-scatter of dL/dt_hat, all-reduce of the flat prompt gradient + loss terms); rank 0 also runs the whole batch
alone and compares; with --golden the sharded step is also held against the full-batch fixture (BASELINE configs[2])."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "multigpu_check.py"), *extra]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    lines = [json.loads(ln) for ln in res.stdout.splitlines() if ln.startswith("{")]
    print("\n".join(json.dumps(ln) for ln in lines))          # shown by `pytest -rA` (profiles/r03_multigpu_pytest.log)
    assert res.returncode == 0, f"rc={res.returncode}\n{lines}\n{res.stderr[-3000:]}"
    return lines


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
def test_two_rank_step_matches_single_gpu_step(prec):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    lines = _run(["--classes", "1000", "--batch", "8", "--prec", prec], 29541)
    assert lines[0]["ok"] and lines[-1]["all_ranks_identical_params_and_ok"], lines


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
def test_two_rank_step_matches_cfg3_golden(prec):
    """Global batch 128 over 2 ranks (64 images + 500 classes each) against tests/golden/cfg3_promptsrc_ce.npz."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    lines = _run(["--golden", "cfg3_promptsrc_ce", "--prec", prec], 29543)
    assert lines[0]["ok"] and lines[-1]["all_ranks_identical_params_and_ok"], lines


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
def test_two_rank_captured_segments_match_eager_and_single_gpu(prec):
    """The step the trainers and bench.py run at world_size > 1: six segment graphs in one pool with NCCL between them,
    the batch arriving on the copy stream.  Three steps on three batches: bit-identical to the eager sharded launches
    on every rank, and within the sharding tolerance of the captured single-GPU step (multigpu_check.py --graphs)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    lines = _run(["--classes", "1000", "--batch", "8", "--prec", prec, "--graphs"], 29545)
    g = [ln for ln in lines if "graphs_ok" in ln]
    assert g and g[0]["graphs_ok"] and g[0]["graph_vs_eager_param_maxabs"] == 0.0, lines
    assert lines[-1]["all_ranks_identical_params_and_ok"], lines


@pytest.mark.parametrize("trainer,classes", [("cocoop", "100"), ("maple", "1000")])
def test_two_rank_captured_segments_cocoop_and_maple(trainer, classes):
    """Same check for the pure batch-sharded CoCoOp step (one graph + all-reduce + SGD graph, `whole`) and the MaPLe
    step (class-sharded text tower with the coupling functions' backward in the last segment)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    lines = _run(["--classes", classes, "--batch", "4", "--prec", "bf16", "--graphs", "--trainer", trainer], 29547)
    g = [ln for ln in lines if "graphs_ok" in ln]
    assert lines[0]["ok"] and g and g[0]["graphs_ok"], lines
    assert lines[-1]["all_ranks_identical_params_and_ok"], lines
