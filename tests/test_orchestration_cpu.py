"""CPU tests of the PRODUCT's host orchestration with the kernel layer swapped for plain-torch restatements
(tests/torch_kernels.py): single process against the reference's golden outputs, and the class-sharded /
data-parallel step under world_size-2 `gloo` against the single-process step."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import COCOOP_REF, cocoop_setup, coop_setup, cos, load_golden, maple_ref_name, maple_setup, promptsrc_setup, ref_name
from torch_kernels import TorchKernels

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))


class _Patch:
    """monkeypatch-like shim for subprocesses."""

    def setattr(self, obj, name, value, raising=True):
        setattr(obj, name, value)


def test_step_orchestration_matches_reference_golden(monkeypatch):
    """engine.py + step.py (launch order, prompt insertion / gradient routing, head + loss plumbing) driven with
    torch math reproduce the unmodified reference's outputs for the tiny PromptSRC case."""
    TorchKernels().install(monkeypatch)
    fx = load_golden("tiny_promptsrc")
    B = fx["label"].shape[0]
    from psrc_b200 import synthetic as S
    step, sd, fixed = promptsrc_setup("tiny", 3, fx["tokens"], "cpu", "fp32", batch=B)
    image, label = S.synthetic_batch(B, fx["tokens"].shape[0], 64, seed=1)
    step.image.copy_(image)
    step.labels.copy_(label.int())
    step.forward_backward()
    g = lambda k: torch.from_numpy(fx[k])  # noqa: E731
    assert float((fixed - g("fixed_embeddings")).norm() / g("fixed_embeddings").norm()) < 2e-5
    assert float((step.logits - g("logits")).abs().max()) < 1e-4
    assert float((step.zs_logits - g("zs_logits")).abs().max()) < 1e-4
    assert abs(float(step.losses.sum()) - float(fx["loss"])) < 2e-5
    assert np.array_equal(step.logits.argmax(1).numpy().astype(np.int32), fx["argmax"])
    for k in step.params:
        assert cos(step.grads[k], fx["grad/" + ref_name(k)]) > 0.99999, k


@pytest.mark.parametrize("name,focal,position", [("tiny_coop_ce", False, "end"), ("tiny_coop_focal", True, "end"),
                                                 ("tiny_coop_middle", False, "middle"),
                                                 ("tiny_coop_front", True, "front")])
def test_coop_step_orchestration_matches_reference_golden(monkeypatch, name, focal, position):
    """CoOp with every CLASS_TOKEN_POSITION (coop.py:267-293; middle / front = per-class row permutation of the
    'end' layout) against the unmodified reference's loss, context gradient and context after one SGD step."""
    TorchKernels().install(monkeypatch)
    from psrc_b200 import synthetic as S
    fx = load_golden(name)
    B, C = fx["label"].shape[0], fx["tokens"].shape[0]
    alpha = S.focal_alpha(fx["shots"].tolist(), C) if focal else None
    step, sd = coop_setup("tiny", fx["tokens"], "cpu", "fp32", B, fx["ctx0"], focal_alpha=alpha, position=position)
    image, label = S.synthetic_batch(B, C, 64, seed=2)
    assert np.array_equal(label.numpy().astype(np.int32), fx["label"])
    step.image.copy_(image)
    step.labels.copy_(label.int())
    step.forward_backward()
    assert float((step.logits - torch.from_numpy(fx["logits"])).abs().max()) < 1e-4
    assert abs(float(step.losses.sum()) - float(fx["loss"])) < 2e-5
    assert cos(step.grads["ctx"], fx["grad/prompt_learner.ctx"]) > 0.99999
    step.init_optimizer(lr=0.002)
    step.sgd()
    assert float((step.params["ctx"] - torch.from_numpy(fx["after/prompt_learner.ctx"])).abs().max()) < 1e-6


@pytest.mark.parametrize("name,focal", [("tiny_cocoop_ce", False), ("tiny_cocoop_focal", True)])
def test_cocoop_step_orchestration_matches_reference_golden(monkeypatch, name, focal):
    """CoCoOp: meta-net -> per-image context shift -> B x C text sequences -> per-image logits, and the backward into
    ctx + meta-net, against the unmodified reference's outputs (cocoop.py:173-198, :235-260)."""
    TorchKernels().install(monkeypatch)
    from psrc_b200 import synthetic as S
    fx = load_golden(name)
    B, C = fx["label"].shape[0], fx["tokens"].shape[0]
    alpha = S.focal_alpha(fx["shots"].tolist(), C) if focal else None
    step, sd, _ = cocoop_setup("tiny", fx["tokens"], "cpu", "fp32", batch=B, focal_alpha=alpha)
    for k in step.params:
        assert float((step.params[k] - torch.from_numpy(fx["init/" + COCOOP_REF[k]])).abs().max()) == 0.0, k
    image, label = S.synthetic_batch(B, C, 64, seed=4)
    assert np.array_equal(label.numpy().astype(np.int32), fx["label"])
    step.image.copy_(image)
    step.labels.copy_(label.int())
    step.forward_backward()
    assert float((step.logits - torch.from_numpy(fx["logits"])).abs().max()) < 1e-4
    assert abs(float(step.losses.sum()) - float(fx["loss"])) < 2e-5
    for k in step.params:
        assert cos(step.grads[k], fx["grad/" + COCOOP_REF[k]]) > 0.99999, k
    step.init_optimizer(lr=0.002)
    step.sgd()
    for k in step.params:
        assert float((step.params[k] - torch.from_numpy(fx["after/" + COCOOP_REF[k]])).abs().max()) < 1e-6, k


@pytest.mark.parametrize("focal", [False, True])
def test_maple_step_orchestration_matches_oracle(monkeypatch, focal):
    """MaPLe on the tiny architecture: coupled vision prompts (one batched Linear launch), prompt-only backward through
    both towers and the coupling Linears, against the oracle restatement (pinned on the reference at ViT-B widths,
    tests/golden/vitb_maple_*.npz -- the reference hard-codes 768/512)."""
    import clip_oracle as O
    TorchKernels().install(monkeypatch)
    from psrc_b200 import synthetic as S
    arch = S.ARCHS["tiny"]
    depth, n_ctx, B = 3, 2, 3
    tokens = S.synthetic_tokens(10, name_lens=[1, 2, 3, 8, 1, 1, 2, 2, 1, 2])
    C = tokens.shape[0]
    alpha = S.focal_alpha(list(range(1, C + 1)), C) if focal else None
    step, sd, mp_ = maple_setup("tiny", depth, tokens, "cpu", "fp32", batch=B, focal_alpha=alpha)
    image, label = S.synthetic_batch(B, C, arch.image_resolution, seed=3)
    step.image.copy_(image)
    step.labels.copy_(label.int())
    step.forward_backward()
    P = {k: step.params[k].detach().clone().requires_grad_(True) for k in step.params}
    o = O.maple_step(sd, torch.as_tensor(tokens), image, label, P["ctx"], [P[f"text.{i}"] for i in range(1, depth)],
                     (P["couple.w.0"], P["couple.b.0"]),
                     [(P[f"couple.w.{i}"], P[f"couple.b.{i}"]) for i in range(1, depth)], n_ctx, arch.vision_heads,
                     arch.transformer_heads, focal_alpha=torch.tensor(alpha) if focal else None, use_focal=focal)
    o["loss"].backward()
    assert float((step.logits - o["logits"].detach()).abs().max()) < 1e-4
    assert abs(float(step.losses.sum()) - float(o["loss"])) < 2e-5
    for k in step.params:
        assert cos(step.grads[k], P[k].grad) > 0.99999, k


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _sharded_worker(rank, world, port, tokens, B_global, focal, out, overlapped=False):
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        TorchKernels().install(_Patch())
        from psrc_b200 import synthetic as S
        C = tokens.shape[0]
        alpha = S.focal_alpha(list(range(1, C + 1)), C) if focal else None
        image, label = S.synthetic_batch(B_global, C, 64, seed=1)
        Bl = B_global // world
        step, sd, fixed = promptsrc_setup("tiny", 3, tokens, "cpu", "fp32", batch=Bl, world=world, rank=rank,
                                          pg=dist.group.WORLD, focal_alpha=alpha)
        step.image.copy_(image[rank * Bl:(rank + 1) * Bl])
        step.labels.copy_(label[rank * Bl:(rank + 1) * Bl].int())
        step.init_optimizer()
        if overlapped:          # the segment order of the multi-GPU fast path: asynchronous collectives between segments
            step.train_step()
        else:
            step.forward_backward()
            step.sgd()
        res = {"grad": step.flat_grad.clone(), "losses": step.losses.clone(), "param": step.flat_param.clone(),
               "logits": step.logits.clone(), "t_hat": step.t_hat.clone(), "classes": (step.c0, step.c1)}
        if rank == 0:  # single-process step on the global batch, same process (same torch math)
            ref, _, _ = promptsrc_setup("tiny", 3, tokens, "cpu", "fp32", batch=B_global, sd=sd, fixed=fixed,
                                        focal_alpha=alpha)
            ref.image.copy_(image)
            ref.labels.copy_(label.int())
            ref.forward_backward()
            ref.init_optimizer()
            ref.sgd()
            res["ref"] = {"grad": ref.flat_grad.clone(), "losses": ref.losses.clone(), "param": ref.flat_param.clone(),
                          "logits": ref.logits.clone(), "t_hat": ref.t_hat.clone()}
        out[rank] = res
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("focal,n_cls,overlapped", [(False, 10, False), (True, 10, False), (False, 9, False),
                                                    (True, 10, True), (False, 9, True)])
def test_class_sharded_data_parallel_step_world2_gloo(focal, n_cls, overlapped):
    """Images data-parallel, class prompts sharded by class: all-gather of t_hat, reduce-scatter of dL/dt_hat,
    all-reduce of the flat prompt gradient and of the loss terms give the single-device step (same gradients,
    same losses, same parameters after SGD) on every rank."""
    from psrc_b200 import synthetic as S
    # n_cls = 9: not a multiple of the world size -> the last shard carries one padding class (equal chunks for the
    # collectives; it never reaches the logits, gets a zero gradient and is excluded from the L1-text term)
    tokens = S.synthetic_tokens(n_cls, name_lens=[1, 2, 3, 8, 1, 1, 2, 2, 1, 2][:n_cls])
    world, B_global = 2, 4
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_sharded_worker, args=(world, _free_port(), tokens, B_global, focal, out, overlapped), nprocs=world,
             join=True)
    ref = out[0]["ref"]
    assert out[0]["classes"] == (0, 5) and out[1]["classes"] == (5, 10)
    for r in range(world):
        o = out[r]
        assert float((o["t_hat"][:n_cls] - ref["t_hat"]).abs().max()) < 1e-6    # all-gathered text features
        Bl = B_global // world
        assert float((o["logits"] - ref["logits"][r * Bl:(r + 1) * Bl]).abs().max()) < 1e-5
        assert float((o["losses"] - ref["losses"]).abs().max()) < 1e-5, (o["losses"], ref["losses"])
        assert cos(o["grad"], ref["grad"]) > 0.999999
        # per-sequence prompt gradients pass through fp16 (autograd of `.half()`, model.py:238): a different fp32
        # summation order may flip single fp16 roundings (one ulp = 2^-14 at 0.06..0.12)
        assert float((o["grad"] - ref["grad"]).abs().max()) < 1e-4 * float(ref["grad"].abs().max())
        assert float((o["param"] - ref["param"]).abs().max()) < 0.0025 * 2e-4 * float(ref["grad"].abs().max()) + 1e-7
    assert float((out[0]["param"] - out[1]["param"]).abs().max()) == 0.0         # ranks stay bit-identical


def _sharded_worker_other(rank, world, port, kind, tokens, B_global, out, overlapped=False):
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        TorchKernels().install(_Patch())
        from psrc_b200 import synthetic as S
        C = tokens.shape[0]
        alpha = S.focal_alpha(list(range(1, C + 1)), C)
        image, label = S.synthetic_batch(B_global, C, 64, seed=1)
        Bl = B_global // world

        def build(batch, w, r, pg):
            if kind == "maple":
                return maple_setup("tiny", 3, tokens, "cpu", "fp32", batch=batch, world=w, rank=r, pg=pg,
                                   focal_alpha=alpha)[0]
            return cocoop_setup("tiny", tokens, "cpu", "fp32", batch=batch, world=w, rank=r, pg=pg,
                                focal_alpha=alpha)[0]
        step = build(Bl, world, rank, dist.group.WORLD)
        step.image.copy_(image[rank * Bl:(rank + 1) * Bl])
        step.labels.copy_(label[rank * Bl:(rank + 1) * Bl].int())
        step.init_optimizer()
        if overlapped:
            step.train_step()
        else:
            step.forward_backward()
            step.sgd()
        res = {"grad": step.flat_grad.clone(), "losses": step.losses.clone(), "param": step.flat_param.clone(),
               "logits": step.logits.clone(), "classes": (step.c0, step.c1)}
        if rank == 0:
            ref = build(B_global, 1, 0, None)
            ref.image.copy_(image)
            ref.labels.copy_(label.int())
            ref.forward_backward()
            ref.init_optimizer()
            ref.sgd()
            res["ref"] = {"grad": ref.flat_grad.clone(), "losses": ref.losses.clone(), "param": ref.flat_param.clone(),
                          "logits": ref.logits.clone()}
        out[rank] = res
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kind,overlapped", [("maple", False), ("cocoop", False), ("maple", True), ("cocoop", True)])
def test_maple_cocoop_sharded_step_world2_gloo(kind, overlapped):
    """MaPLe: class-sharded text tower + data-parallel images, the coupling-Linear gradients receive contributions
    from both sides before the all-reduce.  CoCoOp: pure batch sharding (every rank runs all classes for its images),
    only the all-reduce of ctx + meta-net gradients.  Both must equal the single-process step."""
    from psrc_b200 import synthetic as S
    tokens = S.synthetic_tokens(10, name_lens=[1, 2, 3, 8, 1, 1, 2, 2, 1, 2])
    world, B_global = 2, 4
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_sharded_worker_other, args=(world, _free_port(), kind, tokens, B_global, out, overlapped), nprocs=world,
             join=True)
    ref = out[0]["ref"]
    want = [(0, 10), (0, 10)] if kind == "cocoop" else [(0, 5), (5, 10)]
    assert [out[r]["classes"] for r in range(world)] == want
    Bl = B_global // world
    for r in range(world):
        o = out[r]
        assert float((o["logits"] - ref["logits"][r * Bl:(r + 1) * Bl]).abs().max()) < 1e-5
        assert float((o["losses"] - ref["losses"]).abs().max()) < 1e-5, (o["losses"], ref["losses"])
        assert cos(o["grad"], ref["grad"]) > 0.999999
        # per-sequence prompt gradients pass through fp16 (autograd of `.half()`, model.py:238): a different fp32
        # summation order (2 vs 4 images per GEMM) may flip single fp16 roundings -- the error is quantised in fp16 ulps
        # of one sequence's gradient element (measured 8.0e-5 or 1.6e-4 of max|grad| = one or two flips; identical
        # with exact fp64 accumulation in the restated GEMMs, with or without the last-block pruning): bound = 2^-10
        assert float((o["grad"] - ref["grad"]).abs().max()) < 2.0 ** -10 * float(ref["grad"].abs().max())
    assert float((out[0]["param"] - out[1]["param"]).abs().max()) == 0.0


def _trainer_worker(rank, world, port, out):
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        TorchKernels().install(_Patch())
        import trainer_checks as TC
        from psrc_b200 import synthetic as S
        Bg = 4
        tr, cfg, tokens = TC.make_trainer("PromptSRC", "fp32", device="cpu", **{"DATALOADER.TRAIN_X.BATCH_SIZE": Bg})
        assert tr.model._step.cfg.world_size == world and tr.model._step.cfg.batch == Bg // world
        image, label = S.synthetic_batch(Bg, tokens.shape[0], 64, seed=1)
        losses = []
        for s_ in range(2):                       # Dassl hands EVERY rank the full batch (no DistributedSampler)
            tr.batch_idx = 0
            losses.append(tr.forward_backward({"img": image, "label": label})["loss"])
        assert tr.model._ranks_share_batch        # the first step compared the ranks' label fingerprints
        try:                                      # unseeded loaders: every rank holds another batch -> refused
            tr.model.check_ranks_share_the_batch(label + rank)
            refused = False
        except RuntimeError as ex:
            refused = "seed" in str(ex)
        out[rank] = {"losses": losses, "param": tr.model._step.flat_param.clone(), "refused": refused}
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_trainer_level_two_ranks_slice_the_full_batch_and_match_one_process(monkeypatch):
    """`forward_backward(batch)` of the drop-in trainer under torch.distributed -- every rank receives the
    full BATCH_SIZE from Dassl's loader, trains on its contiguous share through the sharded step, and the two ranks
    end with the parameters of the single-process trainer on the whole batch."""
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_trainer_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    TorchKernels().install(monkeypatch)
    import trainer_checks as TC
    from psrc_b200 import synthetic as S
    tr, cfg, tokens = TC.make_trainer("PromptSRC", "fp32", device="cpu", **{"DATALOADER.TRAIN_X.BATCH_SIZE": 4})
    image, label = S.synthetic_batch(4, tokens.shape[0], 64, seed=1)
    ref_losses = []
    for s_ in range(2):
        tr.batch_idx = 0
        ref_losses.append(tr.forward_backward({"img": image, "label": label})["loss"])
    ref = tr.model._step.flat_param
    assert float((out[0]["param"] - out[1]["param"]).abs().max()) == 0.0
    for r in range(world):
        assert out[r]["refused"], "ranks holding different batches must be refused (ADVICE r2)"
        assert max(abs(a - b) for a, b in zip(out[r]["losses"], ref_losses)) < 1e-5, (out[r]["losses"], ref_losses)
        assert float((out[r]["param"] - ref).abs().max()) < 1e-7, float((out[r]["param"] - ref).abs().max())
