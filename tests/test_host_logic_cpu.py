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
    # 3 towers forward (text, prompted vision, frozen vision) x 3 layers; backward on two of them.  The LAST block of
    # the two vision towers runs the one-query attention (only the class token leaves the tower)
    assert calls.count("attn_fwd") == 7 and calls.count("attn_row_fwd") == 2
    assert calls.count("attn_bwd") == 5 and calls.count("attn_row_bwd") == 1
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


def test_trainer_dropin_builds_with_reference_names(monkeypatch):
    """Host logic of the Dassl-facing trainer with faked kernels: registry names, model registration name,
    reference state-dict keys, prompt parameters bound to the flat buffers, LR schedule."""
    fk = FakeKernels()
    fk.install(monkeypatch)
    import trainer_checks as TC
    from psrc_b200 import trainers as T
    assert {"PromptSRC", "CoOp"} <= set(T.TRAINER_REGISTRY.registered_names())
    tr, cfg, tokens = TC.make_trainer("PromptSRC", "fp32", device="cpu")
    assert tr.get_model_names() == ["VLPromptLearner"]
    st = tr.model._step
    sd = tr.model.state_dict()
    for k in ("prompt_learner.ctx", "prompt_learner.token_prefix", "prompt_learner.token_suffix", "image_encoder.VPT",
              "image_encoder.transformer.resblocks.2.VPT_shallow", "text_encoder.transformer.resblocks.1.VPT_shallow",
              "text_encoder.positional_embedding", "image_encoder.proj", "logit_scale"):
        assert k in sd, k
    trainable = {n for n, p in tr.model.named_parameters() if p.requires_grad}
    assert len(trainable) == 6 and all(("VPT" in n) or n.endswith("ctx") for n in trainable), trainable
    # parameters alias the flat buffer: a state-dict load lands in the buffer the kernels read
    tr.model.load_state_dict({"prompt_learner.ctx": torch.full_like(sd["prompt_learner.ctx"], 0.5)}, strict=False)
    assert float(st.params["ctx"].min()) == 0.5 == float(st.params["ctx"].max())
    assert tr.model.prompt_learner.ctx.data_ptr() == st.params["ctx"].data_ptr()
    # constant warm-up then cosine (WARMUP_RECOUNT): 1e-5, base, base * (1 + cos(pi / T)) / 2, ...
    import math
    lrs = []
    for _ in range(4):
        lrs.append(tr.optim.param_groups[0]["lr"])
        tr.update_lr()
    want = [1e-5, 0.0025, 0.00125 * (1 + math.cos(math.pi / 4)), 0.00125 * (1 + math.cos(2 * math.pi / 4))]
    assert all(abs(a - b) < 1e-12 for a, b in zip(lrs, want)), (lrs, want)
    # CoOp registers only the prompt learner (coop.py:431)
    tr2, _, _ = TC.make_trainer("CoOp", "fp32", device="cpu")
    assert tr2.get_model_names() == ["prompt_learner"] and list(tr2.model._step.params) == ["ctx"]


def test_maple_cocoop_dropins_build_with_reference_names(monkeypatch):
    """MaPLe / CoCoOp host logic with faked kernels: registration names, reference parameter names bound to the flat
    buffer in its order, launch sequence of one fused step."""
    fk = FakeKernels()
    fk.install(monkeypatch)
    import trainer_checks as TC
    from psrc_b200 import trainers as T
    assert {"PromptSRC", "CoOp", "MaPLe", "CoCoOp"} <= set(T.TRAINER_REGISTRY.registered_names())
    tr, cfg, tokens = TC.make_trainer("MaPLe", "fp32", device="cpu")
    assert tr.get_model_names() == ["MultiModalPromptLearner"]
    st = tr.model._step
    names = {n for n, p in tr.model.named_parameters() if p.requires_grad}
    want = {"prompt_learner.ctx", "prompt_learner.proj.weight", "prompt_learner.proj.bias"}
    for i in range(TC.DEPTH - 1):
        want |= {f"prompt_learner.compound_prompts_text.{i}", f"prompt_learner.compound_prompt_projections.{i}.weight",
                 f"prompt_learner.compound_prompt_projections.{i}.bias"}
    assert names == want, names ^ want
    pl = tr.model.prompt_learner
    assert pl.proj.weight.data_ptr() == st.params["couple.w.0"].data_ptr() == st.couple_w.data_ptr()
    assert pl.compound_prompts_text[0].data_ptr() == st.params["text.1"].data_ptr() == st.couple_x[1].data_ptr()
    assert pl.ctx.data_ptr() == st.couple_x.data_ptr()
    n0 = len(fk.calls)
    st.forward_backward()
    calls = fk.calls[n0:]
    assert calls.count("small_linear_fwd") == 1 and calls.count("small_linear_bwd") == 1   # all coupling Linears batched
    assert calls.count("attn_fwd") == 5 and calls.count("attn_bwd") == 5                   # two towers x 3 layers,
    assert calls.count("attn_row_fwd") == 1 and calls.count("attn_row_bwd") == 1           # one-query last vision block
    # CoCoOp: one text pass per image -> B x C sequences in ONE tower launch sequence
    tr2, _, tokens2 = TC.make_trainer("CoCoOp", "fp32", device="cpu")
    assert tr2.get_model_names() == ["prompt_learner"]
    st2 = tr2.model._step
    assert list(st2.params) == ["ctx", "meta.w1", "meta.b1", "meta.w2", "meta.b2"]
    assert st2.tt.n_seq == TC.B * tokens2.shape[0] and st2.eot.shape == (TC.B * tokens2.shape[0],)
    assert tr2.model.prompt_learner.meta_net.linear2.weight.data_ptr() == st2.params["meta.w2"].data_ptr()
    n0 = len(fk.calls)
    st2.forward_backward()
    calls = fk.calls[n0:]
    assert calls.count("small_linear_fwd") == 2 and calls.count("small_linear_bwd") == 2   # meta-net
    assert calls.count("attn_fwd") == 5 and calls.count("attn_row_fwd") == 1               # vision fwd only (last block
    assert calls.count("attn_bwd") == 3                                                    # one-query), text fwd+bwd
    assert calls.count("prompt_grad") == TC.B and calls.count("ctx_bias_grad") == 1
    assert calls.count("group_logits_fwd") == 1 and calls.count("group_logits_bwd") == 1


def test_lookahead_loader_yields_same_batches_with_the_following_one_attached():
    """trainers.LookaheadLoader (drop-in around Dassl's train_loader_x): same batches, same order, same length, each with
    the FOLLOWING batch under "_next" (None for the last), and attribute access falls through to the wrapped loader."""
    from psrc_b200.trainers import LookaheadLoader, _next_of

    class Loader(list):
        dataset = "the-dataset"

    batches = Loader({"img": torch.full((2, 3), float(i)), "label": torch.tensor([i, i])} for i in range(4))
    la = LookaheadLoader(batches)
    assert len(la) == 4 and la.dataset == "the-dataset"
    for epoch in range(2):                      # re-iterable, like a DataLoader
        seen = list(la)
        assert [int(b["label"][0]) for b in seen] == [0, 1, 2, 3]
        for i, b in enumerate(seen):
            nxt = _next_of(b)
            if i < 3:
                assert nxt[0] is batches[i + 1]["img"] and nxt[1] is batches[i + 1]["label"]
            else:
                assert nxt is None
    assert "_next" not in batches[0]            # the loader's own dicts are not modified
    assert _next_of({"img": 1, "label": 2}) is None


def test_train_set_smaller_than_batch_size_builds_the_step_for_the_short_batch(monkeypatch):
    """Dassl drops the last short batch only when the train set holds a full one (data_manager.py:37-48); a
    1-shot x 2-class set with BATCH_SIZE 3 arrives as ONE batch of 2 every epoch and the reference trains on it."""
    fk = FakeKernels()
    fk.install(monkeypatch)
    import trainer_checks as TC
    from psrc_b200 import trainers as T
    real_dm = TC.FakeDM

    class SmallDM(real_dm):
        def __init__(self, classnames):
            super().__init__(classnames)
            self.dataset.train_x = [object(), object()]          # 2 training items < BATCH_SIZE = 3

    monkeypatch.setattr(TC, "FakeDM", SmallDM)
    tr, cfg, tokens = TC.make_trainer("PromptSRC", "fp32", device="cpu")
    assert cfg.DATALOADER.TRAIN_X.BATCH_SIZE == 3 and tr.model._step.cfg.batch == 2
    assert not hasattr(T, "_BATCH_OVERRIDE")     # the size travels as an explicit constructor argument, no module state
    monkeypatch.setattr(TC, "FakeDM", real_dm)
    tr2, _, _ = TC.make_trainer("PromptSRC", "fp32", device="cpu")
    assert tr2.model._step.cfg.batch == 3


def test_init_weights_loads_prompts_and_refuses_changed_frozen_tensors(monkeypatch, tmp_path):
    """`cfg.MODEL.INIT_WEIGHTS` (promptsrc.py:253-254) is honoured for the prompt tensors; a file that would
    change a frozen CLIP tensor (already converted to the kernel layout) is refused, not half-applied."""
    fk = FakeKernels()
    fk.install(monkeypatch)
    import trainer_checks as TC
    tr, cfg, tokens = TC.make_trainer("PromptSRC", "fp32", device="cpu")
    sd = {k: v.clone() for k, v in tr.model.state_dict().items()}
    sd["prompt_learner.ctx"] = torch.full_like(sd["prompt_learner.ctx"], 0.25)
    f = tmp_path / "init.pth.tar"
    torch.save({"state_dict": sd, "epoch": 3}, f)
    tr2, _, _ = TC.make_trainer("PromptSRC", "fp32", device="cpu", **{"MODEL": {"INIT_WEIGHTS": str(f)}})
    assert float(tr2.model._step.params["ctx"].min()) == 0.25 == float(tr2.model._step.params["ctx"].max())
    sd["image_encoder.ln_pre.weight"] = sd["image_encoder.ln_pre.weight"] + 1.0
    torch.save({"state_dict": sd}, f)
    with pytest.raises(NotImplementedError):
        TC.make_trainer("PromptSRC", "fp32", device="cpu", **{"MODEL": {"INIT_WEIGHTS": str(f)}})
