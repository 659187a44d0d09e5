"""N-GPU parity of the class-sharded / data-parallel step against the single-GPU step on the global batch.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29541 \
      tests/multigpu_check.py [--classes 1000] [--batch 16] [--prec fp32]

Every rank runs its shard through NCCL (all-gather of t_hat, reduce-scatter of dL/dt_hat, all-reduce of the flat
prompt gradient); rank 0 additionally runs the whole batch / all classes alone and compares losses, gradients
and the parameters after one SGD step.  Prints one JSON line; exit code 1 on a mismatch.

--graphs: the path the trainers and bench.py actually time at world_size > 1 -- `capture()` (one CUDA graph per
segment between two collectives, one shared pool, NCCL outside, `step.py:_capture_segments`) and `run_host()` replays
with the batch arriving on the copy stream.  From one common state, three steps on three different batches are run
(a) eagerly, sharded, (b) as graph replays, sharded, (c) as graph replays on ONE GPU over the global batch.  (a) vs (b)
are the same launches on the same buffers, so they must agree to the last bit on every rank -- a segment graph that
read a buffer before NCCL had filled it, or a pool slot reused across segments, shows up here; (b) vs (c) is the
sharding parity over several steps (momentum included).
--trainer cocoop | maple: the same for the pure batch-sharded CoCoOp step and the coupled MaPLe step."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

from helpers import cocoop_setup, cos, maple_setup, promptsrc_setup  # noqa: E402


def graph_replay_check(args, step, ref, image0, label0, world, rank, dev):
    """Three steps from one state: sharded eager vs sharded captured segments (bit-identical, every rank) vs the
    captured single-GPU step on the global batch (rank 0).  Returns rank 0's verdict (other ranks: their own (a)=(b))."""
    from psrc_b200 import synthetic as S
    b, Bg, C = args.batch, args.batch * world, step.cfg.n_cls
    res = step.cfg.arch.image_resolution
    batches = []
    for s_ in (2, 3, 4):
        im, lb = S.synthetic_batch(Bg, C, res, seed=s_)
        batches.append((im.pin_memory(), lb.pin_memory()))
    lo = rank * b
    state = (step.flat_param.clone(), step.mom_buf.clone())      # identical on every rank (checked by the caller too)

    def restore(st):
        st.flat_param.copy_(state[0])
        st.mom_buf.copy_(state[1])
        st.invalidate_text()

    # (a) eager launches, the order of run_segments(None)
    eager_losses = []
    for im, lb in batches:
        step.image.copy_(im[lo:lo + b])
        step.labels.copy_(lb[lo:lo + b].int())
        step.train_step()
        eager_losses.append(step.losses.clone())
    p_eager = step.flat_param.clone()
    torch.cuda.synchronize()
    # (b) captured segments, batch arriving through the copy stream (run_host: what trainers / bench.py time)
    restore(step)
    step.capture(warmup=1)                                        # capture() puts parameters and momentum back itself
    assert step._seg_graphs is not None
    graph_losses = []
    for im, lb in batches:
        step.run_host(im[lo:lo + b].contiguous().pin_memory(), lb[lo:lo + b].contiguous().pin_memory())
        graph_losses.append(step.losses.clone())
    torch.cuda.synchronize()
    p_graph = step.flat_param.clone()
    m = {"graphs": len(step._seg_graphs),
         "graph_vs_eager_param_maxabs": float((p_graph - p_eager).abs().max()),
         "graph_vs_eager_loss_maxabs": max(float((a - g).abs().max()) for a, g in zip(eager_losses, graph_losses))}
    ok = m["graph_vs_eager_param_maxabs"] == 0.0 and m["graph_vs_eager_loss_maxabs"] == 0.0
    if rank == 0:
        # (c) one GPU, global batch, its own captured graphs, same three batches from the same state
        restore(ref)
        ref.capture(warmup=1)
        single_losses = []
        for im, lb in batches:
            ref.run_host(im, lb)
            single_losses.append(ref.losses.clone())
        torch.cuda.synchronize()
        d_g, d_s = p_graph - state[0], ref.flat_param - state[0]   # what three SGD steps did to the prompts
        m.update({"steps": len(batches),
                  "graph_vs_single_loss_maxabs": max(float((a - g).abs().max()) for a, g in zip(single_losses, graph_losses)),
                  "graph_vs_single_update_cos": cos(d_g, d_s),
                  "graph_vs_single_update_rel": float((d_g - d_s).norm() / d_s.norm()),
                  "losses_graph": [float(x.sum()) for x in graph_losses],
                  "losses_single": [float(x.sum()) for x in single_losses]})
        tol = dict(loss=1e-4, cos=0.99999) if args.prec == "fp32" else dict(loss=5e-3, cos=0.9995)
        ok = ok and m["graph_vs_single_loss_maxabs"] < tol["loss"] and m["graph_vs_single_update_cos"] > tol["cos"]
        m["graphs_ok"] = ok
        print(json.dumps(m), flush=True)
    elif not ok:
        print(json.dumps({"rank": rank, **m}), flush=True)
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--classes", type=int, default=1000)
    ap.add_argument("--batch", type=int, default=8, help="images per rank")
    ap.add_argument("--prec", default="fp32")
    ap.add_argument("--arch", default="ViT-B/16")
    ap.add_argument("--golden", default="", help="also compare the sharded step with this tests/golden fixture "
                                                  "(cfg3_promptsrc_ce: global batch 128 / world images per rank)")
    ap.add_argument("--graphs", action="store_true", help="also check captured-segment replays (see the docstring)")
    ap.add_argument("--trainer", default="promptsrc", choices=["promptsrc", "cocoop", "maple"])
    args = ap.parse_args()
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from psrc_b200 import synthetic as S
    arch = S.ARCHS[args.arch]
    fx = None
    if args.golden:
        from helpers import load_golden, ref_name
        fx = load_golden(args.golden)
        args.classes = fx["tokens"].shape[0]
        assert fx["label"].shape[0] % world == 0 and not int(fx["focal"])
        args.batch = fx["label"].shape[0] // world
    C = args.classes - args.classes % world
    tokens = S.load_token_fixture()[:C] if fx is None else fx["tokens"]
    Bg = args.batch * world
    image, label = S.synthetic_batch(Bg, C, arch.image_resolution, seed=1)
    depth = 9 if args.arch != "tiny" else 3

    def make(batch, w=1, r=0, pg=None, sd=None, fixed=None):
        """(step, sd, fixed) of the chosen trainer: sharded over `w` ranks, or the single-GPU step (w = 1)."""
        if args.trainer == "promptsrc":
            return promptsrc_setup(args.arch, depth, tokens, dev, args.prec, batch=batch, world=w, rank=r, pg=pg, sd=sd,
                                   fixed=fixed)
        if args.trainer == "cocoop":
            st, sd_, _ = cocoop_setup(args.arch, tokens, dev, args.prec, batch=batch, world=w, rank=r, pg=pg, sd=sd)
        else:
            st, sd_, _ = maple_setup(args.arch, depth, tokens, dev, args.prec, batch=batch, world=w, rank=r, pg=pg, sd=sd)
        return st, sd_, None

    step, sd, fixed = make(args.batch, world, rank, dist.group.WORLD)
    lo = rank * args.batch
    step.image.copy_(image[lo:lo + args.batch])
    step.labels.copy_(label[lo:lo + args.batch].int())
    step.init_optimizer()
    step.train_step()
    torch.cuda.synchronize()
    ok, line = True, None
    if rank == 0:
        ref, _, _ = make(Bg, sd=sd, fixed=fixed)
        ref.image.copy_(image)
        ref.labels.copy_(label.int())
        ref.init_optimizer()
        # gradients of the reference step BEFORE its SGD update
        ref.forward_backward()
        g_ref = ref.flat_grad.clone()
        ref.sgd()
        torch.cuda.synchronize()
        m = {"world": world, "prec": args.prec, "classes": C, "global_batch": Bg,
             "loss_sharded": float(step.losses.sum()), "loss_single": float(ref.losses.sum()),
             "losses_maxabs": float((step.losses - ref.losses).abs().max()),
             "grad_cos": cos(step.flat_grad, g_ref),
             "grad_rel": float((step.flat_grad - g_ref).norm() / g_ref.norm()),
             "param_maxabs": float((step.flat_param - ref.flat_param).abs().max()),
             "t_hat_maxabs": float((step.t_hat - ref.t_hat).abs().max()) if step.share_text else 0.0,
             "logits_maxabs": float((step.logits - ref.logits[:args.batch]).abs().max()), "trainer": args.trainer}
        tol = dict(loss=1e-4, cos=0.99999, t=1e-5) if args.prec == "fp32" else dict(loss=5e-3, cos=0.9995, t=5e-3)
        ok = (m["losses_maxabs"] < tol["loss"] and m["grad_cos"] > tol["cos"] and m["t_hat_maxabs"] < tol["t"])
        if fx is not None:   # the N-GPU step against the oracle / reference outputs of the whole batch
            m["golden"] = args.golden
            m["golden_loss_rel"] = abs(m["loss_sharded"] - float(fx["loss"])) / abs(float(fx["loss"]))
            m["golden_logits_maxabs"] = float((step.logits.cpu() - torch.from_numpy(fx["logits"][:args.batch])).abs().max())
            cs = {k: cos(step.grads[k], fx["grad/" + ref_name(k)]) for k in step.params}
            m["golden_gradcos_worst"] = min(cs.values())
            m["golden_gradcos_worst_key"] = min(cs, key=cs.get)
            gt = dict(loss=1e-5, logits=1e-4, cos=0.9999) if args.prec == "fp32" else dict(loss=1e-2, logits=2e-2, cos=0.999)
            ok = ok and m["golden_loss_rel"] < gt["loss"] and m["golden_logits_maxabs"] < gt["logits"] \
                and m["golden_gradcos_worst"] >= gt["cos"]
        m["ok"] = ok
        line = json.dumps(m)
        print(line, flush=True)
    if args.graphs:
        ok = graph_replay_check(args, step, ref if rank == 0 else None, image, label, world, rank, dev) and ok
    # every rank must hold identical parameters after the step
    p0 = step.flat_param.clone()
    dist.broadcast(p0, 0)
    same = float((p0 - step.flat_param).abs().max()) == 0.0
    flag = torch.tensor([1.0 if (ok and same) else 0.0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(json.dumps({"all_ranks_identical_params_and_ok": bool(flag.item())}), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() else 1)


if __name__ == "__main__":
    main()
