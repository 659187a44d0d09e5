"""N-GPU parity of the class-sharded / data-parallel step against the single-GPU step on the global batch.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29541 \
      tests/multigpu_check.py [--classes 1000] [--batch 16] [--prec fp32]

Every rank runs its shard through NCCL (all-gather of t_hat, reduce-scatter of dL/dt_hat, all-reduce of the flat
prompt gradient); rank 0 additionally runs the whole batch / all classes alone and compares losses, gradients
and the parameters after one SGD step.  Prints one JSON line; exit code 1 on a mismatch."""
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

from helpers import cos, promptsrc_setup  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--classes", type=int, default=1000)
    ap.add_argument("--batch", type=int, default=8, help="images per rank")
    ap.add_argument("--prec", default="fp32")
    ap.add_argument("--arch", default="ViT-B/16")
    ap.add_argument("--golden", default="", help="also compare the sharded step with this tests/golden fixture "
                                                  "(cfg3_promptsrc_ce: global batch 128 / world images per rank)")
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
    step, sd, fixed = promptsrc_setup(args.arch, depth, tokens, dev, args.prec, batch=args.batch, world=world, rank=rank,
                                      pg=dist.group.WORLD)
    lo = rank * args.batch
    step.image.copy_(image[lo:lo + args.batch])
    step.labels.copy_(label[lo:lo + args.batch].int())
    step.init_optimizer()
    step.train_step()
    torch.cuda.synchronize()
    ok, line = True, None
    if rank == 0:
        ref, _, _ = promptsrc_setup(args.arch, depth, tokens, dev, args.prec, batch=Bg, sd=sd, fixed=fixed)
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
             "t_hat_maxabs": float((step.t_hat - ref.t_hat).abs().max()),
             "logits_maxabs": float((step.logits - ref.logits[:args.batch]).abs().max())}
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
