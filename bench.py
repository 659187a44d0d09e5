#!/usr/bin/env python
"""bench.py -- train images/sec of the prompted-CLIP (PromptSRC) step, BASELINE.json cfg #3.

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
  python bench.py --impl reference ...                      CPU arm: the oracle port on host cores

A "step" = one PromptSRC train step over one synthetic batch: prompted ViT-B/16 (4 deep visual prompts,
depth 9) over B images, text transformer over the rank's class shard of C = 1000 prompts, cosine logits,
CE + 25*L1(text) + 10*L1(image) + KL(logits) against the frozen CLIP branch, prompt-only backward, SGD.
One JSON line on stdout (rank 0).  Keys: see DESIGN.md "Measurement".
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "train_images_per_sec"
UNIT = "img/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=128, help="images per GPU (weak) / global (strong)")
    ap.add_argument("--classes", type=int, default=1000)
    ap.add_argument("--prec", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--e2e-breakdown", action="store_true", help="diagnostic timings of the host-facing path (stderr)")
    ap.add_argument("--overlap-zs", type=int, default=1,
                    help="bit 0: frozen tower forward on a second stream (default on); bit 1: text backward on a second stream")
    ap.add_argument("--graph-multi", action="store_true", help="multi-GPU: capture the step incl. NCCL (experimental)")
    ap.add_argument("--no-truncate", action="store_true", help="run all 77 text positions (dense-equivalent)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-kernel-roofline", action="store_true")
    ap.add_argument("--cpu-sample-batch", type=int, default=8)
    ap.add_argument("--cpu-sample-classes", type=int, default=64)
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tf_burst": d["bf16_tflops"], "tf_sustained": d["bf16_tflops_sustained"],
                "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


def gemm_traffic():
    """DRAM read+write bytes per launch of the dominant GEMM flavours from the committed `ncu --set full` capture
    (profiles/gemm_traffic.json); the mean over the captured launches, next to their algorithmic bytes."""
    p = os.path.join(ROOT, "profiles", "gemm_traffic.json")
    if not os.path.exists(p):
        return None
    rows = json.load(open(p))["per_launch"]
    return {"dram_bytes_per_launch": sum(r["dram_read_bytes"] + r["dram_write_bytes"] for r in rows) / len(rows),
            "algorithmic_bytes_per_launch": sum(r["algorithmic_bytes"] for r in rows) / len(rows),
            "launches": len(rows), "source": "profiles/gemm_traffic.json (ncu --set full)"}


# ------------------------------------------------------------------ clocks sampler
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # "under load": samples within the top half of observed power draw
        thr = 0.5 * max(pw)
        load = [s for s, p in zip(sm, pw) if p >= thr] or sm
        return {"sm_mhz": statistics.median(load), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------ CPU arm (oracle port on host cores)
def cpu_oracle_leg(args, steps, warmup):
    """The oracle restatement of the reference PromptSRC step (oracle/clip_oracle.py: fp32 PyTorch + autograd,
    dense 77 text positions exactly like the reference) timed on the host cores on a bounded, proportional
    sample of cfg #3: `cpu_sample_batch` images with `cpu_sample_classes` classes (same classes-per-image
    ratio as 128 : 1000), forward + backward + SGD."""
    import torch
    import clip_oracle as O
    import psrc_b200  # noqa: F401
    from psrc_b200 import synthetic as S
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    arch = S.ARCHS["ViT-B/16"]
    B, C, depth, n_ctx = args.cpu_sample_batch, args.cpu_sample_classes, 9, 4
    sd = S.make_clip_state_dict(arch, seed=0)
    tokens = torch.as_tensor(S.load_token_fixture()[:C])
    image, label = S.synthetic_batch(B, C, arch.image_resolution, seed=1)
    fixed = O.teacher_text_features(sd, tokens, arch.transformer_heads, truncate=int(tokens.argmax(-1).max()) + 1)
    deep = S.make_deep_prompts(arch, n_ctx, n_ctx, depth, depth, seed=100)
    params = [sd["token_embedding.weight"][tokens[0, 1:1 + n_ctx]].clone(), deep["visual.VPT"].clone()]
    params += [deep[f"visual.transformer.resblocks.{i}.VPT_shallow"].clone() for i in range(1, depth)]
    params += [deep[f"transformer.resblocks.{i}.VPT_shallow"].clone() for i in range(1, depth)]
    params = [p.requires_grad_(True) for p in params]
    opt = torch.optim.SGD(params, lr=0.0025, momentum=0.9, weight_decay=5e-4)

    def one():
        ctx, vpt = params[0], params[1]
        dv = {i: params[1 + i] for i in range(1, depth)}
        dtx = {i: params[depth + i] for i in range(1, depth)}
        o = O.promptsrc_step(sd, tokens, image, label, ctx, vpt, dv, dtx, fixed, n_ctx, arch.vision_heads,
                             arch.transformer_heads)
        opt.zero_grad()
        o["loss"].backward()
        opt.step()
        return float(o["loss"].detach())

    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    dt = (time.perf_counter() - t0) / steps
    return {"value": B / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"PromptSRC ViT-B/16 fp32 step (fwd+bwd+SGD), {B} images x {C} classes, 77 text positions, "
                      f"{steps} timed steps after {warmup} warm-up, {dt:.2f} s/step, torch CPU {cores} threads"}, dt


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 3)), max(1, min(args.warmup, 1))
    cb, dt = cpu_oracle_leg(args, steps, warmup)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": workload_config(args, args.gpus, None, cpu=True),
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(args, world, Lt, cpu=False):
    gb = args.batch * world if args.scaling == "weak" else args.batch
    return {"workload": "PromptSRC ViT-B/16 train step, ImageNet-LT-shaped 1000 classes, 4+4 deep V/L prompts depth 9 "
                        "(BASELINE.json configs[2])",
            "global_batch": gb, "per_gpu_batch": gb // world, "classes": args.classes,
            "parallelism": f"dp{world} images x class-sharded text tower (all-gather t_hat / reduce-scatter dt_hat)",
            "text_positions": (77 if cpu else Lt),
            "l2": "per-step working set (saved activations, several GB) exceeds the 126 MB L2; no explicit flush"}


# ------------------------------------------------------------------ B200 arm
def b200_arm(args):
    import torch
    import torch.distributed as dist
    import psrc_b200  # noqa: F401
    from psrc_b200 import cabi, flops, ops, synthetic as S
    from psrc_b200.step import PromptedClipStep, StepConfig, encode_text_frozen

    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl b200 needs a CUDA device: there is no CPU fallback")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise RuntimeError(f"--gpus {args.gpus} must be launched under torchrun with {args.gpus} ranks")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    pg = None
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep stdout to the ONE JSON line
        dist.init_process_group("nccl", device_id=dev)
        pg = dist.group.WORLD
    cabi.load()
    peaks = load_peaks()

    arch = S.ARCHS["ViT-B/16"]
    C = args.classes
    B = args.batch if args.scaling == "weak" else args.batch // world
    tokens = torch.as_tensor(S.load_token_fixture()[:C])
    sd = S.make_clip_state_dict(arch, seed=0)
    shots = S.lt_counts_pareto(C)
    cfg = StepConfig(arch=arch, n_cls=C, batch=B, prec=args.prec, n_ctx_text=4, n_ctx_vision=4, depth_text=9,
                     depth_vision=9, zero_shot_branch=True, world_size=world, rank=rank, process_group=pg,
                     truncate_text=not args.no_truncate)
    fixed = encode_text_frozen(sd, tokens, dev, prec="fp32")               # frozen teacher, init-time
    step = PromptedClipStep(cfg, sd, tokens, dev, fixed_embeddings=fixed)
    deep = S.make_deep_prompts(arch, 4, 4, 9, 9, seed=100)
    with torch.no_grad():
        step.params["ctx"].copy_(sd["token_embedding.weight"][tokens[0, 1:5]])
        step.params["vpt"].copy_(deep["visual.VPT"])
        for i in range(1, 9):
            step.params[f"vision.{i}"].copy_(deep[f"visual.transformer.resblocks.{i}.VPT_shallow"])
            step.params[f"text.{i}"].copy_(deep[f"transformer.resblocks.{i}.VPT_shallow"])
    step.init_optimizer(lr=0.0025, momentum=0.9, weight_decay=5e-4)
    step.overlap_zero_shot = bool(args.overlap_zs & 1)
    step.overlap_backward = bool(args.overlap_zs & 2)
    del sd
    g = torch.Generator().manual_seed(1 + rank)
    image_host = torch.randn(B, 3, arch.image_resolution, arch.image_resolution, generator=g).pin_memory()
    label_host = torch.randint(0, C, (B,), generator=g).pin_memory()       # int64, as Dassl hands it over
    step.image.copy_(image_host)
    step.labels.copy_(label_host.to(torch.int32))
    torch.cuda.synchronize()

    # launches per step (eager pass, counted by the library)
    n0 = cabi.launch_count()
    step.train_step()
    torch.cuda.synchronize()
    launches = cabi.launch_count() - n0
    # multi-GPU: eager by default -- capturing the NCCL collectives was measured at N=2 (24.47 vs 24.57 ms/step: the
    # step is not launch-bound) and the captured communicator did not shut down cleanly, so it stays opt-in
    use_graph = not args.no_graph and (world == 1 or args.graph_multi)
    if use_graph:
        step.capture(warmup=1)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms) / steps

    sampler = ClockSampler(local)
    for _ in range(max(3, args.warmup)):
        step.run()
    if rank == 0:
        sampler.start()
    ms_dev = timed(step.run, args.steps)                                    # inputs resident in HBM
    # end to end through the host-facing call: pinned host batch -> H2D -> step -> SGD -> loss D2H
    # (the next batch's H2D copy is issued behind each step's launch, as trainers.LookaheadLoader does for Dassl's
    # loop: every timed step still copies one full batch from pinned host memory and reads its loss back)
    nxt = (image_host, label_host)
    for _ in range(2):
        step.train_step_host(image_host, label_host, next_host=nxt)
    ms_e2e = timed(lambda: step.train_step_host(image_host, label_host, next_host=nxt), args.steps)
    if args.e2e_breakdown and rank == 0 and world == 1:    # diagnostic: where does e2e - value go on this box?
        again = timed(step.run, args.steps)
        step.stage_host(*nxt)
        nosync = timed(lambda: (step.run_staged(), step.stage_host(*nxt)), args.steps)
        plain = timed(lambda: step.train_step_host(image_host, label_host), args.steps)
        t0 = time.perf_counter()
        for _ in range(20):
            step.stage_host(*nxt)
        torch.cuda.synchronize()
        h2d = (time.perf_counter() - t0) / 20 * 1e3
        print(f"[e2e breakdown] device-resident {ms_dev:.3f} ms, e2e {ms_e2e:.3f}, device-resident again {again:.3f}, "
              f"staged without the per-step loss read {nosync:.3f}, no look-ahead {plain:.3f}, H2D alone {h2d:.3f} ms",
              file=sys.stderr)
    clocks = sampler.stop() if rank == 0 else None
    loss_value = float(step.losses.sum())

    gb = B * world
    fl = flops.step_flops(arch, B, step.Cl, step.Lt, 4, True, True)
    dense = flops.step_flops(arch, B, step.Cl, 77, 4, True, True)
    tf_step = fl["total"] / (ms_dev * 1e-3) / 1e12                          # per rank

    # ---- roofline of the dominant kernel (tcgen05 GEMM): every GEMM shape of the step timed alone with CUDA
    #      events on the launching stream; operands sized like the step's (A + out >> L2 for the big shapes)
    roof = None
    if not args.no_kernel_roofline and args.prec == "bf16":
        shapes = flops.gemm_launches(arch, B, step.Cl, step.Lt, 4, True, True)
        tot_f, tot_ms, n_launch, per_shape = 0.0, 0.0, 0, []
        for (M, N, K, cnt, mode) in shapes:
            if K % 8:
                continue
            a = torch.randn(M, K, device=dev).to(torch.bfloat16)
            w = torch.randn(N, K, device=dev).to(torch.bfloat16)
            bias = torch.zeros(N, device=dev)
            kw = {}
            if mode in ("res", "res_inplace"):        # fp32 residual stream in / out
                o = torch.empty(M, N, device=dev)
                kw["residual"] = o if mode == "res_inplace" else torch.randn(M, N, device=dev)
                kw["alpha"] = 1e-3                    # (in place: keeps the repeated accumulation finite)
            else:
                o = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
                if mode in ("gelu", "gelu_save"):
                    kw["act"] = ops.ACT_QUICKGELU
                    if mode == "gelu_save":
                        kw["out2"] = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
                elif mode == "gelugrad":
                    kw["act"] = ops.ACT_QUICKGELU_GRAD
                    kw["aux"] = torch.randn(M, N, device=dev).to(torch.bfloat16)
            if mode != "gelugrad":
                kw["bias"] = bias
            for _ in range(3):
                ops.gemm(a, w, o, **kw)
            reps = 20
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(reps):
                ops.gemm(a, w, o, **kw)
            e1.record()
            torch.cuda.synchronize()
            t = e0.elapsed_time(e1) / reps
            f = 2.0 * M * N * K
            per_shape.append({"M": M, "N": N, "K": K, "count": cnt, "epilogue": mode, "us": round(t * 1e3, 2),
                              "tflops": round(f / (t * 1e-3) / 1e12, 1)})
            tot_f += f * cnt; tot_ms += t * cnt; n_launch += cnt
            del a, w, o, kw
        ach = tot_f / (tot_ms * 1e-3) / 1e12
        roof = {"bound": "tensor", "kernel": "pair::gemm_pair_kernel (tcgen05.mma.cta_group::2 + TMA + TMEM)", "achieved": round(ach, 1),
                "peak": peaks["tf_burst"], "unit": "TFLOP/s", "frac": round(ach / peaks["tf_burst"], 4),
                "peak_source": peaks["source"] + ", burst bf16 (kernel timed alone)", "traffic": gemm_traffic(),
                "launches_per_step": n_launch, "avg_launch_us": round(tot_ms / n_launch * 1e3, 2),
                "gemm_share_of_step": round(tot_ms / ms_dev, 3), "per_shape": per_shape}

    cpu_b = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_b, _ = cpu_oracle_leg(args, steps=2, warmup=1)

    if rank == 0:
        line = {"metric": METRIC, "value": gb / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(3, args.warmup), "ms_per_step": ms_dev, "higher_is_better": True,
                "scaling": args.scaling, "vs_baseline": None, "dtype": args.prec, "data": "synthetic",
                "config": workload_config(args, world, step.Lt),
                "e2e": {"value": gb / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                        "h2d_bytes_per_step": image_host.numel() * 4 + label_host.numel() * 8,
                        "d2h_bytes_per_step": 16, "api": "PromptedClipStep.train_step_host (pinned host batch; the following batch's H2D is issued under the step)"},
                "gpu_launches": launches * args.steps, "gpu_launches_per_step": launches,
                "cuda_graph": use_graph, "clocks": clocks, "loss": loss_value,
                "step_tflops_per_gpu": round(tf_step, 1),
                "step_mfu_vs_sustained": round(tf_step / peaks["tf_sustained"], 4),
                "step_mfu_vs_nominal_2250": round(tf_step / 2250.0, 4),
                "executed_tflop_per_step_per_gpu": round(fl["total"] / 1e12, 3),
                "dense_equiv_tflop_per_step_per_gpu": round(dense["total"] / 1e12, 3),
                "roofline": roof, "cpu_baseline": cpu_b}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        b200_arm(args)


if __name__ == "__main__":
    main()
