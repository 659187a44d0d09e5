"""Dev tool (GPU box): time every GEMM shape/epilogue of the cfg-3 step with CUDA events, CTA-pair kernel vs
the single-CTA kernel, and check both against torch.  `python tests/gemm_bench.py [quick]`"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import psrc_b200  # noqa: E402,F401
from psrc_b200 import ops  # noqa: E402


def qg(x):
    return x * torch.sigmoid(1.702 * x)


def qgg(x):
    s = torch.sigmoid(1.702 * x)
    return s * (1 + 1.702 * x * (1 - s))


def run(M, N, K, mode, backend, reps=20, check=True):
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(M + N + K)
    a = (torch.randn(M, K, device=dev, generator=g) * 0.5).to(torch.bfloat16)
    w = (torch.randn(N, K, device=dev, generator=g) * (K ** -0.5)).to(torch.bfloat16)
    bias = torch.randn(N, device=dev, generator=g)
    kw, out, ref_fn = {}, None, None
    if mode == "none":
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        kw = dict(bias=bias)
        ref_fn = lambda r: r + bias  # noqa: E731
    elif mode == "gelu":
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        pre = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        kw = dict(bias=bias, act=ops.ACT_QUICKGELU, out2=pre)
        ref_fn = lambda r: qg(r + bias)  # noqa: E731
    elif mode == "gelugrad":
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        u = torch.randn(M, N, device=dev, generator=g).to(torch.bfloat16)
        kw = dict(act=ops.ACT_QUICKGELU_GRAD, aux=u)
        ref_fn = lambda r: r * qgg(u.float())  # noqa: E731
    elif mode == "res":
        res = torch.randn(M, N, device=dev, generator=g)
        out = torch.empty(M, N, device=dev)
        kw = dict(bias=bias, residual=res)
        ref_fn = lambda r: r + bias + res  # noqa: E731
    for _ in range(3 if reps > 1 else 1):
        ops.gemm(a, w, out, backend=backend, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        ops.gemm(a, w, out, backend=backend, **kw)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    err = float("nan")
    if check:
        ref = ref_fn(a.float() @ w.float().t())
        err = float((out.float() - ref).norm() / ref.norm())
        if mode == "gelu":
            err = max(err, float((pre.float() - (a.float() @ w.float().t() + bias)).norm() / ref.norm()))
    return us, err


if __name__ == "__main__":
    quick = "quick" in sys.argv
    Mv, Mt = 25728, 11000
    shapes = [(Mv, 2304, 768, "none"), (Mv, 768, 768, "res"), (Mv, 3072, 768, "gelu"), (Mv, 768, 3072, "res"),
              (Mv, 3072, 768, "gelugrad"), (Mv, 768, 3072, "none"), (Mv, 768, 768, "none"), (Mv, 768, 2304, "none"),
              (Mt, 1536, 512, "none"), (Mt, 512, 512, "res"), (Mt, 2048, 512, "gelu"), (Mt, 512, 2048, "res"),
              (Mt, 2048, 512, "gelugrad"), (Mt, 512, 2048, "none"), (Mt, 512, 512, "none"), (Mt, 512, 1536, "none"),
              (300, 384, 768, "res"), (1000, 200, 72, "gelu"), (4, 512, 768, "none"), (128, 512, 768, "none")]
    if quick:
        shapes = shapes[:4] + shapes[-4:]
    if "tails" in sys.argv:     # pair kernel only, the shapes whose last round is partially filled (PSRC_PAIR_SPLIT A/B)
        Mz, Ms, M8 = 25216, 3216, 1375
        shapes = shapes[:16] + [(Mz, 768, 768, "res"), (Mz, 768, 3072, "res"), (Mz, 2304, 768, "none"),
                                (Mz, 3072, 768, "gelu"), (Ms, 2304, 768, "none"), (Ms, 768, 768, "res"),
                                (Ms, 3072, 768, "gelu"), (Ms, 768, 3072, "res"), (M8, 1536, 512, "none"),
                                (M8, 512, 512, "res"), (M8, 2048, 512, "gelu"), (M8, 512, 2048, "res"),
                                (128, 3072, 768, "gelu"), (128, 768, 3072, "res"), (128, 768, 768, "res"),
                                (128, 512, 768, "none")]
        print(f"PSRC_PAIR_SPLIT={os.environ.get('PSRC_PAIR_SPLIT', 'auto')}")
        for (M, N, K, mode) in shapes:
            u2, e2 = run(M, N, K, mode, ops.GEMM_TCGEN05)
            print(f"{M:6d} {N:5d} {K:5d} {mode:>9} | {u2:9.1f} us {2.0 * M * N * K / u2 / 1e6:7.1f} TF/s err {e2:9.2e}", flush=True)
        sys.exit(0)
    print(f"{'M':>6} {'N':>5} {'K':>5} {'mode':>9} | {'pair us':>9} {'TF':>7} {'err':>9} | {'1cta us':>9} {'TF':>7} {'err':>9}")
    for (M, N, K, mode) in shapes:
        f = 2.0 * M * N * K
        u2, e2 = run(M, N, K, mode, ops.GEMM_TCGEN05)
        u1, e1 = run(M, N, K, mode, ops.GEMM_TCGEN05_1CTA)
        print(f"{M:6d} {N:5d} {K:5d} {mode:>9} | {u2:9.1f} {f / u2 / 1e6:7.1f} {e2:9.2e} | {u1:9.1f} {f / u1 / 1e6:7.1f} {e1:9.2e}",
              flush=True)
