"""Dev tool (GPU box): CUDA-event timings of single kernels at the cfg-3 shapes, cold L2 (a 512 MB buffer is rewritten
between calls, outside the timed region).  `python tests/kernel_bench.py [attn_row] [attn] [ln] ...`"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import psrc_b200  # noqa: E402,F401
from psrc_b200 import ops  # noqa: E402

_flush = None


def timed(fn, reps=8, flush=True):
    global _flush
    if _flush is None:
        _flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        fn()
    ts = []
    for _ in range(reps):
        if flush:
            _flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def bench_attn_row():
    dev = "cuda"
    for (n, L, H, causal, dtp) in ((128, 201, 12, False, torch.bfloat16), (128, 197, 12, False, torch.bfloat16),
                                   (1000, 11, 8, True, torch.float16), (16, 259, 16, False, torch.bfloat16)):
        D = H * 64
        qkv = torch.randn(n * L, 3 * D, device=dev).to(dtp)
        out = torch.empty(n * L, D, device=dev, dtype=dtp)
        lse = torch.empty(n, H, L, device=dev)
        dout = torch.randn(n * L, D, device=dev).to(torch.bfloat16)
        dqkv = torch.empty(n * L, 3 * D, device=dev, dtype=torch.bfloat16)
        idx = (torch.full((n,), L - 1, device=dev, dtype=torch.int32)) if causal else None
        f = timed(lambda: ops.attn_row_fwd(qkv, out, lse, n, L, H, causal, idx=idx))
        b = timed(lambda: ops.attn_row_bwd(qkv, dout, lse, dqkv, n, L, H, causal, idx=idx))
        fw = timed(lambda: ops.attn_row_fwd(qkv, out, lse, n, L, H, causal, idx=idx), flush=False)
        bw = timed(lambda: ops.attn_row_bwd(qkv, dout, lse, dqkv, n, L, H, causal, idx=idx), flush=False)
        small = torch.zeros(64, device=dev)
        ov = timed(lambda: ops.axpy(small, small, 1.0))
        print(f"   warm L2: fwd {fw:7.1f} us  bwd {bw:7.1f} us;  harness overhead (tiny kernel after flush) {ov:5.1f} us")
        rd = 2 * n * L * D * 2
        print(f"attn_row n={n} L={L} H={H}: fwd {f:7.1f} us ({rd / f / 1e3:6.0f} GB/s of K+V)  "
              f"bwd {b:7.1f} us ({(rd + 3 * n * L * D * 2) / b / 1e3:6.0f} GB/s of K+V+dqkv)", flush=True)


def bench_attn():
    dev = "cuda"
    for (n, L, H, causal) in ((128, 201, 12, False), (128, 197, 12, False)):
        D = H * 64
        qkv = torch.randn(n * L, 3 * D, device=dev).to(torch.bfloat16)
        out = torch.empty(n * L, D, device=dev, dtype=torch.bfloat16)
        lse = torch.empty(n, H, L, device=dev)
        dout = torch.randn(n * L, D, device=dev).to(torch.bfloat16)
        dqkv = torch.empty_like(qkv)
        f = timed(lambda: ops.attn_fwd(qkv, out, lse, n, L, H, causal))
        b = timed(lambda: ops.attn_bwd(qkv, out, dout, lse, dqkv, n, L, H, causal))
        fl = 4.0 * n * H * L * L * 64
        print(f"attn n={n} L={L} H={H}: fwd {f:7.1f} us ({fl / f / 1e6:6.1f} TF/s)  bwd {b:7.1f} us "
              f"({2.5 * fl / b / 1e6:6.1f} TF/s)", flush=True)


def bench_ln():
    dev = "cuda"
    for (rows, D) in ((25728, 768), (11000, 512)):
        x = torch.randn(rows, D, device=dev)
        g, bt = torch.randn(D, device=dev), torch.randn(D, device=dev)
        y = torch.empty(rows, D, device=dev, dtype=torch.float16)
        mean, rstd = torch.empty(rows, device=dev), torch.empty(rows, device=dev)
        f = timed(lambda: ops.layernorm_fwd(x, g, bt, y, mean, rstd))
        dy = torch.randn(rows, D, device=dev).to(torch.bfloat16)
        dx = torch.randn(rows, D, device=dev)
        dxa = torch.empty(rows, D, device=dev, dtype=torch.bfloat16)
        b = timed(lambda: ops.layernorm_bwd(dy, x, g, mean, rstd, dx, dx_in=dx, dx_act=dxa))
        print(f"layernorm rows={rows} D={D}: fwd {f:6.1f} us ({rows * D * 6 / f / 1e3:5.0f} GB/s)  bwd {b:6.1f} us "
              f"({rows * D * 16 / b / 1e3:5.0f} GB/s)", flush=True)


GROUPS = {"attn_row": bench_attn_row, "attn": bench_attn, "ln": bench_ln}

if __name__ == "__main__":
    for name in (sys.argv[1:] or list(GROUPS)):
        GROUPS[name]()
