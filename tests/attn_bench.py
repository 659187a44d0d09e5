"""Dev tool (GPU box): time the attention kernels (tcgen05 vs mma.sync) on the cfg-3 shapes with CUDA events."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import psrc_b200  # noqa: E402,F401
from psrc_b200 import ops  # noqa: E402


def run(n_seq, L, H, causal, backend, reps=10):
    dev = "cuda"
    D = H * 64
    qkv = torch.randn(n_seq * L, 3 * D, device=dev).to(torch.bfloat16)
    out = torch.empty(n_seq * L, D, device=dev, dtype=torch.bfloat16)
    lse = torch.empty(n_seq, H, L, device=dev)
    dout = torch.randn(n_seq * L, D, device=dev).to(torch.bfloat16)
    dqkv = torch.empty_like(qkv)
    res = []
    for fn in (lambda: ops.attn_fwd(qkv, out, lse, n_seq, L, H, causal, backend),
               lambda: ops.attn_bwd(qkv, out, dout, lse, dqkv, n_seq, L, H, causal, backend)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        res.append(e0.elapsed_time(e1) / reps * 1e3)
    return res


if __name__ == "__main__":
    for (n, L, H, c) in ((128, 201, 12, False), (128, 197, 12, False), (1000, 77, 8, True), (1000, 11, 8, True), (1000, 16, 8, True), (1000, 24, 8, True)):
        row = f"n={n} L={L} H={H} causal={int(c)}"
        for name, b in (("tcgen05", 3), ("mma.sync", 2), ("one-tile mma", 5)):
            if (b == 3 and not (32 < L <= 256)) or (b == 5 and L > 32):
                continue
            f, bw = run(n, L, H, c, b)
            fl = 4.0 * n * H * L * L * 64 * (0.5 if c else 1.0)
            row += f" | {name}: fwd {f:7.1f} us ({fl / f / 1e6:6.1f} TF) bwd {bw:7.1f} us ({2.5 * fl / bw / 1e6:6.1f} TF)"
        print(row, flush=True)
