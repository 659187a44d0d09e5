"""ncu target: the one-query attention kernels of the last vision block at the cfg-3 shape, one warm + one measured launch.
  ncu --set full --clock-control none --import-source on -k regex:attn_row -c 4 -o gpurun_out/attn_row_full \
      python tests/ncu_attn_row_target.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import psrc_b200  # noqa: E402,F401
from psrc_b200 import ops  # noqa: E402

if __name__ == "__main__":
    n, L, H, dev = 128, 201, 12, "cuda"
    D = H * 64
    qkv = torch.randn(n * L, 3 * D, device=dev).to(torch.bfloat16)
    out = torch.empty(n * L, D, device=dev, dtype=torch.bfloat16)
    lse = torch.empty(n, H, L, device=dev)
    dout = torch.randn(n * L, D, device=dev).to(torch.bfloat16)
    dqkv = torch.empty(n * L, 3 * D, device=dev, dtype=torch.bfloat16)
    for _ in range(2):
        ops.attn_row_fwd(qkv, out, lse, n, L, H, False)
        ops.attn_row_bwd(qkv, dout, lse, dqkv, n, L, H, False)
    torch.cuda.synchronize()
