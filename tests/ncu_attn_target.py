"""ncu target: the vision-tower attention of the cfg-3 step (tcgen05 kernels), warm + measured launch each.
  ncu --set full --clock-control none --import-source on -k regex:attn_ -c 4 -o gpurun_out/attn_full \
      python tests/ncu_attn_target.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
from attn_bench import run  # noqa: E402

if __name__ == "__main__":
    print(run(128, 201, 12, False, 3, reps=1))
    torch.cuda.synchronize()
