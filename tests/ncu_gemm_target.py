"""ncu target: the four vision-tower GEMM flavours of the cfg-3 step, two launches each (warm + measured).
  ncu --set full --clock-control none --import-source on -k regex:gemm_pair -c 8 -o gpurun_out/gemm_full \
      python tests/ncu_gemm_target.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
from gemm_bench import run  # noqa: E402
from psrc_b200 import ops  # noqa: E402

if __name__ == "__main__":
    for (M, N, K, mode) in ((25728, 2304, 768, "none"), (25728, 768, 3072, "res"), (25728, 3072, 768, "gelu"),
                            (25728, 3072, 768, "gelugrad")):
        us, err = run(M, N, K, mode, ops.GEMM_TCGEN05, reps=1, check=False)
        print(M, N, K, mode, f"{us:.1f} us")
    torch.cuda.synchronize()
