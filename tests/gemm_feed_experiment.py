import sys, os
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo")
from gemm_bench import run
from psrc_b200 import ops
for (M, N, K, mode) in ((25728, 2304, 768, "none"), (25728, 768, 3072, "none"), (25728, 3072, 768, "none")):
    us, err = run(M, N, K, mode, ops.GEMM_TCGEN05, check=False)
    print(os.environ.get("PSRC_PAIR_DEBUG_SKIP", "0"), M, N, K, mode, f"{us:.1f} us {2.0*M*N*K/us/1e6:.0f} TF")
