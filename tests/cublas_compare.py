import torch, sys
sys.path.insert(0, '/root/repo')
import psrc_b200
from psrc_b200 import ops
dev='cuda'
def t(fn, reps=30):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)/reps*1e3
for (M,N,K) in [(25728,2304,768),(25728,768,768),(25728,3072,768),(25728,768,3072),(25728,768,2304),(11000,1536,512),(11000,512,512),(11000,2048,512),(11000,512,2048),(8192,8192,8192)]:
    a=torch.randn(M,K,device=dev).to(torch.bfloat16); w=torch.randn(N,K,device=dev).to(torch.bfloat16)
    o=torch.empty(M,N,device=dev,dtype=torch.bfloat16)
    tc=t(lambda: torch.matmul(a,w.t(),out=o))
    to=t(lambda: ops.gemm(a,w,o))
    f=2.0*M*N*K
    print(f"{M:6d} {N:5d} {K:5d} | cuBLAS {tc:8.1f} us {f/tc/1e6:7.1f} TF/s | pair {to:8.1f} us {f/to/1e6:7.1f} TF/s | ratio {tc/to:5.2f}", flush=True)
