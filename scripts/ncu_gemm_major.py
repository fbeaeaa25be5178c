"""ncu driver: the same GEMM (M=32768, N=1024, K=4096) with K-major and with MN-major operands."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from dedloc_b200.ops.gemm import gemm

M, N, K = 32768, 1024, 4096
a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16) * 0.1
w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) * 0.1
at, wt = a.t().contiguous(), w.t().contiguous()
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(2):
    gemm(a, w, out)
    gemm(at, wt, out, a_mn_major=True, b_mn_major=True)
torch.cuda.synchronize()
print("ok")
