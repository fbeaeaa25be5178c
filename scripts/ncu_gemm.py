"""Tiny driver for ncu: a few launches of the tcgen05 GEMM on the ALBERT FFN-up shape."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from dedloc_b200.ops.gemm import gemm

M, N, K = 32768, 4096, 1024
pair = int(sys.argv[1]) if len(sys.argv) > 1 else 0
a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16) * 0.1
w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) * 0.1
bias = torch.randn(N, device="cuda", dtype=torch.bfloat16)
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(4):
    gemm(a, w, out, bias=bias, cta_pair=pair)
torch.cuda.synchronize()
print("ok", out.float().abs().mean().item())
