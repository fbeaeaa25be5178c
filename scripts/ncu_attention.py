"""Tiny driver for ncu: two forward+backward launches of the packed-QKV tcgen05 attention at the ALBERT shape."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from dedloc_b200.ops.attention import packed_attention

B, S, H = 64, 512, 16
qkv = (torch.randn(B, S, 3 * H * 64, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_(True)
dout = (torch.randn(B, S, H * 64, device="cuda") * 0.5).to(torch.bfloat16)
for _ in range(2):
    out = packed_attention(qkv, H)
    out.backward(dout)
    qkv.grad = None
torch.cuda.synchronize()
print("ok", out.float().abs().mean().item())
