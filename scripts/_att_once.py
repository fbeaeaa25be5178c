import os, sys
sys.path.insert(0, "/root/repo")
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
