"""Small-shape pass over every hand-written kernel, for `compute-sanitizer --tool memcheck|racecheck python …`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from dedloc_b200.ops.attention import packed_attention
from dedloc_b200.ops.cross_entropy import cross_entropy_rows
from dedloc_b200.ops.gemm import gemm
from dedloc_b200.ops.layernorm import layer_norm

torch.manual_seed(0)
bf = lambda *s: (torch.randn(*s, device="cuda") * 0.3).to(torch.bfloat16)
M, N, K = 512, 512, 256
a, w, bias, res, aux = bf(M, K), bf(N, K), bf(N), bf(M, N), bf(M, N)
out, pre = torch.empty(M, N, device="cuda", dtype=torch.bfloat16), torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for pair in (1, 2):
    gemm(a, w, out, bias=bias, gelu=True, pre_out=pre, save_dgelu=True, cta_pair=pair)
    gemm(a, w, out, bias=bias, residual=res, cta_pair=pair)
    gemm(a, w, out, mul_aux=aux, cta_pair=pair)
    gemm(a, w.t().contiguous(), out, b_mn_major=True, dgelu_aux=aux, cta_pair=pair)
    wg = torch.zeros(N, K, device="cuda")
    gemm(res, a, wg, a_mn_major=True, b_mn_major=True, accumulate=True, split_k=2, cta_pair=pair)
qkv = bf(2, 256, 3 * 2 * 64).requires_grad_(True)
mask = torch.ones(2, 256, device="cuda")
mask[1, 200:] = 0
packed_attention(qkv, 2, mask).float().square().sum().backward()
packed_attention(qkv, 2).float().sum().backward()
logits = bf(8, 1000).requires_grad_(True)
cross_entropy_rows(logits, torch.randint(0, 1000, (8,), device="cuda")).sum().backward()
x = bf(64, 1024).requires_grad_(True)
ln = torch.nn.LayerNorm(1024, eps=1e-12).to("cuda", torch.bfloat16)
for prm in ln.parameters():
    prm.main_grad = torch.zeros(prm.shape, device="cuda")
layer_norm(x, ln).float().sum().backward()
torch.cuda.synchronize()
print("sanitize smoke ok", float(out.float().abs().mean()))
