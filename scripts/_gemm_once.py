import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dedloc_b200.ops.gemm import gemm
M = 32768
bf = lambda *s: torch.randn(*s, device="cuda", dtype=torch.bfloat16) * 0.1
x, w1, w2, dy = bf(M, 1024), bf(4096, 1024), bf(1024, 4096), bf(M, 1024)
h, pre, dpre = bf(M, 4096), bf(M, 4096), torch.empty(M, 4096, device="cuda", dtype=torch.bfloat16)
b1 = bf(4096)
for name, fn in [("ffn_up plain", lambda: gemm(x, w1, h)), ("ffn_up bias+gelu", lambda: gemm(x, w1, h, bias=b1, gelu=True)), ("ffn_up bias+gelu+pre", lambda: gemm(x, w1, h, bias=b1, gelu=True, pre_out=pre)),
                 ("dgrad plain", lambda: gemm(dy, w2, dpre, b_mn_major=True)), ("dgrad dgelu", lambda: gemm(dy, w2, dpre, b_mn_major=True, dgelu_aux=pre))]:
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    print("==", name, flush=True)
    fn()
    torch.cuda.synchronize()
