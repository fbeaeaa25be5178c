"""Device-timed GEMMs WITH the fused epilogues the ALBERT layer uses, burst (L2 flushed, few iterations) and sustained
(200 back-to-back launches, power-capped clocks) -- the in-situ rates that the plain-GEMM table does not show."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from dedloc_b200.ops.gemm import gemm, pick_split_k

M = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
dev = "cuda"


def sustained(fn, iters=200):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def bf(*shape):
    return torch.randn(*shape, device=dev, dtype=torch.bfloat16) * 0.1


x, w1, w2 = bf(M, 1024), bf(4096, 1024), bf(1024, 4096)
b1, b2 = bf(4096), bf(1024)
h, pre, dpre = torch.empty(M, 4096, device=dev, dtype=torch.bfloat16), torch.empty(M, 4096, device=dev, dtype=torch.bfloat16), torch.empty(M, 4096, device=dev, dtype=torch.bfloat16)
y, res, dy, dx = torch.empty(M, 1024, device=dev, dtype=torch.bfloat16), bf(M, 1024), bf(M, 1024), torch.empty(M, 1024, device=dev, dtype=torch.bfloat16)
g1, g2 = torch.zeros(4096, 1024, device=dev), torch.zeros(1024, 4096, device=dev)
pre.copy_(bf(M, 4096)); h.copy_(bf(M, 4096))
flops = 2.0 * M * 1024 * 4096
cases = {
    "ffn_up plain": lambda: gemm(x, w1, h),
    "ffn_up bias": lambda: gemm(x, w1, h, bias=b1),
    "ffn_up bias+gelu": lambda: gemm(x, w1, h, bias=b1, gelu=True),
    "ffn_up bias+save_pre": lambda: gemm(x, w1, h, bias=b1, pre_out=pre),
    "ffn_up bias+gelu+save_pre": lambda: gemm(x, w1, h, bias=b1, gelu=True, pre_out=pre),
    "ffn_down plain": lambda: gemm(h, w2, y),
    "ffn_down bias": lambda: gemm(h, w2, y, bias=b2),
    "ffn_down bias+residual": lambda: gemm(h, w2, y, bias=b2, residual=res),
    "ffn_down dgrad plain": lambda: gemm(dy, w2, dpre, b_mn_major=True),
    "ffn_up bias+gelu+save_dgelu": lambda: gemm(x, w1, h, bias=b1, gelu=True, pre_out=pre, save_dgelu=True),
    "ffn_down dgrad * gelu'(pre)": lambda: gemm(dy, w2, dpre, b_mn_major=True, dgelu_aux=pre),
    "ffn_down dgrad * saved": lambda: gemm(dy, w2, dpre, b_mn_major=True, mul_aux=pre),
    "ffn_down wgrad": lambda: gemm(dy, h, g2, a_mn_major=True, b_mn_major=True, accumulate=True, split_k=pick_split_k(1024, 4096, M, 148)),
    "ffn_up dgrad": lambda: gemm(dpre, w1, dx, b_mn_major=True),
    "ffn_up wgrad": lambda: gemm(dpre, x, g1, a_mn_major=True, b_mn_major=True, accumulate=True, split_k=pick_split_k(4096, 1024, M, 148)),
    "cublas ffn_up": lambda: torch.matmul(x, w1.t()),
    "cublas ffn_down wgrad": lambda: torch.matmul(dy.t(), h),
}
rows = {}
for name, fn in cases.items():
    t = sustained(fn)
    rows[name] = {"sustained_ms": round(t, 4), "sustained_tflops": round(flops / t / 1e9, 1)}
    print(name, rows[name], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump({"M": M, "rows": rows}, open("gpurun_out/gemm_epilogue_bench.json", "w"), indent=1)
