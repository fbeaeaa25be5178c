"""Sustained rate of the same GEMM with K-major and MN-major operands (the MN-major penalty of the pair kernel)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from dedloc_b200.ops.gemm import gemm

M = 32768
bf = lambda *s: torch.randn(*s, device="cuda", dtype=torch.bfloat16) * 0.1


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


for (N, K) in [(1024, 4096), (4096, 1024), (1024, 1024)]:
    a, w = bf(M, K), bf(N, K)
    wt, at = w.t().contiguous(), a.t().contiguous()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    fl = 2.0 * M * N * K
    for name, fn in [("A K, B K", lambda: gemm(a, w, out)), ("A K, B MN", lambda: gemm(a, wt, out, b_mn_major=True)),
                     ("A MN, B K", lambda: gemm(at, w, out, a_mn_major=True)),
                     ("A MN, B MN", lambda: gemm(at, wt, out, a_mn_major=True, b_mn_major=True))]:
        t = sustained(fn)
        print(f"N={N} K={K} {name}: {t:.4f} ms {fl / t / 1e9:.0f} TFLOP/s", flush=True)
