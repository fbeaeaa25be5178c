"""Device-timed comparison of the tcgen05 GEMM against torch.matmul (cuBLAS) on the ALBERT-large shapes."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from dedloc_b200.ops.gemm import gemm


def time_fn(fn, iters=20, warmup=5, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    tokens = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    shapes = [("qkv", tokens, 3072, 1024), ("attn_out", tokens, 1024, 1024), ("ffn_up", tokens, 4096, 1024),
              ("ffn_down", tokens, 1024, 4096), ("emb_proj", tokens, 1024, 128), ("mlm_dense", tokens, 128, 1024)]
    rows = []
    for name, M, N, K in shapes:
        a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16) * 0.1
        w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) * 0.1
        dy = torch.randn(M, N, device="cuda", dtype=torch.bfloat16) * 0.1
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        dx = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
        wg = torch.zeros(N, K, device="cuda", dtype=torch.float32)
        flops = 2.0 * M * N * K
        from dedloc_b200.ops.gemm import pick_split_k
        sk = pick_split_k(N, K, M, 148)
        res = {"name": name, "M": M, "N": N, "K": K}
        for tag, ours, ref in [
            ("fwd_1cta", lambda: gemm(a, w, out, cta_pair=1), lambda: torch.matmul(a, w.t())),
            ("fwd", lambda: gemm(a, w, out), lambda: torch.matmul(a, w.t())),
            ("dgrad", lambda: gemm(dy, w, dx, b_mn_major=True), lambda: torch.matmul(dy, w)),
            ("wgrad", lambda: gemm(dy, a, wg, a_mn_major=True, b_mn_major=True, accumulate=True, split_k=sk),
             lambda: torch.matmul(dy.t(), a)),
        ]:
            t_o, t_r = time_fn(ours, flush=flush), time_fn(ref, flush=flush)
            res[tag] = {"ours_ms": round(t_o, 4), "cublas_ms": round(t_r, 4), "ours_tflops": round(flops / t_o / 1e9, 1),
                        "cublas_tflops": round(flops / t_r / 1e9, 1)}
        rows.append(res)
        print(json.dumps(res), flush=True)
    with open("gpurun_out/gemm_bench.json", "w") as f:
        json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
