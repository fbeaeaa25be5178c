"""Device-timed attention forward/backward: tcgen05 packed-QKV kernels vs PyTorch SDPA (cuDNN) incl. its layout copies."""
import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from dedloc_b200.ops.attention import packed_attention


def time_fn(fn, iters=10, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


B, S, H = int(sys.argv[1]) if len(sys.argv) > 1 else 64, 512, 16
qkv = (torch.randn(B, S, 3 * H * 64, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_(True)
dout = (torch.randn(B, S, H * 64, device="cuda") * 0.5).to(torch.bfloat16)
flops_fwd = 4.0 * B * H * S * S * 64


def ours_fwd():
    with torch.no_grad():
        return packed_attention(qkv, H)


def ours_fwd_bwd():
    qkv.grad = None
    packed_attention(qkv, H).backward(dout)


def sdpa(x):
    q, k, v = x.view(B, S, 3, H, 64).permute(2, 0, 3, 1, 4)
    return F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(B, S, H * 64)


def ref_fwd():
    with torch.no_grad():
        return sdpa(qkv)


def ref_fwd_bwd():
    qkv.grad = None
    sdpa(qkv).backward(dout)


res = {"B": B, "S": S, "H": H}
for name, fn in (("ours_fwd", ours_fwd), ("sdpa_fwd", ref_fwd), ("ours_fwd_bwd", ours_fwd_bwd), ("sdpa_fwd_bwd", ref_fwd_bwd)):
    res[name + "_ms"] = round(time_fn(fn), 4)
res["ours_fwd_tflops"] = round(flops_fwd / res["ours_fwd_ms"] / 1e9, 1)
res["sdpa_fwd_tflops"] = round(flops_fwd / res["sdpa_fwd_ms"] / 1e9, 1)
res["ours_bwd_ms"] = round(res["ours_fwd_bwd_ms"] - res["ours_fwd_ms"], 4)
res["sdpa_bwd_ms"] = round(res["sdpa_fwd_bwd_ms"] - res["sdpa_fwd_ms"], 4)
print(json.dumps(res))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/attention_bench.json", "w"), indent=1)
