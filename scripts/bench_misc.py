"""Bandwidth-bound helper kernels next to their floor and to the PyTorch expression of the same op.

    python scripts/bench_misc.py [--out gpurun_out/misc_bench.json]

For every op: device time of our kernel (CUDA events, 20 launches after 5 warm-up launches, a 512 MB buffer written
between launches so that nothing is served from L2), the bytes the op MUST move (each input read once, each output
written once), the time those bytes take at the measured copy bandwidth (`MEASURED_PEAKS.json`, fallback 6551 GB/s)
-- the floor -- and the time of the plain PyTorch formulation with the same flush.

    sinkhorn            SwAV assignments, 3 iterations, scores [4096 + queue 3840, 3000] -> [4096, 3000]
                        (vissl/losses/swav_loss.py:177-244); floor = one read of the scores + one write of Q (the
                        kernel keeps the matrix on chip between iterations only if it fits: it does not, so the honest
                        floor for 3 iterations is 1 + 2 x 3 passes; both are printed)
    swav_ce             swapped-prediction cross-entropy, forward (swav_loss.py:275-302)
    accumulate_clip     acc += coef * grad, grad = 0 on the ALBERT-large flat vector (fp32 and bf16 gradient)
    sumsq               global gradient norm
    layer_norm 1024     forward, 65536 rows
    cross_entropy       fused CE forward over [4928, 30000] bf16 logits
    colsum              bias-gradient column sums of a [65536, 4096] bf16 matrix
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F


def hbm_gbs() -> float:
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    try:
        peaks = json.load(open(path))
        for key in ("hbm_gbs", "copy_gbs", "hbm_copy_gbs"):
            if key in peaks:
                return float(peaks[key])
    except Exception:
        pass
    return 6551.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/misc_bench.json")
    args = ap.parse_args()
    from dedloc_b200.ops import cross_entropy, flat_ops, layernorm, sinkhorn, swav_ops

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    flush = torch.empty(512 * 2 ** 20, dtype=torch.uint8, device=dev)
    peak = hbm_gbs()

    def timed(fn, iters=20, warmup=5):
        for _ in range(warmup):
            fn()
        ts = []
        for _ in range(iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        return ts[len(ts) // 2]

    rows = []

    def row(name, ours, torch_fn, nbytes, note=""):
        t_ours, t_torch = timed(ours), timed(torch_fn)
        floor_ms = nbytes / (peak * 1e9) * 1e3
        rows.append({"op": name, "ours_ms": round(t_ours, 4), "torch_ms": round(t_torch, 4),
                     "must_move_mb": round(nbytes / 1e6, 1), "floor_ms": round(floor_ms, 4),
                     "of_hbm_peak": round(floor_ms / t_ours, 3), "vs_torch": round(t_torch / t_ours, 2), "note": note})
        print(json.dumps(rows[-1]), flush=True)

    g = torch.Generator(device="cpu").manual_seed(0)
    # ---- Sinkhorn
    n, n_out, K = 4096 + 3840, 4096, 3000
    scores = (torch.randn(n, K, generator=g) * 0.3).to(dev)
    one_pass = n * K * 4 + n_out * K * 4
    row("sinkhorn_fp32_3iters", lambda: sinkhorn.sinkhorn(scores, 0.03, 3, n_out),
        lambda: sinkhorn.sinkhorn_reference(scores, 0.03, 3, n_out), one_pass,
        note=f"floor = scores read once + Q written once; with the matrix re-read per half-iteration: "
             f"{(n * K * 4 * 7 + n_out * K * 4) / (peak * 1e9) * 1e3:.4f} ms")
    # ---- swapped-prediction CE (forward)
    bs, crops = 512, 8
    sc = (torch.randn(crops * bs, K, generator=g) * 0.3).to(dev).bfloat16()
    q = torch.softmax(torch.randn(2, bs, K, generator=g), dim=-1).to(dev)

    def ce_torch():
        logp = F.log_softmax(sc.float().view(crops, bs, K) / 0.1, dim=-1)
        tot = 0.0
        for i in range(2):
            mask = torch.ones(crops, dtype=torch.bool, device=dev)
            mask[i] = False
            tot = tot - (q[i].unsqueeze(0) * logp[mask]).sum(-1).mean()
        return tot

    row("swav_ce_fwd_bf16", lambda: swav_ops.swav_cross_entropy(sc, q, (0, 1), crops, 0.1), ce_torch,
        sc.numel() * 2 + q.numel() * 4 * (crops - 1) // 1,
        note="q is re-read for each of the 7 other crops (L2 hits when it fits)")
    # ---- flat gradient helpers on the ALBERT-large vector
    nel = 17_848_320 // 2048 * 2048
    acc = torch.zeros(nel, device=dev)
    g32 = torch.randn(nel, device=dev)
    g16 = torch.randn(nel, device=dev).bfloat16()
    nrm = torch.ones(1, device=dev)

    def acc_torch32():
        acc.add_(g32, alpha=0.5)
        g32.zero_()

    def acc_torch16():
        acc.add_(g16, alpha=0.5)
        g16.zero_()

    row("accumulate_clip_fp32", lambda: flat_ops.accumulate_clip(acc, g32, nrm, 1.0, zero_grad=True), acc_torch32,
        nel * 16, note="read acc + grad, write acc + zeroed grad")
    row("accumulate_clip_bf16", lambda: flat_ops.accumulate_clip(acc, g16, nrm, 1.0, zero_grad=True), acc_torch16,
        nel * 12)
    scratch = torch.empty(4096, device=dev)
    row("sumsq_fp32", lambda: flat_ops.sumsq(g32, scratch, nrm), lambda: torch.sum(g32 * g32), nel * 4,
        note="torch materialises the squares")
    # ---- LayerNorm 1024 forward
    x = torch.randn(65536, 1024, generator=g).to(dev).bfloat16()
    ln = torch.nn.LayerNorm(1024).to(dev).bfloat16()
    for p in ln.parameters():
        p.main_grad = torch.zeros_like(p, dtype=torch.float32)
    with torch.no_grad():
        row("layer_norm_1024_fwd", lambda: layernorm.layer_norm(x, ln), lambda: F.layer_norm(x, (1024,), ln.weight, ln.bias, ln.eps),
            x.numel() * 4)
    # ---- fused cross-entropy forward
    logits = (torch.randn(4928, 30000, generator=g) * 2).to(dev).bfloat16()
    labels = torch.randint(0, 30000, (4928,), generator=g).to(dev)
    with torch.no_grad():
        row("cross_entropy_fwd_30000", lambda: cross_entropy.cross_entropy_rows(logits, labels),
            lambda: F.cross_entropy(logits.float(), labels, reduction="none"), logits.numel() * 2)
    # ---- bias-gradient column sums
    dy = torch.randn(65536, 4096, generator=g).to(dev).bfloat16()
    out = torch.zeros(4096, device=dev)
    row("colsum_65536x4096", lambda: layernorm.colsum_accumulate(dy, out), lambda: out.add_(dy.sum(dim=0, dtype=torch.float32)),
        dy.numel() * 2)
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    json.dump({"hbm_gbs": peak, "rows": rows}, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
