"""Sinkhorn-Knopp for SwAV: CUDA kernel wrapper (csrc/sinkhorn.cu) + the PyTorch restatement of the reference
(vissl/losses/swav_loss.py:177-244) that serves as CPU path and oracle."""
from __future__ import annotations

import torch

from .. import lib as native


@torch.no_grad()
def sinkhorn_reference(scores: torch.Tensor, epsilon: float, iters: int, n_out: int, world_size: int = 1) -> torch.Tensor:
    """scores: [n, K] (batch rows first, then queue rows). Returns assignments [n_out, K] (fp32).
    Same arithmetic as the reference: Q = exp(scores/eps - max)^T, normalise, `iters` x (rows to 1/K, columns to
    1/N), final column normalisation; the inf / zero guards are expressed as clamps (no host sync)."""
    s = scores.float() / epsilon
    Q = torch.exp(s - s.max()).t()                      # [K, n]
    Q = torch.nan_to_num(Q, posinf=torch.finfo(Q.dtype).max)
    Q /= Q.sum()
    K, n = Q.shape
    r = torch.full((K,), 1.0 / K, device=Q.device)
    c = torch.full((n,), 1.0 / (n * world_size), device=Q.device)
    for _ in range(iters):
        u = Q.sum(dim=1).clamp_min(1e-37)
        Q *= (r / u).unsqueeze(1)
        Q *= (c / Q.sum(dim=0).clamp_min(1e-37)).unsqueeze(0)
    Q = (Q / Q.sum(dim=0, keepdim=True).clamp_min(1e-37)).t()
    return Q[:n_out].contiguous()


_SCRATCH = {}


@torch.no_grad()
def sinkhorn(scores: torch.Tensor, epsilon: float, iters: int, n_out: int, world_size: int = 1) -> torch.Tensor:
    if not scores.is_cuda:
        return sinkhorn_reference(scores, epsilon, iters, n_out, world_size)
    assert scores.dim() == 2 and scores.stride(1) == 1 and scores.dtype in (torch.bfloat16, torch.float32)
    n, K = scores.shape
    out = torch.empty(n_out, K, dtype=torch.float32, device=scores.device)
    need = int(native.lib().dedloc_sinkhorn_scratch_floats(K, iters))
    key = (scores.device.index, need)
    if key not in _SCRATCH:
        _SCRATCH[key] = torch.empty(need, dtype=torch.float32, device=scores.device)
    scratch = _SCRATCH[key]
    with torch.cuda.device(scores.device):
        native.check(native.lib().dedloc_sinkhorn(
            scores.data_ptr(), int(scores.dtype == torch.bfloat16), scores.stride(0), n, K, n_out, float(epsilon),
            int(iters), float(world_size), out.data_ptr(), out.stride(0), scratch.data_ptr(),
            torch.cuda.current_stream(scores.device).cuda_stream), "sinkhorn")
    return out
