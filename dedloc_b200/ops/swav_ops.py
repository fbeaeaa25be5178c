"""SwAV head / loss on this repo's kernels: tcgen05 GEMMs for the projection head, the prototype
layer and the queue re-scoring (ops/gemm.py), a fused L2-normalise and a fused swapped-prediction cross-entropy
(csrc/swav_ops.cu). PyTorch restatements of the same ops are the CPU path and the test oracles (models/swav.py)."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from .. import lib as native


def _st(t):
    return torch.cuda.current_stream(t.device).cuda_stream


class _L2Norm128(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, eps):
        x2 = x.contiguous()
        y = torch.empty_like(x2)
        with torch.cuda.device(x.device):
            native.check(native.lib().dedloc_l2norm128_fwd(x2.data_ptr(), y.data_ptr(), x2.shape[0], float(eps), _st(x)),
                         "l2norm128_fwd")
        ctx.save_for_backward(x2)
        ctx.eps = eps
        return y

    @staticmethod
    def backward(ctx, dy):
        (x2,) = ctx.saved_tensors
        dy2 = dy.contiguous()
        dx = torch.empty_like(x2)
        with torch.cuda.device(dy.device):
            native.check(native.lib().dedloc_l2norm128_bwd(x2.data_ptr(), dy2.data_ptr(), dx.data_ptr(), x2.shape[0],
                                                           float(ctx.eps), _st(dy)), "l2norm128_bwd")
        return dx, None


def l2_normalize128(x: torch.Tensor, eps: float = 1e-12) -> torch.Tensor:
    """F.normalize(x, dim=1) for [M, 128] bf16."""
    assert x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and x.shape[1] == 128
    return _L2Norm128.apply(x, eps)


class _SwavCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scores, q, assign_crop, num_crops, temperature):
        rows, K = scores.shape
        bs = rows // num_crops
        assert scores.stride(1) == 1 and q.dtype == torch.float32 and q.is_contiguous() and q.shape[1:] == (bs, K)
        loss_row = torch.empty(rows, dtype=torch.float32, device=scores.device)
        lse = torch.empty(rows, dtype=torch.float32, device=scores.device)
        with torch.cuda.device(scores.device):
            native.check(native.lib().dedloc_swav_ce_fwd(
                scores.data_ptr(), int(scores.dtype == torch.bfloat16), scores.stride(0), q.data_ptr(), q.shape[0],
                assign_crop.data_ptr(), num_crops, bs, K, float(temperature), loss_row.data_ptr(), lse.data_ptr(),
                _st(scores)), "swav_ce_fwd")
        ctx.save_for_backward(scores, q, assign_crop, lse)
        ctx.dims = (num_crops, bs, K, temperature)
        # - 1 / (bs * (num_crops - 1) * n_assign) * sum over rows
        ctx.norm = -1.0 / (bs * (num_crops - 1) * q.shape[0])
        return loss_row.sum() * ctx.norm

    @staticmethod
    def backward(ctx, g):
        scores, q, assign_crop, lse = ctx.saved_tensors
        num_crops, bs, K, temperature = ctx.dims
        dscores = torch.empty_like(scores)
        scale = (g.float() * ctx.norm).reshape(1).contiguous()
        with torch.cuda.device(scores.device):
            native.check(native.lib().dedloc_swav_ce_bwd(
                scores.data_ptr(), int(scores.dtype == torch.bfloat16), scores.stride(0), q.data_ptr(), q.shape[0],
                assign_crop.data_ptr(), num_crops, bs, K, float(temperature), lse.data_ptr(), scale.data_ptr(),
                dscores.data_ptr(), _st(scores)), "swav_ce_bwd")
        return dscores, None, None, None, None


def swav_ce_supported(scores: torch.Tensor) -> bool:
    return (scores.is_cuda and scores.dtype in (torch.bfloat16, torch.float32) and scores.dim() == 2
            and scores.stride(1) == 1 and scores.shape[1] <= 4096 and native.available())


def swav_cross_entropy(scores: torch.Tensor, q: torch.Tensor, crops_for_assign: Sequence[int], num_crops: int,
                       temperature: float) -> torch.Tensor:
    """mean over assign crops i, other crops p and samples b of  - sum_k q_i[b,k] log_softmax(scores[p,b] / T)[k]
    (vissl/losses/swav_loss.py:275-302). scores: [num_crops * bs, K]; q: [len(crops_for_assign), bs, K] fp32."""
    assign = torch.tensor(list(crops_for_assign), dtype=torch.int32, device=scores.device)
    return _SwavCE.apply(scores, q.contiguous(), assign, num_crops, temperature)
