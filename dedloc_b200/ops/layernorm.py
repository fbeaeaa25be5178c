"""LayerNorm forward/backward kernels (csrc/layernorm.cu) behind an autograd function. Parameter gradients are
accumulated into fp32 `main_grad` buffers, like the fused linears (ALBERT's shared layer is applied 24 times)."""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .. import lib as native


def _st(t):
    return torch.cuda.current_stream(t.device).cuda_stream


def supported(x: torch.Tensor, weight: torch.Tensor) -> bool:
    H = x.shape[-1]
    return (x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 and H % 256 == 0
            and 256 <= H <= 2048 and H // 256 in (1, 2, 3, 4, 8) and getattr(weight, "main_grad", None) is not None)


class _LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps, wgrad_out, bgrad_out):
        H = x.shape[-1]
        x2 = x.reshape(-1, H)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M = x2.shape[0]
        y = torch.empty_like(x2)
        mean = torch.empty(M, dtype=torch.float32, device=x.device)
        rstd = torch.empty(M, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            native.check(native.lib().dedloc_layernorm_fwd(x2.data_ptr(), weight.data_ptr(), bias.data_ptr(),
                                                           y.data_ptr(), mean.data_ptr(), rstd.data_ptr(), M, H,
                                                           float(eps), _st(x)), "layernorm_fwd")
        ctx.save_for_backward(x2, weight, mean, rstd)
        ctx.wgrad_out, ctx.bgrad_out, ctx.shape = wgrad_out, bgrad_out, x.shape
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, weight, mean, rstd = ctx.saved_tensors
        M, H = x2.shape
        dy2 = dy.reshape(M, H)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = torch.empty_like(x2)
        with torch.cuda.device(x2.device):
            native.check(native.lib().dedloc_layernorm_bwd(dy2.data_ptr(), x2.data_ptr(), weight.data_ptr(),
                                                           mean.data_ptr(), rstd.data_ptr(), dx.data_ptr(),
                                                           ctx.wgrad_out.data_ptr(), ctx.bgrad_out.data_ptr(), M, H,
                                                           _st(x2)), "layernorm_bwd")
        return dx.view(ctx.shape), None, None, None, None, None


def layer_norm(x: torch.Tensor, ln: torch.nn.LayerNorm, impl: str = "native") -> torch.Tensor:
    if impl == "native" and supported(x, ln.weight):
        return _LayerNormFn.apply(x, ln.weight, ln.bias, ln.eps, ln.weight.main_grad, ln.bias.main_grad)
    return F.layer_norm(x, ln.normalized_shape, ln.weight, ln.bias, ln.eps)


def colsum_accumulate(dy: torch.Tensor, out: torch.Tensor) -> None:
    """out[n] += sum_m dy[m, n] (bf16 -> fp32)."""
    assert dy.dim() == 2 and dy.stride(1) == 1 and dy.dtype == torch.bfloat16 and out.dtype == torch.float32
    M, N = dy.shape
    with torch.cuda.device(dy.device):
        native.check(native.lib().dedloc_colsum_accumulate(dy.data_ptr(), out.data_ptr(), M, N, dy.stride(0), _st(dy)),
                     "colsum_accumulate")
