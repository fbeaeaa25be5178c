"""Attention on the packed QKV layout (csrc/attention_fwd.cu)."""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch

from .. import lib as native


def attention_fwd(qkv: torch.Tensor, num_heads: int, mask: Optional[torch.Tensor] = None,
                  need_lse: bool = True) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """qkv: [B, S, 3*H*64] bf16 contiguous (= [B, S, 3, H, 64]); mask: [B, S] (non-zero = attend).
    Returns (context [B, S, H*64] bf16, lse [B, H, S] fp32)."""
    assert qkv.is_cuda and qkv.dtype == torch.bfloat16 and qkv.is_contiguous() and qkv.dim() == 3
    B, S, three_hd = qkv.shape
    D = three_hd // (3 * num_heads)
    out = torch.empty(B, S, num_heads * D, dtype=torch.bfloat16, device=qkv.device)
    lse = torch.empty(B, num_heads, S, dtype=torch.float32, device=qkv.device) if need_lse else None
    m = None
    if mask is not None:
        m = mask if (mask.dtype == torch.uint8 and mask.is_contiguous()) else (mask != 0).to(torch.uint8).contiguous()
    with torch.cuda.device(qkv.device):
        native.check(native.lib().dedloc_attention_fwd(
            qkv.data_ptr(), m.data_ptr() if m is not None else None, out.data_ptr(),
            lse.data_ptr() if lse is not None else None, B, S, num_heads, D, 1.0 / math.sqrt(D),
            torch.cuda.current_stream(qkv.device).cuda_stream), "attention_fwd")
    return out, lse


_BWD_SCRATCH = {}


def attention_bwd(qkv: torch.Tensor, out: torch.Tensor, dout: torch.Tensor, lse: torch.Tensor, num_heads: int,
                  mask_u8: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Gradient w.r.t. the packed qkv ([B, S, 3*H*64] bf16)."""
    B, S, three_hd = qkv.shape
    D = three_hd // (3 * num_heads)
    if not dout.is_contiguous():
        dout = dout.contiguous()
    dqkv = torch.empty_like(qkv)
    need = int(native.lib().dedloc_attention_bwd_scratch_floats(S))
    key = (qkv.device.index, need)
    if key not in _BWD_SCRATCH:
        _BWD_SCRATCH[key] = torch.empty(need, dtype=torch.float32, device=qkv.device)
    with torch.cuda.device(qkv.device):
        native.check(native.lib().dedloc_attention_bwd(
            qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(),
            mask_u8.data_ptr() if mask_u8 is not None else None, dqkv.data_ptr(), _BWD_SCRATCH[key].data_ptr(),
            B, S, num_heads, D, 1.0 / math.sqrt(D), torch.cuda.current_stream(qkv.device).cuda_stream), "attention_bwd")
    return dqkv


class _PackedAttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, num_heads, mask_u8):
        out, lse = attention_fwd(qkv, num_heads, mask_u8)
        ctx.save_for_backward(qkv, out, lse, mask_u8 if mask_u8 is not None else torch.empty(0, device=qkv.device))
        ctx.num_heads, ctx.has_mask = num_heads, mask_u8 is not None
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, out, lse, mask_u8 = ctx.saved_tensors
        return attention_bwd(qkv, out, dout, lse, ctx.num_heads, mask_u8 if ctx.has_mask else None), None, None


def packed_attention(qkv: torch.Tensor, num_heads: int, mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Differentiable attention on the packed QKV projection: [B, S, 3*H*64] -> [B, S, H*64]."""
    mask_u8 = (mask != 0).to(torch.uint8).contiguous() if mask is not None else None
    if not qkv.is_contiguous():
        qkv = qkv.contiguous()
    return _PackedAttentionFn.apply(qkv, num_heads, mask_u8)


def supported(qkv: torch.Tensor, num_heads: int) -> bool:
    if not (qkv.is_cuda and qkv.dtype == torch.bfloat16 and qkv.dim() == 3):
        return False
    B, S, three_hd = qkv.shape
    return three_hd == 3 * num_heads * 64 and S % 128 == 0 and 128 <= S <= 512


class _AttentionBlockFn(torch.autograd.Function):
    """y = (attention(x Wqkv^T + bqkv)) Wo^T + bo + x : ALBERT's whole self-attention sub-layer before its LayerNorm
    (HF AlbertAttention.forward) as ONE autograd node. The point is the backward: the gradient that reaches `x`
    through the residual connection is added inside the epilogue of the QKV dgrad GEMM instead of by a separate
    elementwise pass over [tokens, hidden] (autograd's AccumulateGrad of the two branches)."""

    @staticmethod
    def forward(ctx, x, w_qkv, b_qkv, w_o, b_o, mask_u8, num_heads, g_wqkv, g_bqkv, g_wo, g_bo):
        from .gemm import gemm

        B, S, H = x.shape
        x2 = x.reshape(B * S, H)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        qkv = torch.empty(B * S, w_qkv.shape[0], dtype=x.dtype, device=x.device)
        gemm(x2, w_qkv, qkv, bias=b_qkv)
        att, lse = attention_fwd(qkv.view(B, S, -1), num_heads, mask_u8)
        y = torch.empty(B * S, w_o.shape[0], dtype=x.dtype, device=x.device)
        gemm(att.view(B * S, -1), w_o, y, bias=b_o, residual=x2)
        ctx.save_for_backward(x2, w_qkv, w_o, qkv, att, lse,
                              mask_u8 if mask_u8 is not None else torch.empty(0, device=x.device))
        ctx.has_mask, ctx.num_heads, ctx.shape = mask_u8 is not None, num_heads, (B, S, H)
        ctx.grads = (g_wqkv, g_bqkv, g_wo, g_bo)
        ctx.wqkv_t, ctx.wo_t = getattr(w_qkv, "weight_t", None), getattr(w_o, "weight_t", None)
        return y.view(B, S, -1)

    @staticmethod
    def backward(ctx, dy):
        from .gemm import _bias_grad, _run_wgrad, _sms, dgrad, gemm, pick_split_k

        x2, w_qkv, w_o, qkv, att, lse, mask_u8 = ctx.saved_tensors
        g_wqkv, g_bqkv, g_wo, g_bo = ctx.grads
        B, S, H = ctx.shape
        M = B * S
        sms = _sms(x2.device)
        dy2 = dy.reshape(M, -1)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        att2 = att.view(M, -1)
        datt = torch.empty_like(att2)
        dgrad(dy2, w_o, ctx.wo_t, datt)

        def grads_o():
            gemm(dy2, att2, g_wo, a_mn_major=True, b_mn_major=True, accumulate=True,
                 split_k=pick_split_k(w_o.shape[0], w_o.shape[1], M, sms))
            _bias_grad(dy2, g_bo)

        _run_wgrad(x2.device, grads_o, dy2, att2)
        dqkv = attention_bwd(qkv.view(B, S, -1), att, datt.view(B, S, -1), lse, ctx.num_heads,
                             mask_u8 if ctx.has_mask else None).view(M, -1)
        dx = torch.empty_like(x2)
        dgrad(dqkv, w_qkv, ctx.wqkv_t, dx, residual=dy2)         # + the gradient of the residual branch

        def grads_qkv():
            gemm(dqkv, x2, g_wqkv, a_mn_major=True, b_mn_major=True, accumulate=True,
                 split_k=pick_split_k(w_qkv.shape[0], w_qkv.shape[1], M, sms))
            _bias_grad(dqkv, g_bqkv)

        _run_wgrad(x2.device, grads_qkv, dqkv, x2)
        return dx.view(B, S, H), None, None, None, None, None, None, None, None, None, None


def attention_block(x: torch.Tensor, qkv_lin, out_lin, num_heads: int, mask: Optional[torch.Tensor] = None):
    """Fused self-attention sub-layer incl. the residual add: `qkv_lin` / `out_lin` carry `.weight/.bias` with fp32
    `.main_grad` buffers (weight gradients accumulate there, as in ops.gemm.linear)."""
    mask_u8 = (mask != 0).to(torch.uint8).contiguous() if mask is not None else None
    return _AttentionBlockFn.apply(x, qkv_lin.weight, qkv_lin.bias, out_lin.weight, out_lin.bias, mask_u8, num_heads,
                                   qkv_lin.weight.main_grad, qkv_lin.bias.main_grad, out_lin.weight.main_grad,
                                   out_lin.bias.main_grad)
