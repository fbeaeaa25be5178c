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
        m = (mask != 0).to(torch.uint8).contiguous()
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
