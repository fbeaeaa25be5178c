"""Flat gradient-path kernels (csrc/flat_ops.cu): global L2 norm, clipped accumulation, wire encoders."""
from __future__ import annotations

from typing import Optional

import torch

from .. import lib as native
from ..utils.compression import CompressionType


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def sumsq(x: torch.Tensor, scratch: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """out[0] = sum(x^2) for a flat bf16/fp32 CUDA tensor whose length is a multiple of 8. Asynchronous."""
    assert x.is_cuda and x.is_contiguous() and x.dtype in (torch.bfloat16, torch.float32)
    with torch.cuda.device(x.device):
        native.check(native.lib().dedloc_sumsq(x.data_ptr(), int(x.dtype == torch.bfloat16), x.numel(),
                                               scratch.data_ptr(), scratch.numel(), out.data_ptr(), _stream(x)),
                     "sumsq")
    return out


def accumulate_clip(acc: torch.Tensor, grad: torch.Tensor, normsq: Optional[torch.Tensor], max_norm: float,
                    zero_grad: bool = True) -> None:
    """acc += min(1, max_norm/(sqrt(normsq)+1e-6)) * grad; optionally grad = 0 (SURVEY C9 + H3 step 2)."""
    assert acc.dtype == torch.float32 and acc.numel() == grad.numel()
    with torch.cuda.device(acc.device):
        native.check(native.lib().dedloc_accumulate_clip(
            acc.data_ptr(), grad.data_ptr(), int(grad.dtype == torch.bfloat16), acc.numel(),
            normsq.data_ptr() if normsq is not None else None, float(max_norm), int(zero_grad), _stream(acc)),
            "accumulate_clip")


def encode(codec: CompressionType, src: torch.Tensor, scale: float, payload: torch.Tensor,
           aux: torch.Tensor) -> None:
    """payload/aux = codec(src * scale); src is flat fp32 with a length that is a multiple of 2048."""
    assert src.dtype == torch.float32 and src.is_contiguous()
    with torch.cuda.device(src.device):
        native.check(native.lib().dedloc_encode(int(codec), src.data_ptr(), src.numel(), float(scale),
                                                payload.data_ptr(), aux.data_ptr(), _stream(src)), "encode")
