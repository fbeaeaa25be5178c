"""Fused masked-LM cross-entropy on bf16 logits (csrc/cross_entropy.cu).

`cross_entropy_rows(logits, labels)` == `F.cross_entropy(logits.float(), labels, reduction="none")` for bf16 CUDA
logits, without the fp32 copy of the [rows, vocab] matrix: one pass for (loss, log-sum-exp), one pass for the
gradient, which is written as bf16 for the decoder's dgrad / wgrad GEMMs. Labels outside [0, V) give loss 0 and a
pure-softmax gradient scaled by the incoming row gradient (callers mask those rows with a zero weight). Reference
call site: HF AlbertForPreTraining's `CrossEntropyLoss()(prediction_scores.view(-1, V), labels.view(-1))`, driven by
/root/reference/albert/run_trainer.py:56-70.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .. import lib as native


def supported(logits: torch.Tensor) -> bool:
    return (logits.is_cuda and logits.dtype == torch.bfloat16 and logits.dim() == 2 and logits.stride(1) == 1
            and logits.stride(0) % 8 == 0 and logits.data_ptr() % 16 == 0 and native.available())


class _CrossEntropyRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels):
        rows, V = logits.shape
        loss = torch.empty(rows, dtype=torch.float32, device=logits.device)
        lse = torch.empty(rows, dtype=torch.float32, device=logits.device)
        labels = labels.to(torch.int64).contiguous()
        with torch.cuda.device(logits.device):
            native.check(native.lib().dedloc_ce_fwd(
                logits.data_ptr(), logits.stride(0), labels.data_ptr(), rows, V, loss.data_ptr(), lse.data_ptr(),
                torch.cuda.current_stream(logits.device).cuda_stream), "ce_fwd")
        ctx.save_for_backward(logits, labels, lse)
        return loss

    @staticmethod
    def backward(ctx, dloss):
        logits, labels, lse = ctx.saved_tensors
        rows, V = logits.shape
        dlogits = torch.empty_like(logits)
        scale = dloss.to(torch.float32).contiguous()
        with torch.cuda.device(logits.device):
            native.check(native.lib().dedloc_ce_bwd(
                logits.data_ptr(), logits.stride(0), labels.data_ptr(), lse.data_ptr(), scale.data_ptr(),
                dlogits.data_ptr(), dlogits.stride(0), rows, V,
                torch.cuda.current_stream(logits.device).cuda_stream), "ce_bwd")
        return dlogits, None


def cross_entropy_rows(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """Per-row cross-entropy (fp32). Falls back to PyTorch for CPU / non-bf16 / unaligned inputs."""
    if supported(logits):
        return _CrossEntropyRows.apply(logits, labels)
    return F.cross_entropy(logits.float(), labels, reduction="none")
