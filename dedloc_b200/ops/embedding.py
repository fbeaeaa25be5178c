"""ALBERT's embedding front end and the 128-wide LayerNorm as single kernels (csrc/embedding.cu).

`embed_layer_norm` = LayerNorm(word[ids] + type[token_types] + pos[0..S)) in one kernel; its backward does the
LayerNorm backward and the three embedding backwards at once, adding into the fp32 `main_grad` rows with red.add
(torch's embedding backward sorts the indices and was the second-largest kernel of a micro-step after the GEMMs).
`layer_norm128` is the same LayerNorm for a [M, 128] input (MLM head). Both need bf16 tensors whose parameters carry
fp32 `main_grad` buffers (the fused-optimizer layout); otherwise the callers use PyTorch.
"""
from __future__ import annotations

import torch

from .. import lib as native


def _st(t):
    return torch.cuda.current_stream(t.device).cuda_stream


def _has_main_grads(*params) -> bool:
    return all(getattr(p, "main_grad", None) is not None for p in params)


def embed_supported(emb_module, input_ids: torch.Tensor) -> bool:
    w = emb_module.word_embeddings.weight
    return (input_ids.is_cuda and input_ids.dtype == torch.int64 and input_ids.dim() == 2
            and w.dtype == torch.bfloat16 and w.shape[1] == 128 and native.available()
            and _has_main_grads(w, emb_module.position_embeddings.weight, emb_module.token_type_embeddings.weight,
                                emb_module.LayerNorm.weight, emb_module.LayerNorm.bias))


class _EmbedLNFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, tts, word, pos, typ, gamma, beta, eps, pad_id, grads):
        B, S = ids.shape
        M, E = B * S, word.shape[1]
        if S > pos.shape[0]:
            raise ValueError(f"sequence length {S} exceeds the {pos.shape[0]} position embeddings")
        if tts.shape != ids.shape:
            raise ValueError("input_ids and token_type_ids must have one shape")
        ids1 = ids.reshape(-1).contiguous()
        tts1 = tts.reshape(-1).contiguous()
        y = torch.empty(M, E, dtype=torch.bfloat16, device=ids.device)
        mean = torch.empty(M, dtype=torch.float32, device=ids.device)
        rstd = torch.empty(M, dtype=torch.float32, device=ids.device)
        with torch.cuda.device(ids.device):
            native.check(native.lib().dedloc_embed_ln_fwd(
                ids1.data_ptr(), tts1.data_ptr(), word.data_ptr(), pos.data_ptr(), typ.data_ptr(), gamma.data_ptr(),
                beta.data_ptr(), y.data_ptr(), mean.data_ptr(), rstd.data_ptr(), M, S, E, float(eps), _st(ids)),
                "embed_ln_fwd")
        ctx.save_for_backward(ids1, tts1, word, pos, typ, gamma, mean, rstd)
        ctx.grads, ctx.dims, ctx.pad_id = grads, (B, S, E), pad_id
        return y.view(B, S, E)

    @staticmethod
    def backward(ctx, dy):
        ids1, tts1, word, pos, typ, gamma, mean, rstd = ctx.saved_tensors
        B, S, E = ctx.dims
        gword, gpos, gtype, dgamma, dbeta = ctx.grads
        dy2 = dy.reshape(B * S, E)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        with torch.cuda.device(dy.device):
            native.check(native.lib().dedloc_embed_ln_bwd(
                dy2.data_ptr(), ids1.data_ptr(), tts1.data_ptr(), word.data_ptr(), pos.data_ptr(), typ.data_ptr(),
                gamma.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gword.data_ptr(), gpos.data_ptr(), gtype.data_ptr(),
                dgamma.data_ptr(), dbeta.data_ptr(), B * S, S, E, int(ctx.pad_id), int(typ.shape[0]), _st(dy)),
                "embed_ln_bwd")
        return (None,) * 10


def embed_layer_norm(emb_module, input_ids: torch.Tensor, token_type_ids: torch.Tensor) -> torch.Tensor:
    """The whole of AlbertEmbeddings.forward (without dropout) in one kernel. Gradients go to `.main_grad`."""
    w, p, t = (emb_module.word_embeddings.weight, emb_module.position_embeddings.weight,
               emb_module.token_type_embeddings.weight)
    ln = emb_module.LayerNorm
    pad = emb_module.word_embeddings.padding_idx
    grads = (w.main_grad, p.main_grad, t.main_grad, ln.weight.main_grad, ln.bias.main_grad)
    if token_type_ids.dtype != torch.int64 or token_type_ids.shape != input_ids.shape:      # the kernel reads int64
        token_type_ids = token_type_ids.to(torch.int64).expand_as(input_ids)
    # the output must require grad so that autograd calls backward: hang it on the word table
    return _EmbedLNFn.apply(input_ids, token_type_ids, w, p, t, ln.weight, ln.bias, ln.eps,
                            -1 if pad is None else pad, grads)


class _LN128Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, eps, wgrad, bgrad):
        x2 = x.reshape(-1, 128)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M = x2.shape[0]
        y = torch.empty_like(x2)
        mean = torch.empty(M, dtype=torch.float32, device=x.device)
        rstd = torch.empty(M, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            native.check(native.lib().dedloc_ln128_fwd(x2.data_ptr(), gamma.data_ptr(), beta.data_ptr(), y.data_ptr(),
                                                       mean.data_ptr(), rstd.data_ptr(), M, float(eps), _st(x)),
                         "ln128_fwd")
        ctx.save_for_backward(x2, gamma, mean, rstd)
        ctx.grads, ctx.shape = (wgrad, bgrad), x.shape
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, gamma, mean, rstd = ctx.saved_tensors
        dy2 = dy.reshape(-1, 128)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = torch.empty_like(x2)
        with torch.cuda.device(dy.device):
            native.check(native.lib().dedloc_ln128_bwd(dy2.data_ptr(), x2.data_ptr(), gamma.data_ptr(), mean.data_ptr(),
                                                       rstd.data_ptr(), dx.data_ptr(), ctx.grads[0].data_ptr(),
                                                       ctx.grads[1].data_ptr(), x2.shape[0], _st(dy)), "ln128_bwd")
        return dx.view(ctx.shape), None, None, None, None, None


def ln128_supported(x: torch.Tensor, ln: torch.nn.LayerNorm) -> bool:
    return (x.is_cuda and x.dtype == torch.bfloat16 and x.shape[-1] == 128 and ln.weight.dtype == torch.bfloat16
            and native.available() and _has_main_grads(ln.weight, ln.bias))


def layer_norm128(x: torch.Tensor, ln: torch.nn.LayerNorm) -> torch.Tensor:
    return _LN128Fn.apply(x, ln.weight, ln.bias, ln.eps, ln.weight.main_grad, ln.bias.main_grad)
