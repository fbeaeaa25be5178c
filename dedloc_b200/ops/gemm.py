"""tcgen05/TMEM/TMA GEMM (csrc/gemm_tcgen05.cu) and the autograd functions built on it.

`gemm()` is the raw launcher. `linear()` / `linear_gelu()` / `linear_residual()` are differentiable and cover
ALBERT's QKV, attention-out, FFN-up (+gelu_new) and FFN-down projections plus the SwAV head (SURVEY C2, C4,
C5, C6, C13): forward is one kernel with the bias / activation / residual fused in the epilogue, dgrad is one
kernel (optionally with gelu' fused), and wgrad accumulates split-K partial products straight into an fp32
main-gradient buffer with red.add -- which is exactly what ALBERT's 24-fold weight sharing needs.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from .. import lib as native

_SM_COUNT = {}


def _sms(device) -> int:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _SM_COUNT:
        _SM_COUNT[idx] = torch.cuda.get_device_properties(idx).multi_processor_count
    return _SM_COUNT[idx]


def gemm(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, *, a_mn_major: bool = False,
         b_mn_major: bool = False, bias: Optional[torch.Tensor] = None, gelu: bool = False,
         dgelu_aux: Optional[torch.Tensor] = None, residual: Optional[torch.Tensor] = None,
         pre_out: Optional[torch.Tensor] = None, accumulate: bool = False, alpha: float = 1.0,
         save_dgelu: bool = False, mul_aux: Optional[torch.Tensor] = None,
         split_k: int = 1, block_n: int = 0, max_ctas: int = 0, cta_pair: int = 0) -> torch.Tensor:
    """out[M,N] = epilogue(A . B^T).

    a: [M,K] (K-major) or [K,M] (a_mn_major); b: [N,K] (K-major, nn.Linear weight) or [K,N] (b_mn_major).
    out: bf16 [M,N]; or fp32 [M,N] (store, or `accumulate=True` for out += with red.add and optional split-K).
    """
    assert a.is_cuda and a.dim() == 2 and b.dim() == 2 and out.dim() == 2
    assert a.dtype == b.dtype and a.dtype in (torch.bfloat16, torch.float16)
    assert a.stride(1) == 1 and b.stride(1) == 1 and out.stride(1) == 1
    M, N = out.shape
    K = a.shape[0] if a_mn_major else a.shape[1]
    assert (a.shape[1] if a_mn_major else a.shape[0]) == M, "A shape mismatch"
    assert (b.shape == (K, N)) if b_mn_major else (b.shape == (N, K)), "B shape mismatch"
    g = native.GemmArgs()
    g.A, g.B, g.C = a.data_ptr(), b.data_ptr(), out.data_ptr()
    g.M, g.N, g.K = M, N, K
    g.lda, g.ldb, g.ldc = a.stride(0), b.stride(0), out.stride(0)
    g.a_mn_major, g.b_mn_major = int(a_mn_major), int(b_mn_major)
    g.is_fp16 = int(a.dtype == torch.float16)
    epi = 0
    if bias is not None:
        epi |= native.EPI_BIAS
        if bias.dtype == torch.float32:
            epi |= native.EPI_BIAS_F32
        else:
            assert bias.dtype == torch.bfloat16
        g.bias = bias.data_ptr()
    if gelu:
        epi |= native.EPI_GELU
    if dgelu_aux is not None:
        epi |= native.EPI_DGELU
        g.aux_in, g.ld_aux = dgelu_aux.data_ptr(), dgelu_aux.stride(0)
    if residual is not None:
        epi |= native.EPI_RESIDUAL
        g.residual, g.ld_residual = residual.data_ptr(), residual.stride(0)
    if mul_aux is not None:
        assert dgelu_aux is None
        epi |= native.EPI_MUL_AUX
        g.aux_in, g.ld_aux = mul_aux.data_ptr(), mul_aux.stride(0)
    if pre_out is not None:
        epi |= native.EPI_SAVE_PRE
        if save_dgelu:          # pre_out receives gelu'(pre) (one tanh for both), for a multiply-only backward
            assert gelu
            epi |= native.EPI_SAVE_DGELU
        g.pre_out, g.ld_pre = pre_out.data_ptr(), pre_out.stride(0)
    g.epilogue = epi
    if out.dtype == torch.bfloat16:
        g.out_kind = native.OUT_BF16
        assert not accumulate
    else:
        assert out.dtype == torch.float32
        g.out_kind = native.OUT_F32_ATOMIC if accumulate else native.OUT_F32
    g.alpha = alpha
    g.split_k, g.block_n, g.max_ctas, g.cta_pair = split_k, block_n, max_ctas, cta_pair
    with torch.cuda.device(a.device):
        native.check(native.lib().dedloc_gemm(C.byref(g), torch.cuda.current_stream(a.device).cuda_stream), "gemm")
    return out


def pick_split_k(m_out: int, n_out: int, k_red: int, sms: int, block_n: int = 256) -> int:
    """Split the reduction so that (output tiles x splits) fills whole waves of CTA pairs: weight gradients have few
    output tiles (e.g. 3072 x 1024 = 48 tiles of 256 x 256 for 74 pairs). The smallest split within 2 % of the best
    wave efficiency wins (every extra split is another fp32 red.add pass over the output)."""
    pairs = max(sms // 2, 1)
    tiles = ((m_out + 255) // 256) * ((n_out + 255) // 256)
    kb = max(k_red // 64, 1)
    best_s, best_eff = 1, 0.0
    for s in range(1, min(kb, 16) + 1):
        work = tiles * s
        eff = work / (((work + pairs - 1) // pairs) * pairs)
        if eff > best_eff + 0.02:
            best_s, best_eff = s, eff
    return best_s


def dgrad(dy2: torch.Tensor, weight: torch.Tensor, weight_t: Optional[torch.Tensor], dx: torch.Tensor, **epilogue):
    """dx[M, in] = dy[M, out] . W[out, in] (+ epilogue). With a K-major copy of W^T ([in, out], refreshed by the model
    once per forward -- cheap for ALBERT's shared layer) the B operand is K-major; MN-major operands cost 6-13 % each in
    the tcgen05 kernel (profiles/README.md)."""
    if weight_t is not None:
        return gemm(dy2, weight_t, dx, **epilogue)
    return gemm(dy2, weight, dx, b_mn_major=True, **epilogue)


# ---------------------------------------------------------------- weight / bias gradients off the critical path
# Nothing downstream of a layer's backward consumes dW / db before the optimizer, so (opt-in) they are launched on a
# low-priority side stream: the bandwidth-bound kernels of the critical path (LayerNorm backward, bias-gradient
# column sums, attention) then share the SMs with the compute-bound weight-gradient GEMMs instead of queueing behind
# them. All of them accumulate into the same shared-layer main_grad buffers (red.add) and stay ordered on the one
# side stream. The caller must `join_wgrad_stream()` before reading the gradients (PeerTrainer does).
_OVERLAP = {"enabled": False, "streams": {}, "refs": []}


def overlap_wgrad(enabled: bool) -> None:
    _OVERLAP["enabled"] = bool(enabled)


def _side_stream(device) -> "torch.cuda.Stream":
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _OVERLAP["streams"]:
        _OVERLAP["streams"][idx] = torch.cuda.Stream(device=idx, priority=0)   # main work runs at default priority too;
    return _OVERLAP["streams"][idx]                                             # kernels already resident keep their SMs


def _run_wgrad(device, fn, *keep_alive) -> None:
    """Run `fn` (weight / bias gradient launches) on the side stream when overlap is on, else inline."""
    if not _OVERLAP["enabled"]:
        fn()
        return
    side, cur = _side_stream(device), torch.cuda.current_stream(device)
    side.wait_stream(cur)
    with torch.cuda.stream(side):
        fn()
    # the operands were allocated on the main stream: keep them referenced until the join, so that the caching
    # allocator (also inside a CUDA-graph capture) cannot hand their memory to a later main-stream tensor
    _OVERLAP["refs"].append(keep_alive)


def join_wgrad_stream(device=None) -> None:
    """Make the current stream wait for every outstanding side-stream gradient launch."""
    for idx, side in _OVERLAP["streams"].items():
        if device is None or idx == (device.index if device.index is not None else torch.cuda.current_device()):
            torch.cuda.current_stream(idx).wait_stream(side)
    _OVERLAP["refs"].clear()


class _LinearFn(torch.autograd.Function):
    """y = act(x W^T + b) (+ residual). Weight/bias gradients are ACCUMULATED into `.main_grad` (fp32)."""

    @staticmethod
    def forward(ctx, x, weight, bias, act: str, residual, wgrad_out, bgrad_out):
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M, K = x2.shape
        N = weight.shape[0]
        out = torch.empty(M, N, dtype=x.dtype, device=x.device)
        pre = None
        if act == "gelu":
            pre = torch.empty(M, N, dtype=x.dtype, device=x.device)
        res2 = residual.reshape(M, N) if residual is not None else None
        gemm(x2, weight, out, bias=bias, gelu=(act == "gelu"), pre_out=pre, residual=res2)
        ctx.save_for_backward(x2, weight, pre)
        ctx.act = act
        ctx.has_bias = bias is not None
        ctx.has_res = residual is not None
        ctx.wgrad_out, ctx.bgrad_out = wgrad_out, bgrad_out
        ctx.weight_t = getattr(weight, "weight_t", None)
        ctx.x_shape = x.shape
        return out.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, weight, pre = ctx.saved_tensors
        M, K = x2.shape
        N = weight.shape[0]
        dy2 = dy.reshape(M, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dres = dy if ctx.has_res else None
        if ctx.act == "gelu":
            # dpre = dy * gelu'(pre): small elementwise pass, then both GEMMs consume dpre
            from .elementwise import dgelu_mul
            dy2 = dgelu_mul(dy2, pre)
        # dgrad: dx[M,K] = dy[M,N] . W[N,K]  -> B operand is W stored [N,K] = [K_red, N_out] MN-major
        dx = torch.empty(M, K, dtype=x2.dtype, device=x2.device)
        dgrad(dy2, weight, ctx.weight_t, dx)
        # wgrad: dW[N,K] += dy^T[N,M] . x[M,K]  -> both operands MN-major, reduction over tokens
        wg, bg, has_bias = ctx.wgrad_out, ctx.bgrad_out, ctx.has_bias
        sk = pick_split_k(N, K, M, _sms(x2.device))

        def grads():
            gemm(dy2, x2, wg, a_mn_major=True, b_mn_major=True, accumulate=True, split_k=sk)
            if has_bias:
                _bias_grad(dy2, bg)

        _run_wgrad(x2.device, grads, dy2, x2)
        return dx.view(ctx.x_shape), None, None, None, dres, None, None


def _bias_grad(dy2: torch.Tensor, out: torch.Tensor) -> None:
    if dy2.dtype == torch.bfloat16 and dy2.shape[1] % 8 == 0:
        from .layernorm import colsum_accumulate

        colsum_accumulate(dy2, out)
    else:
        out.add_(dy2.sum(dim=0, dtype=torch.float32))


class _FFNFn(torch.autograd.Function):
    """y = gelu_new(x W1^T + b1) W2^T + b2 (+ residual): ALBERT's feed-forward block (SURVEY C5 + C6) as ONE
    autograd node, so that gelu' is fused into the epilogue of the dgrad GEMM of the second linear
    (dpre = (dy W2) * gelu'(pre)) instead of being a separate pass over the [tokens, 4096] activation."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, residual, g_w1, g_b1, g_w2, g_b2):
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M, K = x2.shape
        I, N = w1.shape[0], w2.shape[0]
        # the epilogue saves gelu'(pre) instead of pre: the backward epilogue is then a multiply (no tanh per element
        # in an epilogue that already is the bottleneck of that GEMM)
        pre = torch.empty(M, I, dtype=x.dtype, device=x.device)
        h = torch.empty(M, I, dtype=x.dtype, device=x.device)
        gemm(x2, w1, h, bias=b1, gelu=True, pre_out=pre, save_dgelu=True)
        out = torch.empty(M, N, dtype=x.dtype, device=x.device)
        res2 = residual.reshape(M, N) if residual is not None else None
        gemm(h, w2, out, bias=b2, residual=res2)
        ctx.save_for_backward(x2, w1, w2, pre, h)
        ctx.grads = (g_w1, g_b1, g_w2, g_b2)
        ctx.w1t, ctx.w2t = getattr(w1, "weight_t", None), getattr(w2, "weight_t", None)
        ctx.has_res, ctx.x_shape = residual is not None, x.shape
        # residual IS the input (ALBERT: FFN(a) + a): its gradient is added in the epilogue of the last dgrad GEMM
        ctx.res_is_x = residual is not None and residual.data_ptr() == x.data_ptr() and residual.shape == x.shape
        return out.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, w1, w2, pre, h = ctx.saved_tensors
        g_w1, g_b1, g_w2, g_b2 = ctx.grads
        M, K = x2.shape
        I, N = w1.shape[0], w2.shape[0]
        dy2 = dy.reshape(M, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        sms = _sms(x2.device)
        dpre = torch.empty(M, I, dtype=x2.dtype, device=x2.device)
        dgrad(dy2, w2, ctx.w2t, dpre, mul_aux=pre)                               # (dy W2) * gelu'(pre), saved in forward

        def grads2():
            gemm(dy2, h, g_w2, a_mn_major=True, b_mn_major=True, accumulate=True, split_k=pick_split_k(N, I, M, sms))
            _bias_grad(dy2, g_b2)

        _run_wgrad(x2.device, grads2, dy2, h)
        dx = torch.empty(M, K, dtype=x2.dtype, device=x2.device)
        dgrad(dpre, w1, ctx.w1t, dx, residual=dy2 if ctx.res_is_x else None)

        def grads1():
            gemm(dpre, x2, g_w1, a_mn_major=True, b_mn_major=True, accumulate=True, split_k=pick_split_k(I, K, M, sms))
            _bias_grad(dpre, g_b1)

        _run_wgrad(x2.device, grads1, dpre, x2)
        dres = None if ctx.res_is_x else (dy if ctx.has_res else None)
        return dx.view(ctx.x_shape), None, None, None, None, dres, None, None, None, None


def ffn(x, lin1, lin2, residual=None):
    """Fused feed-forward block; `lin1` / `lin2` carry `.weight/.bias` with fp32 `.main_grad` buffers."""
    return _FFNFn.apply(x, lin1.weight, lin1.bias, lin2.weight, lin2.bias, residual, lin1.weight.main_grad,
                        lin1.bias.main_grad, lin2.weight.main_grad, lin2.bias.main_grad)


def linear(x, weight, bias=None, *, act: str = "none", residual=None, wgrad_out=None, bgrad_out=None):
    """Differentiable fused linear. `wgrad_out`/`bgrad_out` are fp32 buffers that receive dW / db (+=)."""
    assert wgrad_out is not None, "fused linear accumulates weight gradients into an fp32 main_grad buffer"
    return _LinearFn.apply(x, weight, bias, act, residual, wgrad_out, bgrad_out)
