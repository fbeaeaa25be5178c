"""tcgen05 GEMM numerics against an fp32 PyTorch reference: every operand layout, epilogue and tail case."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rand(shape, seed, scale=1.0, dtype=torch.bfloat16):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dtype).cuda()


def _check(out, ref, what, rtol=2e-2, atol=2e-2, tile=64):
    """Element-wise bound relative to the largest reference value, AND a per-tile bound relative to that tile's own
    norm: a wrong tile of small values passes the first check, not the second. The tile tolerance is 5x the rounding
    of the output format (bf16: 2^-9 per element; fp32 outputs only differ by accumulation order)."""
    is_fp32 = out.dtype == torch.float32
    out, ref = out.float(), ref.float()
    err = (out - ref).abs().max().item()
    denom = ref.abs().max().item() + 1e-6
    assert torch.allclose(out, ref, rtol=rtol, atol=atol * denom), f"{what}: max err {err:.4g} (ref max {denom:.4g})"
    if out.dim() != 2:
        return
    M, N = out.shape
    pm, pn = (-M) % tile, (-N) % tile
    d = torch.nn.functional.pad(out - ref, (0, pn, 0, pm))
    r = torch.nn.functional.pad(ref, (0, pn, 0, pm))
    tiles = lambda x: x.reshape((M + pm) // tile, tile, (N + pn) // tile, tile).pow(2).sum(dim=(1, 3)).sqrt()  # noqa: E731
    d_norm, r_norm = tiles(d), tiles(r)
    tol = 2e-3 if is_fp32 else 1e-2
    bad = d_norm > tol * r_norm + 1e-3 * denom * tile
    assert not bool(bad.any()), (f"{what}: tile {tuple(bad.nonzero()[0].tolist())} is off by "
                                 f"{float((d_norm / (r_norm + 1e-12))[bad].max()):.3g} of its own norm")


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(256, 256, 128), (384, 512, 320), (128, 128, 64), (200, 264, 192), (1024, 3072, 1024)])
def test_gemm_layouts(native, a_mn, b_mn, M, N, K):
    from dedloc_b200.ops.gemm import gemm

    if a_mn and M % 8:
        pytest.skip("MN-major A needs M % 8 == 0")
    a = _rand((K, M) if a_mn else (M, K), 1, 0.5)
    b = _rand((K, N) if b_mn else (N, K), 2, 0.5)
    a_log = a.float().t() if a_mn else a.float()
    b_log = b.float().t() if b_mn else b.float()
    ref = a_log @ b_log.t()
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    gemm(a, b, out, a_mn_major=a_mn, b_mn_major=b_mn)
    torch.cuda.synchronize()
    _check(out, ref, f"layout a_mn={a_mn} b_mn={b_mn} {M}x{N}x{K}")
    out32 = torch.empty(M, N, dtype=torch.float32, device="cuda")
    gemm(a, b, out32, a_mn_major=a_mn, b_mn_major=b_mn)
    _check(out32, ref, "fp32 out", rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(512, 512, 128), (768, 1024, 320), (1000, 776, 192), (4096, 3072, 1024)])
def test_gemm_cta_pair_layouts(native, a_mn, b_mn, M, N, K):
    """cta_group::2 kernel (256 x 256 tile per CTA pair): every operand layout, M/N tails, bias+gelu epilogue."""
    from dedloc_b200.ops.elementwise import gelu_new
    from dedloc_b200.ops.gemm import gemm

    a = _rand((K, M) if a_mn else (M, K), 21, 0.5)
    b = _rand((K, N) if b_mn else (N, K), 22, 0.5)
    bias = _rand((N,), 23, 0.5)
    ref = (a.float().t() if a_mn else a.float()) @ (b.float() if b_mn else b.float().t())
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    gemm(a, b, out, a_mn_major=a_mn, b_mn_major=b_mn, cta_pair=2)
    torch.cuda.synchronize()
    _check(out, ref, f"2cta a_mn={a_mn} b_mn={b_mn} {M}x{N}x{K}")
    gemm(a, b, out, a_mn_major=a_mn, b_mn_major=b_mn, cta_pair=2, bias=bias, gelu=True)
    _check(out, gelu_new(ref + bias.float()), "2cta bias+gelu")
    acc = torch.ones(M, N, dtype=torch.float32, device="cuda")
    gemm(a, b, acc, a_mn_major=a_mn, b_mn_major=b_mn, cta_pair=2, accumulate=True, split_k=2 if K >= 128 else 1)
    _check(acc, 1.0 + ref, "2cta split-k accumulate", rtol=1e-3, atol=1e-3)


def test_gemm_block_n_128_and_fp16(native):
    from dedloc_b200.ops.gemm import gemm

    M, N, K = 512, 128, 1024
    a, b = _rand((M, K), 3, 0.3), _rand((N, K), 4, 0.3)
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    gemm(a, b, out)
    _check(out, a.float() @ b.float().t(), "bn128")
    out = torch.empty(M, 512, dtype=torch.bfloat16, device="cuda")
    b2 = _rand((512, K), 5, 0.3)
    gemm(a, b2, out, block_n=128)
    _check(out, a.float() @ b2.float().t(), "bn128 forced")
    ah, bh = a.to(torch.float16), b2.to(torch.float16)
    out32 = torch.empty(M, 512, dtype=torch.float32, device="cuda")
    gemm(ah, bh, out32)
    _check(out32, ah.float() @ bh.float().t(), "fp16 inputs", rtol=1e-3, atol=1e-3)


def test_gemm_epilogues(native):
    from dedloc_b200.ops.elementwise import gelu_new, gelu_new_grad
    from dedloc_b200.ops.gemm import gemm

    M, N, K = 640, 1024, 256
    a, b = _rand((M, K), 6, 0.3), _rand((N, K), 7, 0.3)
    bias = _rand((N,), 8, 0.5)
    res = _rand((M, N), 9, 1.0)
    base = a.float() @ b.float().t()
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")

    gemm(a, b, out, bias=bias)
    _check(out, base + bias.float(), "bias bf16")
    gemm(a, b, out, bias=bias.float())
    _check(out, base + bias.float(), "bias fp32")
    pre = torch.empty_like(out)
    gemm(a, b, out, bias=bias, gelu=True, pre_out=pre)
    _check(pre, base + bias.float(), "saved pre-activation")
    _check(out, gelu_new(base + bias.float()), "gelu")
    gemm(a, b, out, bias=bias, residual=res)
    _check(out, base + bias.float() + res.float(), "bias+residual")
    aux = _rand((M, N), 10, 1.0)
    gemm(a, b, out, dgelu_aux=aux)
    _check(out, base * gelu_new_grad(aux.float()), "dgelu")
    gemm(a, b, out, alpha=0.125)
    _check(out, base * 0.125, "alpha")
    # derivative saved by the forward epilogue (one tanh for gelu and gelu'), multiply-only backward epilogue
    dact = torch.empty_like(out)
    for pair in (0, 1, 2):     # auto, 1-CTA, CTA pair
        gemm(a, b, out, bias=bias, gelu=True, pre_out=dact, save_dgelu=True, cta_pair=pair)
        _check(out, gelu_new(base + bias.float()), f"gelu with saved derivative (cta_pair={pair})")
        _check(dact, gelu_new_grad(base + bias.float()), f"saved gelu' (cta_pair={pair})")
        gemm(a, b, out, mul_aux=aux, cta_pair=pair)
        _check(out, base * aux.float(), f"mul_aux (cta_pair={pair})")
    # wide tile path of the CTA-pair kernel (coalesced row-operand prefetch needs full 128-column halves)
    M2, N2 = 512, 512
    a2, b2, aux2, res2 = _rand((M2, K), 11, 0.3), _rand((N2, K), 12, 0.3), _rand((M2, N2), 13, 1.0), _rand((M2, N2), 14, 1.0)
    out2 = torch.empty(M2, N2, dtype=torch.bfloat16, device="cuda")
    base2 = a2.float() @ b2.float().t()
    gemm(a2, b2, out2, mul_aux=aux2, cta_pair=2)
    _check(out2, base2 * aux2.float(), "mul_aux pair kernel")
    gemm(a2, b2, out2, dgelu_aux=aux2, cta_pair=2)
    _check(out2, base2 * gelu_new_grad(aux2.float()), "dgelu pair kernel")
    gemm(a2, b2, out2, residual=res2, cta_pair=2)
    _check(out2, base2 + res2.float(), "residual pair kernel")


def test_gemm_wgrad_splitk_accumulate(native):
    from dedloc_b200.ops.gemm import gemm

    tokens, N, K = 4096, 512, 256
    dy, x = _rand((tokens, N), 11, 0.1), _rand((tokens, K), 12, 0.1)
    wg = torch.ones(N, K, dtype=torch.float32, device="cuda")
    ref = 1.0 + dy.float().t() @ x.float()
    gemm(dy, x, wg, a_mn_major=True, b_mn_major=True, accumulate=True, split_k=8)
    _check(wg, ref, "wgrad split-k", rtol=1e-3, atol=1e-3)
    gemm(dy, x, wg, a_mn_major=True, b_mn_major=True, accumulate=True, split_k=1)
    _check(wg, ref + dy.float().t() @ x.float(), "wgrad accumulate twice", rtol=1e-3, atol=1e-3)


def test_fused_linear_autograd(native):
    from dedloc_b200.ops.elementwise import gelu_new
    from dedloc_b200.ops.gemm import linear

    B, S, K, N = 4, 128, 256, 512
    x = _rand((B, S, K), 13, 0.5).requires_grad_(True)
    w, bias = _rand((N, K), 14, 0.05), _rand((N,), 15, 0.1)
    res = _rand((B, S, N), 16, 0.5).requires_grad_(True)
    dy = _rand((B, S, N), 17, 0.1)
    for act in ("none", "gelu"):
        wg = torch.zeros(N, K, device="cuda")
        bg = torch.zeros(N, device="cuda")
        x.grad = None
        res.grad = None
        y = linear(x, w, bias, act=act, residual=res, wgrad_out=wg, bgrad_out=bg)
        y.backward(dy)
        xr = x.detach().float().requires_grad_(True)
        wr = w.float().requires_grad_(True)
        br = bias.float().requires_grad_(True)
        rr = res.detach().float().requires_grad_(True)
        pre = xr @ wr.t() + br
        yr = (gelu_new(pre) if act == "gelu" else pre) + rr
        yr.backward(dy.float())
        _check(y, yr, f"fwd {act}")
        _check(x.grad, xr.grad, f"dx {act}", rtol=3e-2, atol=3e-2)
        _check(wg, wr.grad, f"dw {act}", rtol=3e-2, atol=3e-2)
        _check(bg, br.grad, f"db {act}", rtol=3e-2, atol=3e-2)
        _check(res.grad, rr.grad, f"dres {act}")


def test_gemm_rejects_bad_arguments(native):
    from dedloc_b200.lib import NativeError
    from dedloc_b200.ops.gemm import gemm

    a, b = _rand((128, 72), 1), _rand((128, 72), 2)
    out = torch.empty(128, 128, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(NativeError):
        gemm(a[:, :68], b[:, :68], out)  # K not a multiple of 8 (rows of the K tail are not 16-byte multiples)


def test_gemm_k_tail_is_zero_filled(native):
    """K need not be a multiple of the 64-wide K block: TMA zero-fills the out-of-bounds part of the last box (the
    vocabulary reduction of the MLM decoder's dgrad has K = 30000)."""
    from dedloc_b200.ops.gemm import gemm

    for pair in (1, 2):
        M, N, K = 512, 384, 200
        a, b = _rand((M, K), 21, 0.3), _rand((N, K), 22, 0.3)
        out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        gemm(a, b, out, cta_pair=pair)
        _check(out, a.float() @ b.float().t(), f"K tail, K-major (cta_pair={pair})")
        bt = b.t().contiguous()          # [K, N]
        gemm(a, bt, out, b_mn_major=True, cta_pair=pair)
        _check(out, a.float() @ b.float().t(), f"K tail, MN-major B (cta_pair={pair})")
        at = a.t().contiguous()          # [K, M]
        wg = torch.zeros(M, N, dtype=torch.float32, device="cuda")
        gemm(at, bt, wg, a_mn_major=True, b_mn_major=True, accumulate=True, split_k=2, cta_pair=pair)
        _check(wg, a.float() @ b.float().t(), f"K tail, both MN-major split-K (cta_pair={pair})", rtol=2e-2, atol=2e-2)
