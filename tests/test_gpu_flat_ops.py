"""GPU numerics of the flat gradient-path kernels and wire codecs against fp32 PyTorch references."""
import pytest
import torch

from dedloc_b200.utils import compression as comp
from dedloc_b200.utils.compression import CompressionType

pytestmark = pytest.mark.gpu

N = 2048 * 70  # 2 full 8-bit codec blocks + a partial one


def _data(n=N, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, generator=g) * scale


def test_sumsq_and_accumulate_clip(native):
    from dedloc_b200.ops import flat_ops

    for dtype in (torch.bfloat16, torch.float32):
        x = _data(seed=1).to(dtype).cuda()
        scratch = torch.zeros(2048, device="cuda")
        out = torch.zeros(1, device="cuda")
        flat_ops.sumsq(x, scratch, out)
        ref = (x.float() ** 2).sum()
        assert torch.allclose(out[0], ref, rtol=1e-4), (out, ref)

        acc = _data(seed=2).cuda()
        acc_ref = acc.clone()
        g = x.clone()
        max_norm = 1.0
        flat_ops.accumulate_clip(acc, g, out, max_norm, zero_grad=True)
        coef = min(1.0, max_norm / (ref.sqrt().item() + 1e-6))
        acc_ref += coef * x.float()
        assert torch.allclose(acc, acc_ref, rtol=1e-5, atol=1e-6)
        assert g.abs().max().item() == 0.0
        # no clipping when max_norm <= 0
        g2 = x.clone()
        flat_ops.accumulate_clip(acc, g2, None, 0.0, zero_grad=False)
        acc_ref += x.float()
        assert torch.allclose(acc, acc_ref, rtol=1e-5, atol=1e-6)
        assert torch.equal(g2, x)
        # lengths that end inside an unrolled step / inside a block (any multiple of 8 is legal)
        for n in (8, 1000, 8 * 1237, 4 * 256 * 4 * 3 + 24):
            xs = x[:n].clone()
            a = torch.arange(n, dtype=torch.float32, device="cuda")
            guard = torch.full((n + 64,), 7.0, device="cuda")
            guard[:n] = a
            flat_ops.accumulate_clip(guard[:n], xs, None, 0.0, zero_grad=True)
            assert torch.allclose(guard[:n], a + x[:n].float(), rtol=1e-5, atol=1e-6), n
            assert float(guard[n:].min()) == 7.0 == float(guard[n:].max()), "wrote past the end"
            assert xs.abs().max().item() == 0.0


TOL = {  # max abs error / max abs value
    CompressionType.NONE: 0.0,
    CompressionType.FLOAT16: 1e-3,
    CompressionType.BFLOAT16: 8e-3,
    CompressionType.MEANSTD_16BIT: 2e-3,
    CompressionType.UNIFORM_8BIT: 0.45,   # values beyond 3 sigma collapse into the edge buckets by design
    CompressionType.QUANTILE_8BIT: 0.5,   # tails are coarse by construction; mean error is checked below
    CompressionType.BLOCKWISE_FP8: 0.07,
}


@pytest.mark.parametrize("codec", list(CompressionType))
def test_gpu_encode_matches_reference_decode(native, codec):
    """Encode on the GPU, decode with the CPU reference: the result must be as close to the input as the CPU
    codec's own round trip (same bucket geometry), and 16-bit/fp8 payloads must match the CPU encoder bit for
    bit up to rounding ties."""
    from dedloc_b200.ops import flat_ops

    x = _data(seed=3, scale=0.02)
    scale = 0.5
    payload = torch.zeros(4 * N, dtype=torch.uint8, device="cuda")
    aux = torch.zeros(max(comp.aux_floats(codec, N), 1), device="cuda")
    flat_ops.encode(codec, x.cuda(), scale, payload, aux)
    torch.cuda.synchronize()
    nbytes = N * comp.payload_bytes_per_elem(codec)
    enc = comp.Encoded(codec, payload[:nbytes].cpu(), aux[:comp.aux_floats(codec, N)].cpu() if comp.aux_floats(codec, N) else None, N)
    dec = comp.decode(enc)
    ref = x * scale
    err = (dec - ref).abs().max().item() / ref.abs().max().item()
    assert err <= TOL[codec] + 1e-9, (codec, err)
    cpu_dec = comp.decode(comp.encode(x, codec, scale))
    cpu_err = (cpu_dec - ref).abs().mean().item()
    gpu_err = (dec - ref).abs().mean().item()
    assert gpu_err <= 1.3 * cpu_err + 1e-9, (codec, gpu_err, cpu_err)


def test_nonfinite_propagates_through_codecs(native):
    from dedloc_b200.ops import flat_ops

    for codec in CompressionType:
        x = _data(seed=4)
        x[12345] = float("nan")
        payload = torch.zeros(4 * N, dtype=torch.uint8, device="cuda")
        aux = torch.zeros(max(comp.aux_floats(codec, N), 1), device="cuda")
        flat_ops.encode(codec, x.cuda(), 1.0, payload, aux)
        nbytes = N * comp.payload_bytes_per_elem(codec)
        na = comp.aux_floats(codec, N)
        dec = comp.decode(comp.Encoded(codec, payload[:nbytes].cpu(), aux[:na].cpu() if na else None, N))
        assert not torch.isfinite(dec).all(), codec
