"""Property tests of the wire codecs (utils/compression.py, the oracle of csrc/codec.cuh): for arbitrary finite
inputs every codec round-trips through bytes, keeps its stated error bound, and is idempotent where it should be.
(the reference has no codec tests; hypothesis is how this repo covers the input space.)"""
import numpy as np
import pytest
import torch
from hypothesis import given, settings
from hypothesis import strategies as st

from dedloc_b200.utils import compression as comp
import os

from dedloc_b200.utils.compression import CompressionType

# deterministic in CI (same examples every run); HYPOTHESIS_CAMPAIGN=N runs N random examples per test instead
_N = int(os.environ.get("HYPOTHESIS_CAMPAIGN", "0"))
SETTINGS = dict(max_examples=_N or 40, deadline=None, derandomize=not _N)


@st.composite
def vectors(draw):
    n_chunks = draw(st.integers(1, 3))
    scale = draw(st.sampled_from([1e-6, 1e-3, 1.0, 30.0]))
    seed = draw(st.integers(0, 2 ** 31 - 1))
    kind = draw(st.sampled_from(["normal", "uniform", "sparse", "constant", "heavy_tail"]))
    g = torch.Generator().manual_seed(seed)
    n = n_chunks * comp.CHUNK
    if kind == "normal":
        x = torch.randn(n, generator=g)
    elif kind == "uniform":
        x = torch.rand(n, generator=g) * 2 - 1
    elif kind == "sparse":
        x = torch.randn(n, generator=g) * (torch.rand(n, generator=g) < 0.05)
    elif kind == "constant":
        x = torch.full((n,), float(draw(st.floats(-2, 2))))
    else:
        x = torch.randn(n, generator=g) ** 3
    return (x * scale).float()


def _roundtrip(x, codec):
    enc = comp.encode(x, codec)
    wire = comp.from_bytes(comp.to_bytes(enc))
    assert wire.codec == codec and wire.numel == enc.numel
    y, y2 = comp.decode(enc), comp.decode(wire)
    assert torch.equal(y, y2), "serialisation changed the decoded values"
    assert y.shape == x.shape and torch.isfinite(y).all()
    return y


@given(vectors())
@settings(**SETTINGS)
def test_none_is_exact(x):
    assert torch.equal(_roundtrip(x, CompressionType.NONE), x)


@given(vectors())
@settings(**SETTINGS)
def test_float16_and_bfloat16_relative_error(x):
    y = _roundtrip(x, CompressionType.FLOAT16)
    big = x.abs() >= 6.2e-5            # fp16 normal range
    assert torch.all((y - x).abs()[big] <= 2 ** -11 * x.abs()[big] * 1.001)
    assert torch.all((y - x).abs()[~big] <= 2 ** -24 * 1.001)       # subnormal spacing
    assert torch.equal(_roundtrip(y, CompressionType.FLOAT16), y)    # idempotent
    yb = _roundtrip(x, CompressionType.BFLOAT16)
    assert torch.all((yb - x).abs() <= 2 ** -8 * x.abs() * 1.001 + 1e-38)
    assert torch.equal(_roundtrip(yb, CompressionType.BFLOAT16), yb)


@given(vectors())
@settings(**SETTINGS)
def test_float16_clamps_instead_of_overflowing(x):
    y = _roundtrip(x * 1e6, CompressionType.FLOAT16)
    assert y.abs().max() <= comp.FP16_MAX


@given(vectors())
@settings(**SETTINGS)
def test_meanstd_error_scales_with_the_chunk_std(x):
    y = _roundtrip(x, CompressionType.MEANSTD_16BIT)
    for c in range(x.numel() // comp.CHUNK):
        xs, ys = x[c * comp.CHUNK:(c + 1) * comp.CHUNK], y[c * comp.CHUNK:(c + 1) * comp.CHUNK]
        sd = max(float(xs.std(unbiased=False)), 1e-6)
        dev = (xs - xs.mean()).abs()
        # fp16 of the normalised value: 2^-11 relative to |x - mean|, plus fp16 subnormal spacing in units of the std
        assert torch.all((ys - xs).abs() <= 2 ** -10 * dev + 2 ** -22 * sd + 1e-7 * xs.abs().max())


@given(vectors())
@settings(**SETTINGS)
def test_blockwise_fp8_error_scales_with_the_block_maximum(x):
    y = _roundtrip(x, CompressionType.BLOCKWISE_FP8)
    xb, yb = x.view(-1, comp.FP8_BLOCK), y.view(-1, comp.FP8_BLOCK)
    amax = xb.abs().amax(dim=1, keepdim=True)
    # e4m3: 3 mantissa bits -> 2^-4 relative; values far below the block maximum fall into the subnormal grid
    assert torch.all((yb - xb).abs() <= 2 ** -4 * xb.abs() + amax * 2 ** -9 + 1e-30)


@pytest.mark.parametrize("codec", [CompressionType.UNIFORM_8BIT, CompressionType.QUANTILE_8BIT])
@given(x=vectors())
@settings(**SETTINGS)
def test_8bit_codebooks(codec, x):
    y = _roundtrip(x, codec)
    n_levels = len(torch.unique(y))
    assert n_levels <= 256 * ((x.numel() + comp.CODEC_BLOCK - 1) // comp.CODEC_BLOCK)
    # a codebook entry is the MEAN of its bucket, so the block mean survives and the error never exceeds the spread
    assert abs(float(y.mean()) - float(x.mean())) <= 1e-3 * float(x.abs().max()) + 1e-12
    tiny = 1e-4 * float(x.abs().max()) + 1e-12           # fp32 rounding of the bucket means
    assert float((y - x).abs().max()) <= float(x.max() - x.min()) + tiny
    if float(x.std()) > 0:
        # every value is replaced by the mean of its bucket, so the error cannot exceed the spread of the data (the
        # within-bucket variance is at most the total variance); heavy tails that UNIFORM clips at 3 sigma get close
        assert float((y - x).pow(2).mean().sqrt()) <= float(x.std()) + tiny


@given(vectors(), st.floats(0.05, 4.0))
@settings(**SETTINGS)
def test_scale_is_applied_before_encoding(x, s):
    for codec in (CompressionType.NONE, CompressionType.FLOAT16, CompressionType.BLOCKWISE_FP8):
        a = comp.decode(comp.encode(x, codec, scale=s))
        b = comp.decode(comp.encode(x * s, codec))
        assert torch.allclose(a, b, rtol=1e-6, atol=1e-12)
