"""Wire codecs of the averaging path: CPU/PyTorch reference implementations.

Behavioural spec: hivemind 0.9.9 `CompressionType` as used at /root/reference/albert/run_trainer.py:256 and
/root/reference/swav/ClassyVision/classy_vision/optim/sgd_collaborative.py:150. These functions are
(1) the CPU data path of `parallel/cpu_backend.py`, and (2) the oracle for the CUDA codecs in csrc/codec.cuh,
which use the same block geometry (2048-element chunks, 65536-element codebook blocks, 128-element fp8 blocks).
"""
from __future__ import annotations

import enum
from dataclasses import dataclass
from typing import Optional

import torch

CHUNK = 2048
CODEC_BLOCK = 32 * CHUNK
FP8_BLOCK = 128
FP16_MAX = 65504.0
FP8_MAX = 448.0


class CompressionType(enum.IntEnum):
    NONE = 0
    MEANSTD_16BIT = 1
    FLOAT16 = 2
    QUANTILE_8BIT = 3
    UNIFORM_8BIT = 4
    BFLOAT16 = 5        # B200-native addition
    BLOCKWISE_FP8 = 6   # B200-native addition (e4m3 + fp32 scale per 128 elements)

    @classmethod
    def Value(cls, name: str) -> "CompressionType":
        """Protobuf-enum style lookup kept for API parity with `hivemind.utils.CompressionType.Value`."""
        try:
            return cls[name.upper()]
        except KeyError as e:
            raise ValueError(f"unknown compression type {name!r}; choose from {[m.name for m in cls]}") from e

    @classmethod
    def Name(cls, value: int) -> str:
        return cls(value).name


@dataclass
class Encoded:
    codec: CompressionType
    payload: torch.Tensor            # uint8 view of the encoded bytes
    aux: Optional[torch.Tensor]      # fp32 side information
    numel: int

    def nbytes(self) -> int:
        return self.payload.numel() + (0 if self.aux is None else 4 * self.aux.numel())


def payload_bytes_per_elem(codec: CompressionType) -> int:
    if codec == CompressionType.NONE:
        return 4
    if codec in (CompressionType.FLOAT16, CompressionType.BFLOAT16, CompressionType.MEANSTD_16BIT):
        return 2
    return 1


def aux_floats(codec: CompressionType, n: int) -> int:
    if codec == CompressionType.MEANSTD_16BIT:
        return 2 * (n // CHUNK)
    if codec in (CompressionType.UNIFORM_8BIT, CompressionType.QUANTILE_8BIT):
        return 256 * ((n + CODEC_BLOCK - 1) // CODEC_BLOCK)
    if codec == CompressionType.BLOCKWISE_FP8:
        return n // FP8_BLOCK
    return 0


def pad_to_chunk(x: torch.Tensor) -> torch.Tensor:
    n = x.numel()
    pad = (-n) % CHUNK
    if pad:
        x = torch.cat([x.reshape(-1), x.new_zeros(pad)])
    return x.reshape(-1)


def _bucket_means(x: torch.Tensor, q: torch.Tensor) -> torch.Tensor:
    sums = torch.zeros(256, dtype=torch.float32).index_add_(0, q.long(), x)
    cnt = torch.zeros(256, dtype=torch.float32).index_add_(0, q.long(), torch.ones_like(x))
    return torch.where(cnt > 0, sums / cnt.clamp_min(1), torch.zeros_like(sums))


def encode(x: torch.Tensor, codec: CompressionType, scale: float = 1.0) -> Encoded:
    """Encode a flat fp32 tensor whose length is a multiple of CHUNK (use `pad_to_chunk`)."""
    codec = CompressionType(codec)
    x = (x.detach().to(torch.float32).reshape(-1).cpu() * scale).contiguous()
    n = x.numel()
    assert n % CHUNK == 0, "pad the vector to a multiple of 2048 elements first"
    if codec == CompressionType.NONE:
        return Encoded(codec, x.view(torch.uint8).clone(), None, n)
    if codec == CompressionType.FLOAT16:
        y = x.clamp(-FP16_MAX, FP16_MAX).to(torch.float16)
        return Encoded(codec, y.view(torch.uint8).clone(), None, n)
    if codec == CompressionType.BFLOAT16:
        return Encoded(codec, x.to(torch.bfloat16).view(torch.uint8).clone(), None, n)
    if codec == CompressionType.MEANSTD_16BIT:
        xc = x.view(-1, CHUNK)
        mean = xc.mean(dim=1, keepdim=True)
        var = ((xc * xc).mean(dim=1, keepdim=True) - mean * mean).clamp_min(0)
        sd = var.sqrt().clamp_min(1e-6)
        y = ((xc - mean) / sd).to(torch.float16)
        aux = torch.cat([mean, sd], dim=1).reshape(-1).contiguous()
        return Encoded(codec, y.reshape(-1).view(torch.uint8).clone(), aux, n)
    if codec == CompressionType.BLOCKWISE_FP8:
        xb = x.view(-1, FP8_BLOCK)
        amax = xb.abs().amax(dim=1, keepdim=True)
        # (a block of denormals: amax / 448 underflows to 0 and x / 0 would be inf / NaN; FLT_MIN keeps it finite)
        s = torch.where(amax > 0, (amax / FP8_MAX).clamp_min(torch.finfo(torch.float32).tiny), torch.ones_like(amax))
        y = (xb / s).to(torch.float8_e4m3fn)
        return Encoded(codec, y.reshape(-1).view(torch.uint8).clone(), s.reshape(-1).contiguous(), n)
    # 8-bit codebook codecs, one codebook per 65536-element block
    payload = torch.empty(n, dtype=torch.uint8)
    nblocks = (n + CODEC_BLOCK - 1) // CODEC_BLOCK
    aux = torch.zeros(nblocks * 256, dtype=torch.float32)
    for b in range(nblocks):
        blk = x[b * CODEC_BLOCK:(b + 1) * CODEC_BLOCK]
        if codec == CompressionType.UNIFORM_8BIT:
            mean = blk.mean()
            sd = ((blk * blk).mean() - mean * mean).clamp_min(0).sqrt()
            step = 6.0 * sd / 256.0
            if step > 0:
                q = torch.round((blk - mean) / step) + 128
            else:
                q = torch.full_like(blk, 128.0)
            q = q.clamp(0, 255).to(torch.uint8)
        else:  # QUANTILE_8BIT: exact quantile borders (the CUDA encoder estimates them from a histogram)
            srt, _ = torch.sort(blk)
            idx = (torch.arange(1, 256, dtype=torch.float64) * blk.numel() / 256).long().clamp(0, blk.numel() - 1)
            borders = srt[idx]
            q = torch.bucketize(blk, borders, right=True).clamp(0, 255).to(torch.uint8)
        payload[b * CODEC_BLOCK:b * CODEC_BLOCK + blk.numel()] = q
        aux[b * 256:(b + 1) * 256] = _bucket_means(blk, q)
    return Encoded(codec, payload, aux, n)


def decode(enc: Encoded) -> torch.Tensor:
    codec, n = enc.codec, enc.numel
    if codec == CompressionType.NONE:
        return enc.payload.view(torch.float32).clone()
    if codec == CompressionType.FLOAT16:
        return enc.payload.view(torch.float16).to(torch.float32)
    if codec == CompressionType.BFLOAT16:
        return enc.payload.view(torch.bfloat16).to(torch.float32)
    if codec == CompressionType.MEANSTD_16BIT:
        y = enc.payload.view(torch.float16).to(torch.float32).view(-1, CHUNK)
        aux = enc.aux.view(-1, 2)
        return (y * aux[:, 1:2] + aux[:, 0:1]).reshape(-1)
    if codec == CompressionType.BLOCKWISE_FP8:
        y = enc.payload.view(torch.float8_e4m3fn).to(torch.float32).view(-1, FP8_BLOCK)
        return (y * enc.aux.view(-1, 1)).reshape(-1)
    out = torch.empty(n, dtype=torch.float32)
    nblocks = (n + CODEC_BLOCK - 1) // CODEC_BLOCK
    for b in range(nblocks):
        q = enc.payload[b * CODEC_BLOCK:(b + 1) * CODEC_BLOCK].long()
        out[b * CODEC_BLOCK:b * CODEC_BLOCK + q.numel()] = enc.aux[b * 256:(b + 1) * 256][q]
    return out


def to_bytes(enc: Encoded) -> bytes:
    """Serialise for the socket transport of the CPU backend."""
    import msgpack

    return msgpack.packb({
        "c": int(enc.codec), "n": enc.numel, "p": enc.payload.numpy().tobytes(),
        "a": b"" if enc.aux is None else enc.aux.numpy().tobytes(),
    }, use_bin_type=True)


def from_bytes(raw: bytes) -> Encoded:
    import msgpack
    import numpy as np

    # bytes from another peer: every field is checked before anything is allocated from it
    try:
        d = msgpack.unpackb(raw, raw=False)
        codec, n, p, a = CompressionType(d["c"]), d["n"], d["p"], d["a"]
    except Exception as e:
        raise ValueError(f"malformed encoded tensor: {e!r}") from e
    if not isinstance(n, int) or isinstance(n, bool) or n <= 0 or n % CHUNK or not isinstance(p, (bytes, bytearray)) \
            or not isinstance(a, (bytes, bytearray)):
        raise ValueError("malformed encoded tensor: bad field types or element count")
    if len(p) != n * payload_bytes_per_elem(codec) or len(a) != 4 * aux_floats(codec, n):
        raise ValueError("malformed encoded tensor: payload / side information do not match the element count")
    payload = torch.from_numpy(np.frombuffer(p, dtype=np.uint8).copy())
    aux = torch.from_numpy(np.frombuffer(a, dtype=np.float32).copy()) if a else None
    return Encoded(codec, payload, aux, n)
