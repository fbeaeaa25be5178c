"""Host-side pieces of the compute path that need no GPU: the flat parameter layout, the CPU codecs (the oracle of
the device encoders / in-kernel decoders), the split-K heuristic of the weight-gradient GEMMs and the PyTorch
fallbacks every op takes for CPU tensors."""
import math

import pytest
import torch
import torch.nn.functional as F

from dedloc_b200.parallel.flat import CHUNK, FlatLayout
from dedloc_b200.utils import compression as comp
from dedloc_b200.utils.compression import CompressionType


def test_flat_layout_is_chunk_aligned_and_round_trips():
    shapes = [(30000, 128), (1024,), (3, 5), (), (4096, 1024)]
    layout = FlatLayout([torch.Size(s) for s in shapes])
    assert layout.total % CHUNK == 0 and layout.n_chunks == layout.total // CHUNK
    assert all(o % CHUNK == 0 for o in layout.offsets)          # a chunk never straddles two tensors
    g = torch.Generator().manual_seed(0)
    tensors = [torch.randn(s, generator=g) for s in shapes]
    flat = layout.pack(tensors)
    for a, b in zip(layout.unpack(flat), tensors):
        assert torch.equal(a, b)
    # padding stays zero and every chunk knows its tensor
    mask = torch.ones(layout.total, dtype=torch.bool)
    for o, n in zip(layout.offsets, layout.numels):
        mask[o:o + n] = False
    assert flat[mask].abs().sum() == 0
    for t in range(layout.n_tensors):
        lo, hi = int(layout.tensor_cbeg[t]), int(layout.tensor_cbeg[t + 1])
        assert (layout.chunk_tid[lo:hi] == t).all()


@pytest.mark.parametrize("shares", [[1, 1, 1, 1], [1, 0, 1], [0, 0, 1], [3, 1], [1] * 8])
def test_shard_ranges_cover_every_chunk_exactly_once(shares):
    layout = FlatLayout([torch.Size((1000, 77)), torch.Size((5,)), torch.Size((4096, 3))])
    ranges = layout.shard_ranges(shares)
    assert len(ranges) == len(shares)
    covered = 0
    for (b, e), s in zip(ranges, shares):
        assert b <= e
        if s == 0:
            assert b == e                                        # client-mode members own nothing
        covered += e - b
    assert covered == layout.n_chunks
    owned = [(b, e) for (b, e) in ranges if e > b]
    assert owned[0][0] == 0 and owned[-1][1] == layout.n_chunks
    for (_, e0), (b1, _) in zip(owned, owned[1:]):
        assert e0 == b1


@pytest.mark.parametrize("codec,tol", [(CompressionType.NONE, 0.0), (CompressionType.FLOAT16, 1e-3),
                                       (CompressionType.BFLOAT16, 4e-3), (CompressionType.MEANSTD_16BIT, 1e-3),
                                       (CompressionType.UNIFORM_8BIT, 0.06), (CompressionType.QUANTILE_8BIT, 0.06),
                                       (CompressionType.BLOCKWISE_FP8, 0.05)])
def test_cpu_codecs_round_trip(codec, tol):
    g = torch.Generator().manual_seed(1)
    x = comp.pad_to_chunk(torch.randn(3 * CHUNK + 17, generator=g) * 0.02)
    enc = comp.encode(x, codec)
    y = comp.decode(comp.from_bytes(comp.to_bytes(enc)))
    assert y.shape == x.shape and y.dtype == torch.float32
    err = ((y - x).pow(2).mean().sqrt() / x.pow(2).mean().sqrt()).item()     # relative RMS error of the codec
    assert err <= tol + 1e-12, (codec, err)
    assert enc.payload.numel() == x.numel() * comp.payload_bytes_per_elem(codec)


def test_pick_split_k_fills_waves_with_few_splits():
    from dedloc_b200.ops.gemm import pick_split_k

    for (m_out, n_out, k_red) in [(3072, 1024, 32768), (1024, 1024, 32768), (4096, 1024, 65536), (30000, 128, 4928)]:
        s = pick_split_k(m_out, n_out, k_red, 148)
        tiles = math.ceil(m_out / 256) * math.ceil(n_out / 256)
        assert 1 <= s <= min(16, max(k_red // 64, 1))
        work = tiles * s
        eff = work / (math.ceil(work / 74) * 74)
        best = max((tiles * c) / (math.ceil(tiles * c / 74) * 74) for c in range(1, min(16, k_red // 64) + 1))
        assert eff >= best - 0.021, (m_out, n_out, s, eff, best)  # within the heuristic's 2 % of the best fill
    assert pick_split_k(16384, 4096, 1024, 148) == 1 or pick_split_k(16384, 4096, 1024, 148) <= 2


def test_ops_fall_back_to_pytorch_on_cpu():
    from dedloc_b200.ops.attention import supported as attention_supported
    from dedloc_b200.ops.cross_entropy import cross_entropy_rows
    from dedloc_b200.ops.layernorm import layer_norm

    g = torch.Generator().manual_seed(2)
    logits = torch.randn(6, 50, generator=g)
    labels = torch.randint(0, 50, (6,), generator=g)
    assert torch.allclose(cross_entropy_rows(logits, labels), F.cross_entropy(logits, labels, reduction="none"))
    ln = torch.nn.LayerNorm(256)
    x = torch.randn(4, 256, generator=g)
    assert torch.allclose(layer_norm(x, ln), ln(x))
    assert not attention_supported(torch.zeros(1, 128, 3 * 2 * 64, dtype=torch.bfloat16), 2)   # CPU tensor


def test_public_api_surface_of_the_reference_is_importable():
    """every symbol the reference's scripts import from hivemind has a counterpart with the same call
    shape (constructor keywords, method names, attributes)."""
    import inspect

    from dedloc_b200 import DHT, CollaborativeOptimizer, CompressionType, get_dht_time
    from dedloc_b200.collab.auth import AccessToken, TokenAuthorizerBase  # noqa: F401
    from dedloc_b200.dht import BytesWithPublicKey, RecordValidatorBase, RSASignatureValidator, SchemaValidator  # noqa: F401
    from dedloc_b200.dht.validation import RSAPrivateKey, RSAPublicKey  # noqa: F401
    from dedloc_b200.p2p import P2P  # noqa: F401
    from dedloc_b200.utils.logging import get_logger

    dht_kw = inspect.signature(DHT.__init__).parameters
    assert all(k in dht_kw for k in ("start", "initial_peers", "listen", "listen_on", "endpoint", "record_validators"))
    assert {"key", "value", "expiration_time", "subkey", "return_future"} <= set(inspect.signature(DHT.store).parameters)
    assert {"key", "latest"} <= set(inspect.signature(DHT.get).parameters)
    opt_kw = inspect.signature(CollaborativeOptimizer.__init__).parameters
    for k in ("opt", "dht", "prefix", "target_batch_size", "batch_size_per_step", "scheduler", "compression_type",
              "compression", "throughput", "bandwidth", "client_mode", "auxiliary", "allow_state_sharing", "verbose", "start",
              "averaging_expiration", "averaging_timeout", "metadata_expiration", "target_group_size",
              "min_refresh_period", "max_refresh_period", "default_refresh_period", "expected_drift_peers",
              "expected_drift_rate", "performance_ema_alpha", "listen_on"):
        assert k in opt_kw, k
    for attr in ("step", "zero_grad", "step_aux", "load_state_from_peers", "state_dict", "load_state_dict",
                 "is_synchronized", "collaboration_state", "shutdown"):
        assert hasattr(CollaborativeOptimizer, attr), attr
    assert [CompressionType.Value(n) for n in ("NONE", "FLOAT16", "MEANSTD_16BIT", "UNIFORM_8BIT", "QUANTILE_8BIT")]
    assert get_dht_time() > 0 and get_logger("dedloc.test") is get_logger("dedloc.test")
