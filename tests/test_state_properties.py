"""Property tests (hypothesis): the pickle-free state format survives arbitrary nesting / dtypes and never turns
arbitrary bytes into anything but a ValueError; the flat layout packs, views and shards consistently."""
import os

import msgpack
import torch
from hypothesis import given, settings
from hypothesis import strategies as st

from dedloc_b200.parallel.flat import CHUNK, FlatLayout
from dedloc_b200.utils.serialization import pack_state, unpack_state

# deterministic in CI (same examples every run); HYPOTHESIS_CAMPAIGN=N runs N random examples per test instead
_N = int(os.environ.get("HYPOTHESIS_CAMPAIGN", "0"))
SETTINGS = dict(max_examples=_N or 60, deadline=None, derandomize=not _N)
DTYPES = [torch.float32, torch.float64, torch.float16, torch.bfloat16, torch.int64, torch.int32, torch.int16,
          torch.int8, torch.uint8, torch.bool]


@st.composite
def tensors(draw):
    dtype = draw(st.sampled_from(DTYPES))
    shape = draw(st.lists(st.integers(0, 4), min_size=0, max_size=3))
    seed = draw(st.integers(0, 2 ** 31 - 1))
    g = torch.Generator().manual_seed(seed)
    base = torch.randint(-100, 100, shape or [], generator=g) if shape else torch.tensor(draw(st.integers(-100, 100)))
    return (base > 0) if dtype == torch.bool else base.abs().to(dtype) if dtype == torch.uint8 else base.to(dtype)


scalars = st.one_of(st.none(), st.booleans(), st.integers(-2 ** 40, 2 ** 40), st.floats(allow_nan=False),
                    st.text(max_size=8), st.binary(max_size=8))
keys = st.one_of(st.integers(-100, 100), st.text(max_size=6))
trees = st.recursive(st.one_of(scalars, tensors()),
                     lambda kids: st.one_of(st.lists(kids, max_size=4), st.lists(kids, max_size=3).map(tuple),
                                            st.dictionaries(keys, kids, max_size=4)), max_leaves=12)


def _same(a, b):
    if isinstance(a, torch.Tensor):
        return isinstance(b, torch.Tensor) and a.dtype == b.dtype and a.shape == b.shape and torch.equal(a, b)
    if isinstance(a, dict):
        return isinstance(b, dict) and list(a.keys()) == list(b.keys()) and all(_same(a[k], b[k]) for k in a)
    if isinstance(a, (list, tuple)):
        return type(a) is type(b) and len(a) == len(b) and all(_same(x, y) for x, y in zip(a, b))
    return type(a) is type(b) and a == b


@given(trees)
@settings(**SETTINGS)
def test_state_round_trip(tree):
    wire = msgpack.packb(pack_state(tree), use_bin_type=True)
    back = unpack_state(msgpack.unpackb(wire, raw=False, strict_map_key=False))
    assert _same(tree, back), (tree, back)


wire_values = st.recursive(
    st.one_of(st.none(), st.booleans(), st.integers(-5, 2 ** 33), st.floats(allow_nan=False), st.text(max_size=6),
              st.binary(max_size=16)),
    lambda kids: st.one_of(
        st.lists(kids, max_size=3),
        st.dictionaries(st.sampled_from(["__tensor__", "__dict__", "__list__", "shape", "data", "tuple", "x"]), kids,
                        max_size=4)), max_leaves=10)


@given(wire_values)
@settings(**SETTINGS)
def test_arbitrary_messages_either_decode_to_plain_data_or_raise_value_error(msg):
    try:
        out = unpack_state(msg, max_tensor_bytes=1 << 20)
    except ValueError:
        return

    def plain(x):
        if isinstance(x, torch.Tensor):
            return x.numel() * max(x.element_size(), 1) <= 1 << 20
        if isinstance(x, dict):
            return all(plain(k) and plain(v) for k, v in x.items())
        if isinstance(x, (list, tuple)):
            return all(plain(v) for v in x)
        return x is None or isinstance(x, (bool, int, float, str, bytes))

    assert plain(out)


@given(st.lists(st.lists(st.integers(0, 70), min_size=0, max_size=3), min_size=1, max_size=6),
       st.lists(st.sampled_from([0.0, 1.0, 1.0, 2.5]), min_size=1, max_size=6))
@settings(**SETTINGS)
def test_flat_layout_pack_views_and_shards(shapes, shares):
    layout = FlatLayout([torch.Size(s) for s in shapes])
    assert layout.total % CHUNK == 0 and layout.n_chunks == layout.total // CHUNK
    tensors_ = [torch.arange(torch.Size(s).numel(), dtype=torch.float32).reshape(s) + 1000 * i
                for i, s in enumerate(shapes)]
    flat = layout.pack(tensors_)
    for v, t in zip(layout.views(flat), tensors_):
        assert v.shape == t.shape and torch.equal(v, t)
    for a, b in zip(layout.unpack(flat), tensors_):
        assert torch.equal(a, b)
    # every tensor starts on a chunk boundary and chunk -> tensor ids are consistent
    for t, off in enumerate(layout.offsets):
        assert off % CHUNK == 0 and int(layout.chunk_tid[off // CHUNK]) == t
    assert int(layout.tensor_cbeg[-1]) == layout.n_chunks
    if sum(shares) > 0:
        ranges = layout.shard_ranges(shares)
        assert ranges[0][0] == 0 and max(e for _, e in ranges) == layout.n_chunks
        covered = sum(e - b for b, e in ranges)
        assert covered == layout.n_chunks and all(b <= e for b, e in ranges)
        assert all(e == b for (b, e), s in zip(ranges, shares) if s == 0)
        prev = 0
        for b, e in ranges:
            if e > b:
                assert b == prev
                prev = e


@given(st.sampled_from(list(__import__("dedloc_b200.utils.compression", fromlist=["CompressionType"]).CompressionType)),
       st.integers(1, 3), st.integers(0, 2 ** 31 - 1), st.lists(st.tuples(st.integers(0, 10 ** 6), st.integers(0, 255)),
                                                                max_size=6), st.integers(-40, 40))
@settings(**SETTINGS)
def test_encoded_tensor_messages_from_the_network(codec, n_chunks, seed, flips, cut):
    """utils/compression.from_bytes on what a peer may send: a valid message round-trips; a corrupted one either raises
    ValueError or still describes exactly the number of elements it claims (the reducer then checks that number)."""
    from dedloc_b200.utils import compression as comp

    x = torch.randn(n_chunks * comp.CHUNK, generator=torch.Generator().manual_seed(seed))
    raw = comp.to_bytes(comp.encode(x, codec))
    again = comp.from_bytes(raw)
    assert again.numel == x.numel() and torch.equal(comp.decode(again), comp.decode(comp.encode(x, codec)))
    bad = bytearray(raw)
    for pos, val in flips:
        bad[pos % len(bad)] = val
    bad = bytes(bad[:len(bad) + cut]) if cut < 0 else bytes(bad) + bytes(cut)
    try:
        enc = comp.from_bytes(bad)
    except ValueError:
        return
    assert enc.numel > 0 and enc.numel % comp.CHUNK == 0
    assert enc.payload.numel() == enc.numel * comp.payload_bytes_per_elem(enc.codec)
    assert comp.decode(enc).numel() == enc.numel
