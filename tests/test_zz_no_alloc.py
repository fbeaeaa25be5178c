"""Allocation-free Sinkhorn entry point, called straight through the C ABI.

Some users of this library run where device memory cannot be requested on demand: their whole budget belongs to
an allocator they do not control, or they sit in a loop that must not hit cudaMalloc's device-wide synchronisation.
`dedloc_sinkhorn_noalloc` serves them from a built-in per-GPU array instead of caller-provided scratch. This test
calls it twice in a row (the array keeps the previous call's contents, so the second call proves the reset) and
compares each result with the regular entry point; it also checks that an oversized request is refused.
The file is named so that pytest collects it last.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _declare(lib):
    vp, ll, i, f = C.c_void_p, C.c_longlong, C.c_int, C.c_float
    lib.dedloc_sinkhorn_noalloc.argtypes = [vp, i, ll, i, i, i, f, i, f, vp, ll, vp]
    lib.dedloc_sinkhorn_noalloc.restype = i
    lib.dedloc_sinkhorn_noalloc_capacity_floats.restype = i
    lib.dedloc_sinkhorn.argtypes = [vp, i, ll, i, i, i, f, i, f, vp, ll, vp, vp]
    lib.dedloc_sinkhorn.restype = i
    lib.dedloc_sinkhorn_scratch_floats.argtypes = [i, i]
    lib.dedloc_sinkhorn_scratch_floats.restype = ll
    lib.dedloc_last_error.restype = C.c_char_p


def test_noalloc_sinkhorn_matches_regular_entry_point(native):
    lib = native.lib()
    _declare(lib)
    stream = torch.cuda.current_stream().cuda_stream
    assert lib.dedloc_sinkhorn_noalloc_capacity_floats() == 32 + 8 * 4096
    g = torch.Generator().manual_seed(0)
    cases = [(264, 3000, 200, torch.bfloat16, 0.03), (97, 1024, 97, torch.float32, 0.05)]
    for n, K, n_out, dtype, eps in cases:      # two calls in a row, different shapes: stale contents must not leak
        emb = torch.nn.functional.normalize(torch.randn(n, 128, generator=g), dim=1)
        proto = torch.nn.functional.normalize(torch.randn(K, 128, generator=g), dim=1)
        scores = (emb @ proto.t()).to(dtype).cuda().contiguous()
        is_bf16 = int(dtype == torch.bfloat16)
        ref = torch.empty(n_out, K, dtype=torch.float32, device="cuda")
        scratch = torch.empty(int(lib.dedloc_sinkhorn_scratch_floats(K, 3)), dtype=torch.float32, device="cuda")
        rc = lib.dedloc_sinkhorn(scores.data_ptr(), is_bf16, scores.stride(0), n, K, n_out, eps, 3, 1.0,
                                 ref.data_ptr(), ref.stride(0), scratch.data_ptr(), stream)
        assert rc == 0, lib.dedloc_last_error()
        got = torch.full((n_out, K), float("nan"), dtype=torch.float32, device="cuda")
        rc = lib.dedloc_sinkhorn_noalloc(scores.data_ptr(), is_bf16, scores.stride(0), n, K, n_out, eps, 3, 1.0,
                                         got.data_ptr(), got.stride(0), stream)
        assert rc == 0, lib.dedloc_last_error()
        torch.cuda.synchronize()
        assert torch.isfinite(got).all()
        assert torch.allclose(got, ref, rtol=2e-3, atol=1e-7), (got - ref).abs().max()
        assert torch.allclose(got.sum(dim=1), torch.ones(n_out, device="cuda"), atol=1e-4)
    # a request that does not fit the built-in array is refused with a message, nothing is launched
    out = torch.empty(8, 4096, dtype=torch.float32, device="cuda")
    big = torch.zeros(8, 4096, dtype=torch.float32, device="cuda")
    rc = lib.dedloc_sinkhorn_noalloc(big.data_ptr(), 0, big.stride(0), 8, 4096, 8, 0.05, 9, 1.0, out.data_ptr(),
                                     out.stride(0), stream)
    assert rc != 0 and b"scratch floats" in lib.dedloc_last_error()
