"""Push-based, stage-pipelined all-reduce + optimizer kernel (csrc/fused_allreduce_v2.cu) on ONE GPU.

All members of a group run as one cooperative launch over P buffer sets on the same device (ranks that wait on each
other must not be separate launches on one GPU -- B200_PROFILING.md), so the flags, the staged payloads, the
{value, tag} norm words and the bounded waits are the production protocol. Oracles are the CPU codecs +
optim/lamb.py + optim/lars.py, as for the first kernel (tests/test_gpu_fused_allreduce.py).
"""
import pytest
import torch

from dedloc_b200.utils import compression as comp
from dedloc_b200.utils.compression import CompressionType

pytestmark = pytest.mark.gpu

SHAPES = [(300, 128), (128,), (1024, 64), (64,), (3000, 128), (7,), (513, 1024)]
WD = [0.01, 0.0, 0.01, 0.0, 0.01, 0.0, 0.01]


def _setup(P, seed=0, with_state=True, shapes=SHAPES):
    from dedloc_b200.parallel.flat import FlatLayout
    from dedloc_b200.parallel.fused_allreduce import PeerBuffers

    layout = FlatLayout([torch.Size(s) for s in shapes])
    bufs = [PeerBuffers(layout, with_state=with_state) for _ in range(P)]
    g = torch.Generator().manual_seed(seed)
    params = [torch.randn(s, generator=g) * 0.05 for s in shapes]
    grads = [[torch.randn(s, generator=g) * 0.01 for s in shapes] for _ in range(P)]
    return layout, bufs, params, grads


def _owners(mask):
    return [(0, 1) if o else (0, 0) for o in mask]


@pytest.mark.parametrize("codec", list(CompressionType))
@pytest.mark.parametrize("fly", [False, True])
def test_weighted_mean_all_codecs(native, codec, fly, monkeypatch):
    if codec in (CompressionType.FLOAT16, CompressionType.UNIFORM_8BIT):
        monkeypatch.setenv("DEDLOC_AR2_SCHEDULE", "concurrent")
    monkeypatch.setenv("DEDLOC_AR2_STAGE_MB", "0.25")        # several stages even for this small vector
    from dedloc_b200.ops import flat_ops
    from dedloc_b200.parallel.fused_allreduce import OptHyper, fused_allreduce_v2_emulated

    if fly and codec not in (CompressionType.NONE, CompressionType.FLOAT16, CompressionType.BFLOAT16):
        pytest.skip("codec needs the two-pass encoder")
    P = 4
    layout, bufs, _, grads = _setup(P, seed=1, with_state=False)
    weights = [3.0, 1.0, 0.0, 2.0]          # member 2 is an auxiliary peer (weight 0) but owns a share
    shards = _owners([1, 1, 1, 0])          # member 3 is a client-mode peer: owns nothing
    decoded = []
    for r in range(P):
        bufs[r].acc.copy_(layout.pack(grads[r]).cuda())
        flat_ops.encode(codec, bufs[r].acc, 0.5, bufs[r].send, bufs[r].aux)
        cpu = comp.Encoded(codec, bufs[r].send[: layout.total * comp.payload_bytes_per_elem(codec)].cpu(),
                           bufs[r].aux[: comp.aux_floats(codec, layout.total)].cpu()
                           if comp.aux_floats(codec, layout.total) else None, layout.total)
        decoded.append(comp.decode(cpu))
        if fly:
            bufs[r].send.zero_()            # the kernel must encode from the accumulators itself
    wsum = sum(weights)
    ref = sum(w / wsum * d for w, d in zip(weights, decoded))
    for rnd in range(2):                    # a second round reuses flags / staging with a new tag
        for b in bufs:
            b.p32.zero_()
            b.p16.zero_()
        fused_allreduce_v2_emulated(bufs, weights, shards, codec, OptHyper(mode=0), tag=101 + rnd, src_f32=fly,
                                    scale=0.5)
        torch.cuda.synchronize()
        for r in range(P):
            assert bufs[r].read_status()["code"] == 0, bufs[r].read_status()
            assert torch.allclose(bufs[r].p32.cpu(), ref, rtol=1e-5, atol=1e-7), (codec, r)
            assert torch.allclose(bufs[r].p16.float().cpu(), ref, rtol=1e-2, atol=1e-5)


def _ref_lamb(params, grads_avg, steps, lr, wd_list, eps=1e-6, clamp=10000.0):
    from dedloc_b200.optim import Lamb

    ps = [p.clone().requires_grad_(True) for p in params]
    groups = [{"params": [p], "weight_decay": wd} for p, wd in zip(ps, wd_list)]
    opt = Lamb(groups, lr=lr, betas=(0.9, 0.999), eps=eps, weight_decay=0.01, clamp_value=clamp, debias=True)
    for g in grads_avg[:steps]:
        for p, gi in zip(ps, g):
            p.grad = gi.clone()
        opt.step()
    return ps, opt


@pytest.mark.parametrize("schedule", ["sequential", "concurrent"])
@pytest.mark.parametrize("replicate", [False, True])
def test_fused_lamb_matches_reference(native, replicate, schedule, monkeypatch):
    monkeypatch.setenv("DEDLOC_AR2_SCHEDULE", schedule)     # phases one after the other / four roles side by side
    from dedloc_b200 import lib
    from dedloc_b200.optim import lamb_bias_correction
    from dedloc_b200.parallel.fused_allreduce import OptHyper, fused_allreduce_v2_emulated

    monkeypatch.setenv("DEDLOC_AR2_STAGE_MB", "0.25")
    P = 3
    layout, bufs, params, _ = _setup(P, seed=2)
    weights = [1.0, 2.0, 1.0]
    shards = _owners([1, 1, 1])
    flat_p = layout.pack(params)
    for b in bufs:
        b.p32.copy_(flat_p.cuda())
        b.p16.copy_(flat_p.cuda())
        b.t_wd.copy_(torch.tensor(WD))
    g = torch.Generator().manual_seed(5)
    all_avg = []
    lr = 0.00176
    for step in range(1, 4):
        grads = [[torch.randn(s, generator=g) * 0.01 for s in SHAPES] for _ in range(P)]
        for r in range(P):
            bufs[r].acc.copy_(layout.pack(grads[r]).cuda())
        wsum = sum(weights)
        all_avg.append([sum(w / wsum * grads[r][i] for r, w in enumerate(weights)) for i in range(len(SHAPES))])
        hyper = OptHyper(mode=lib.AR_LAMB, lr=lr, beta1=0.9, beta2=0.999, eps=1e-6,
                         bias_correction=lamb_bias_correction(step, 0.9, 0.999), clamp_value=10000.0,
                         replicate_state=replicate)
        fused_allreduce_v2_emulated(bufs, weights, shards, CompressionType.NONE, hyper, tag=200 + step)
        torch.cuda.synchronize()
        for b in bufs:
            assert b.read_status()["code"] == 0, b.read_status()
    ref_p, ref_opt = _ref_lamb(params, all_avg, 3, lr, WD)
    for r in range(P):
        assert torch.equal(bufs[r].p32, bufs[0].p32), "fp32 masters must be bit-identical on every member"
        assert torch.equal(bufs[r].p16, bufs[0].p16)
        got = layout.views(bufs[r].p32.cpu())
        for a, b_ in zip(got, ref_p):
            # the update direction crosses the wire as fp16: 2^-11 relative error on a step of ~1e-3 |p|
            assert torch.allclose(a, b_.detach(), rtol=2e-5, atol=2e-7)
        got16 = layout.views(bufs[r].p16.float().cpu())
        for a, b_ in zip(got16, ref_p):
            assert torch.allclose(a, b_.detach(), rtol=1e-2, atol=1e-4)
    # moments: with the owner (interleaved chunks), or on every member when replicate_state is set
    chunk = 2048
    for r in range(P):
        m, v = bufs[r].m.cpu(), bufs[r].v.cpu()
        mine = (torch.arange(layout.n_chunks) % P) == r
        sel = mine.repeat_interleave(chunk) if not replicate else torch.ones(layout.total, dtype=torch.bool)
        ref_m = layout.pack([ref_opt.state[p_]["exp_avg"] for p_ in ref_p])
        ref_v = layout.pack([ref_opt.state[p_]["exp_avg_sq"] for p_ in ref_p])
        assert torch.allclose(m[sel], ref_m[sel], rtol=1e-5, atol=1e-9)
        assert torch.allclose(v[sel], ref_v[sel], rtol=1e-4, atol=1e-12)
        if not replicate:
            assert m[~sel].abs().max().item() == 0.0      # nobody else's moments were touched
    flat = bufs[0].p32.cpu()
    mask = torch.ones(layout.total, dtype=torch.bool)
    for o, n in zip(layout.offsets, layout.numels):
        mask[o:o + n] = False
    assert flat[mask].abs().max().item() == 0.0           # padding stays exactly zero


def test_fused_lars_matches_reference_and_rownorm_and_freeze(native):
    from dedloc_b200 import lib
    from dedloc_b200.optim import build_lars_sgd
    from dedloc_b200.parallel.fused_allreduce import OptHyper, fused_allreduce_v2_emulated

    P = 2
    layout, bufs, params, _ = _setup(P, seed=3)
    weights = [1.0, 1.0]
    shards = _owners([1, 1])
    wd = [1e-6, 0.0, 1e-6, 0.0, 1e-6, 0.0, 1e-6]
    flags = [0, 0, 0, 0, lib.TF_ROWNORM128, lib.TF_FROZEN, 0]  # tensor 4 plays the prototypes, 5 is frozen
    for b in bufs:
        b.p32.copy_(layout.pack(params).cuda())
        b.t_wd.copy_(torch.tensor(wd))
        b.t_flags.copy_(torch.tensor(flags, dtype=torch.int32))
    ps = [p.clone().requires_grad_(True) for p in params]
    opt = build_lars_sgd([{"params": [p], "weight_decay": w} for p, w in zip(ps, wd)], lr=0.6, momentum=0.9,
                         trust_coefficient=0.001, clip=False, eps=1e-8)
    g = torch.Generator().manual_seed(9)
    for step in range(3):
        grads = [[torch.randn(s, generator=g) * 0.01 for s in SHAPES] for _ in range(P)]
        for r in range(P):
            bufs[r].acc.copy_(layout.pack(grads[r]).cuda())
        hyper = OptHyper(mode=lib.AR_LARS, lr=0.6, momentum=0.9, trust=0.001, eps=1e-8, replicate_state=False)
        fused_allreduce_v2_emulated(bufs, weights, shards, CompressionType.NONE, hyper, tag=300 + step)
        torch.cuda.synchronize()
        for i, p in enumerate(ps):
            p.grad = None if i == 5 else 0.5 * (grads[0][i] + grads[1][i])
        opt.step()
        with torch.no_grad():
            ps[4].div_(ps[4].norm(dim=1, keepdim=True).clamp_min(1e-12))
    assert torch.equal(bufs[0].p32, bufs[1].p32)
    for r in range(P):
        assert bufs[r].read_status()["code"] == 0
        got = layout.views(bufs[r].p32.cpu())
        for i, (a, b_) in enumerate(zip(got, ps)):
            # the momentum crosses the wire as bf16 (2^-8 relative) and is scaled by lr = 0.6
            assert torch.allclose(a, b_.detach(), rtol=2e-3, atol=2e-6), i
    assert torch.equal(layout.views(bufs[0].p32.cpu())[5], params[5])


def test_nonfinite_tensor_is_skipped_by_everyone(native):
    from dedloc_b200 import lib
    from dedloc_b200.parallel.fused_allreduce import OptHyper, fused_allreduce_v2_emulated

    P = 2
    layout, bufs, params, grads = _setup(P, seed=4)
    flat_p = layout.pack(params).cuda()
    for b in bufs:
        b.p32.copy_(flat_p)
    grads[1][2][5, 5] = float("nan")
    for r in range(P):
        bufs[r].acc.copy_(layout.pack(grads[r]).cuda())
    hyper = OptHyper(mode=lib.AR_LAMB, lr=1e-3, bias_correction=1.0, replicate_state=False)
    fused_allreduce_v2_emulated(bufs, [1.0, 1.0], _owners([1, 1]), CompressionType.FLOAT16, hyper, tag=400)
    torch.cuda.synchronize()
    assert torch.equal(bufs[0].p32, bufs[1].p32)
    for b in bufs:
        st = b.read_status()
        assert st["name"] == "skipped_nonfinite" and st["nonfinite_chunks"] == 1, st
        got, ref = layout.views(b.p32.cpu()), layout.views(flat_p.cpu())
        assert torch.equal(got[2], ref[2])                       # the poisoned tensor did not move ...
        assert layout.views(b.m.cpu())[2].abs().max().item() == 0.0
        assert torch.isfinite(b.p32).all()
        for i in (0, 4, 6):
            assert not torch.equal(got[i], ref[i])               # ... the others stepped


def test_absent_member_times_out_cleanly(native):
    """A member that never shows up (here: it runs with another round tag) makes every wait expire: all members
    return `abort_before_reduce` with their state untouched, nobody hangs."""
    import ctypes as C
    import time

    from dedloc_b200 import lib
    from dedloc_b200.parallel import fused_allreduce as fa

    P = 3
    layout, bufs, params, grads = _setup(P, seed=6)
    flat_p = layout.pack(params).cuda()
    for r, b in enumerate(bufs):
        b.p32.copy_(flat_p)
        b.acc.copy_(layout.pack(grads[r]).cuda())
    hyper = fa.OptHyper(mode=lib.AR_LAMB, lr=1e-3, bias_correction=1.0, replicate_state=False)
    cap = fa._max_ctas(hyper.mode, CompressionType.FLOAT16)
    ctas = cap // P
    arr = (lib.AR2Params * P)()
    bases = [b.symm.ptr for b in bufs]
    for r in range(P):
        g = fa.GroupSpec(bases=bases, me=r, weights=[1.0] * P, shards=_owners([1] * P), slots=list(range(P)),
                         tag=500 if r != 2 else 501, timeout_s=0.25)
        fa._fill_v2(arr[r], bufs[r], g, CompressionType.FLOAT16, hyper, ctas, True, 1.0, True, True)
    t0 = time.time()
    lib.check(lib.lib().dedloc_fused_allreduce_v2(arr, P, bufs[0].params_dev.data_ptr(),
                                                  torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert time.time() - t0 < 5.0
    for b in bufs:
        assert b.read_status()["name"] == "abort_before_reduce", b.read_status()
        assert torch.equal(b.p32, flat_p)
        assert b.m.abs().max().item() == 0.0


@pytest.mark.parametrize("schedule", ["sequential", "concurrent"])
def test_flag_protocol_soak(native, schedule, monkeypatch):
    """Thousands of back-to-back rounds with a fresh tag each and members that start up to ~50 us apart: a stale flag, a
    norm word of the previous round or a counter that did not reset would let a member run ahead on old data. The
    gradient scale changes every round, so every round has a different right answer; errors are counted on the
    device and read once."""
    from dedloc_b200.parallel.fused_allreduce import OptHyper, fused_allreduce_v2_emulated

    monkeypatch.setenv("DEDLOC_AR2_SCHEDULE", schedule)
    monkeypatch.setenv("DEDLOC_AR2_STAGE_MB", "0.25")
    P, rounds = 4, 1500
    shapes = [(96, 128), (64,), (640, 64), (3,)]
    layout, bufs, _, grads = _setup(P, seed=11, with_state=False, shapes=shapes)
    weights = [1.0, 2.0, 3.0, 2.0]
    for r in range(P):
        bufs[r].acc.copy_(layout.pack(grads[r]).cuda())
    wsum = sum(weights)
    base = sum(w / wsum * layout.pack(grads[r]) for r, w in enumerate(weights)).cuda()
    bad = torch.zeros((), dtype=torch.int64, device="cuda")
    for i in range(rounds):
        scale = 1.0 + (i % 7) * 0.25
        fused_allreduce_v2_emulated(bufs, weights, _owners([1] * P), CompressionType.NONE, OptHyper(mode=0),
                                    tag=10_000 + i, scale=scale, debug_flags=(3 << 8))
        ref = base * scale
        for b in (bufs[i % P],):                       # one member per round keeps the host side cheap
            bad += ((b.p32 - ref).abs().max() > 1e-6).long()
    torch.cuda.synchronize()
    assert int(bad) == 0
    for b in bufs:
        assert b.read_status()["code"] == 0
