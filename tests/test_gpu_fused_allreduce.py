"""Fused all-reduce + optimizer kernel on ONE GPU with emulated ranks.

P buffer sets live on the same device; phase A is launched for every rank, then phase B for every rank, with
the cross-GPU flag waits disabled (ranks that wait on each other must not be separate launches on one GPU --
B200_PROFILING.md). The device code is the production code; only the barriers differ. Oracles are the CPU
codecs + optim/lamb.py + optim/lars.py.
"""
import math

import pytest
import torch

from dedloc_b200.utils import compression as comp
from dedloc_b200.utils.compression import CompressionType

pytestmark = pytest.mark.gpu

SHAPES = [(300, 128), (128,), (1024, 64), (64,), (3000, 128), (7,)]
WD = [0.01, 0.0, 0.01, 0.0, 0.01, 0.0]


def _setup(P, seed=0, with_state=True):
    from dedloc_b200.parallel.flat import FlatLayout
    from dedloc_b200.parallel.fused_allreduce import PeerBuffers

    layout = FlatLayout([torch.Size(s) for s in SHAPES])
    bufs = [PeerBuffers(layout, with_state=with_state) for _ in range(P)]
    g = torch.Generator().manual_seed(seed)
    params = [torch.randn(s, generator=g) * 0.05 for s in SHAPES]
    grads = [[torch.randn(s, generator=g) * 0.01 for s in SHAPES] for _ in range(P)]
    return layout, bufs, params, grads


def _run(bufs, layout, codec, hyper, weights, shards, tag, **kw):
    from dedloc_b200.parallel.fused_allreduce import GroupSpec, fused_allreduce

    P = len(bufs)
    bases = [b.symm.ptr for b in bufs]
    phases = (1, 2) if hyper.mode != 0 else (1,)
    for ph in phases:
        for r in range(P):
            group = GroupSpec(bases=bases, me=r, weights=weights, shards=shards, slots=list(range(P)), tag=tag)
            fused_allreduce(bufs[r], group, codec, hyper, phases=ph, xgpu_barriers=False, **kw)
    torch.cuda.synchronize()


@pytest.mark.parametrize("codec", list(CompressionType))
def test_weighted_mean_all_codecs(native, codec):
    from dedloc_b200.ops import flat_ops
    from dedloc_b200.parallel.fused_allreduce import OptHyper

    P = 4
    layout, bufs, _, grads = _setup(P, seed=1, with_state=False)
    weights = [3.0, 1.0, 0.0, 2.0]  # member 2 is an auxiliary peer (weight 0) but owns a shard
    shards = layout.shard_ranges([1, 1, 1, 0])  # member 3 is a client-mode peer: owns nothing
    decoded = []
    for r in range(P):
        bufs[r].acc.copy_(layout.pack(grads[r]).cuda())
        flat_ops.encode(codec, bufs[r].acc, 0.5, bufs[r].send, bufs[r].aux)
        cpu = comp.Encoded(codec, bufs[r].send[: layout.total * comp.payload_bytes_per_elem(codec)].cpu(),
                           bufs[r].aux[: comp.aux_floats(codec, layout.total)].cpu()
                           if comp.aux_floats(codec, layout.total) else None, layout.total)
        decoded.append(comp.decode(cpu))
    wsum = sum(weights)
    ref = sum(w / wsum * d for w, d in zip(weights, decoded))
    _run(bufs, layout, codec, OptHyper(mode=0), weights, shards, tag=11)
    for r in range(P):
        assert torch.allclose(bufs[r].p32.cpu(), ref, rtol=1e-5, atol=1e-7), (codec, r)
        assert torch.allclose(bufs[r].p16.float().cpu(), ref, rtol=1e-2, atol=1e-5)
        assert bufs[r].read_status()["code"] == 0


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


def test_fused_lamb_matches_reference(native):
    from dedloc_b200 import lib
    from dedloc_b200.ops import flat_ops
    from dedloc_b200.optim import lamb_bias_correction
    from dedloc_b200.parallel.fused_allreduce import OptHyper

    P = 3
    layout, bufs, params, _ = _setup(P, seed=2)
    weights = [1.0, 2.0, 1.0]
    shards = layout.shard_ranges([1, 1, 1])
    flat_p = layout.pack(params)
    for b in bufs:
        b.p32.copy_(flat_p.cuda())
        b.t_wd.copy_(torch.tensor(WD))
    g = torch.Generator().manual_seed(5)
    all_avg = []
    lr = 0.00176
    for step in range(1, 4):
        grads = [[torch.randn(s, generator=g) * 0.01 for s in SHAPES] for _ in range(P)]
        for r in range(P):
            bufs[r].acc.copy_(layout.pack(grads[r]).cuda())
            flat_ops.encode(CompressionType.NONE, bufs[r].acc, 1.0, bufs[r].send, bufs[r].aux)
        wsum = sum(weights)
        all_avg.append([sum(w / wsum * grads[r][i] for r, w in enumerate(weights)) for i in range(len(SHAPES))])
        hyper = OptHyper(mode=lib.AR_LAMB, lr=lr, beta1=0.9, beta2=0.999, eps=1e-6,
                         bias_correction=lamb_bias_correction(step, 0.9, 0.999), clamp_value=10000.0)
        _run(bufs, layout, CompressionType.NONE, hyper, weights, shards, tag=20 + step)
    ref_p, ref_opt = _ref_lamb(params, all_avg, 3, lr, WD)
    for r in range(P):
        got = layout.views(bufs[r].p32.cpu())
        for a, b_ in zip(got, ref_p):
            assert torch.allclose(a, b_.detach(), rtol=2e-5, atol=1e-7)
        got16 = layout.views(bufs[r].p16.float().cpu())
        for a, b_ in zip(got16, ref_p):
            assert torch.allclose(a, b_.detach(), rtol=1e-2, atol=1e-4)
        # optimizer moments are replicated on every member
        m = layout.views(bufs[r].m.cpu())
        v = layout.views(bufs[r].v.cpu())
        for i, p_ in enumerate(ref_p):
            st = ref_opt.state[p_]
            assert torch.allclose(m[i], st["exp_avg"], rtol=1e-5, atol=1e-9)
            assert torch.allclose(v[i], st["exp_avg_sq"], rtol=1e-4, atol=1e-12)
    # padding stays exactly zero
    flat = bufs[0].p32.cpu()
    mask = torch.ones(layout.total, dtype=torch.bool)
    for o, n in zip(layout.offsets, layout.numels):
        mask[o:o + n] = False
    assert flat[mask].abs().max().item() == 0.0


def test_fused_lars_matches_reference_and_rownorm_and_freeze(native):
    from dedloc_b200 import lib
    from dedloc_b200.ops import flat_ops
    from dedloc_b200.optim import build_lars_sgd
    from dedloc_b200.parallel.fused_allreduce import OptHyper

    P = 2
    layout, bufs, params, _ = _setup(P, seed=3)
    weights = [1.0, 1.0]
    shards = layout.shard_ranges([1, 1])
    wd = [1e-6, 0.0, 1e-6, 0.0, 1e-6, 0.0]
    flags = [0, 0, 0, 0, lib.TF_ROWNORM128, lib.TF_FROZEN]  # tensor 4 plays the prototypes, 5 is frozen
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
            flat_ops.encode(CompressionType.NONE, bufs[r].acc, 1.0, bufs[r].send, bufs[r].aux)
        hyper = OptHyper(mode=lib.AR_LARS, lr=0.6, momentum=0.9, trust=0.001, eps=1e-8)
        _run(bufs, layout, CompressionType.NONE, hyper, weights, shards, tag=40 + step)
        for i, p in enumerate(ps):
            p.grad = None if i == 5 else 0.5 * (grads[0][i] + grads[1][i])
        opt.step()
        with torch.no_grad():
            ps[4].div_(ps[4].norm(dim=1, keepdim=True).clamp_min(1e-12))
    for r in range(P):
        got = layout.views(bufs[r].p32.cpu())
        for i, (a, b_) in enumerate(zip(got, ps)):
            assert torch.allclose(a, b_.detach(), rtol=3e-5, atol=1e-7), i
    assert torch.equal(layout.views(bufs[0].p32.cpu())[5], params[5])


def test_nonfinite_gradient_skips_update(native):
    from dedloc_b200 import lib
    from dedloc_b200.ops import flat_ops
    from dedloc_b200.parallel.fused_allreduce import OptHyper

    P = 2
    layout, bufs, params, grads = _setup(P, seed=4)
    flat_p = layout.pack(params).cuda()
    for b in bufs:
        b.p32.copy_(flat_p)
    grads[1][2][5, 5] = float("nan")
    for r in range(P):
        bufs[r].acc.copy_(layout.pack(grads[r]).cuda())
        flat_ops.encode(CompressionType.FLOAT16, bufs[r].acc, 1.0, bufs[r].send, bufs[r].aux)
    hyper = OptHyper(mode=lib.AR_LAMB, lr=1e-3, bias_correction=1.0)
    _run(bufs, layout, CompressionType.FLOAT16, hyper, [1.0, 1.0], layout.shard_ranges([1, 1]), tag=60)
    for b in bufs:
        assert b.read_status()["name"] == "skipped_nonfinite"
        assert torch.equal(b.p32, flat_p)
        assert b.m.abs().max().item() == 0.0
