"""Baseline B4: SwAV ResNet-50 the way the reference trains it WITHOUT the collaborative optimizer -- torch
DistributedDataParallel over NCCL (swav/ClassyVision/classy_vision/generic/distributed_util.py:219-244 wraps the model
the same way), torchvision ResNet-50 trunk, the SwAV head / Sinkhorn / swapped-prediction loss written in plain PyTorch
(restating swav/vissl/vissl/models/heads/swav_prototypes_head.py:64-110 and swav/vissl/vissl/losses/swav_loss.py:177-302)
and SGD-momentum under LARC (swav_1node_resnet_submit.yaml:96-107). None of this repository's kernels, models or
optimizers are on this path: it is the bar `scripts/bench_swav.py` compares against.

    python scripts/bench_swav.py --impl ddp [--batch 64]                 # one GPU
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_swav.py --impl ddp --gpus N
"""
from __future__ import annotations

import os
from typing import List

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


class TorchSwAVHead(nn.Module):
    def __init__(self, dims=(2048, 2048, 128), num_prototypes: int = 3000):
        super().__init__()
        self.projection = nn.Sequential(nn.Linear(dims[0], dims[1]), nn.BatchNorm1d(dims[1]), nn.ReLU(inplace=True),
                                        nn.Linear(dims[1], dims[2]))
        self.prototypes = nn.Linear(dims[2], num_prototypes, bias=False)

    def forward(self, feats: torch.Tensor):
        z = F.normalize(self.projection(feats), dim=1, p=2)
        return z, self.prototypes(z)


class TorchSwAV(nn.Module):
    """Trunk run once per resolution on the concatenated crops of that resolution (multi-crop forward)."""

    def __init__(self):
        super().__init__()
        import torchvision

        trunk = torchvision.models.resnet50(weights=None)
        trunk.fc = nn.Identity()
        self.trunk = trunk
        self.head = TorchSwAVHead()

    def forward(self, crops: List[torch.Tensor]):
        feats, start = [], 0
        sizes = [c.shape[-1] for c in crops]
        while start < len(crops):
            end = start
            while end < len(crops) and sizes[end] == sizes[start]:
                end += 1
            feats.append(self.trunk(torch.cat(crops[start:end])))
            start = end
        return self.head(torch.cat(feats))


@torch.no_grad()
def sinkhorn(scores: torch.Tensor, epsilon: float = 0.03, iters: int = 3, world: int = 1) -> torch.Tensor:
    """swav_loss.py:196-243 (distributed_sinkhornknopp): scores [B, K] -> assignments [B, K]."""
    q = torch.exp(scores.float() / epsilon).t()                # [K, B]
    total = q.sum()
    if world > 1:
        dist.all_reduce(total)
    q /= total
    k, b = q.shape
    b_total = b * world
    for _ in range(iters):
        rows = q.sum(dim=1, keepdim=True)
        if world > 1:
            dist.all_reduce(rows)
        q /= rows * k
        q /= q.sum(dim=0, keepdim=True) * b_total
    return (q * b_total).t()


def swav_loss(scores: torch.Tensor, batch: int, num_crops: int = 8, assign=(0, 1), temperature: float = 0.1,
              world: int = 1) -> torch.Tensor:
    loss = 0.0
    for i in assign:
        q = sinkhorn(scores[i * batch:(i + 1) * batch].detach(), world=world)
        sub = 0.0
        for p in range(num_crops):
            if p == i:
                continue
            logp = F.log_softmax(scores[p * batch:(p + 1) * batch].float() / temperature, dim=1)
            sub = sub - torch.mean(torch.sum(q * logp, dim=1))
        loss = loss + sub / (num_crops - 1)
    return loss / len(assign)


class LARC:
    """apex.parallel.LARC(clip=False) restated: per-tensor trust ratio folded into the gradient before SGD."""

    def __init__(self, optimizer: torch.optim.Optimizer, trust_coefficient: float = 0.001, eps: float = 1e-8):
        self.optim, self.trust, self.eps = optimizer, trust_coefficient, eps

    @torch.no_grad()
    def step(self):
        for group in self.optim.param_groups:
            wd = group.get("weight_decay", 0.0)
            params = [p for p in group["params"] if p.grad is not None]
            if not params:
                continue
            p_norms = torch._foreach_norm(params)
            g_norms = torch._foreach_norm([p.grad for p in params])
            for p, pn, gn in zip(params, p_norms, g_norms):
                ratio = torch.where((pn > 0) & (gn > 0), self.trust * pn / (gn + wd * pn + self.eps), torch.ones_like(pn))
                p.grad.add_(p, alpha=wd).mul_(ratio)
            group["_wd"], group["weight_decay"] = wd, 0.0
        self.optim.step()
        for group in self.optim.param_groups:
            if "_wd" in group:
                group["weight_decay"] = group.pop("_wd")


def run(args) -> dict:
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    on_gpu = torch.cuda.is_available() and args.device != "cpu"
    device = torch.device("cuda", local) if on_gpu else torch.device("cpu")
    if on_gpu:
        torch.cuda.set_device(device)
        torch.backends.cudnn.benchmark = True
    if world > 1:
        dist.init_process_group("nccl" if on_gpu else "gloo", **({"device_id": device} if on_gpu else {}))
    torch.manual_seed(0)
    model = TorchSwAV().to(device)
    if on_gpu:
        model = model.to(memory_format=torch.channels_last)
    if world > 1:
        model = nn.SyncBatchNorm.convert_sync_batchnorm(model) if args.sync_bn else model
        model = nn.parallel.DistributedDataParallel(model, device_ids=[local] if on_gpu else None,
                                                    gradient_as_bucket_view=True)
    base = torch.optim.SGD(model.parameters(), lr=0.6, momentum=0.9, weight_decay=1e-6, nesterov=False)
    opt = LARC(base)
    sizes = [(2, args.big), (6, args.small)]
    batches = [[torch.randn(args.batch, 3, s, s, device=device) for n, s in sizes for _ in range(n)] for _ in range(2)]
    if on_gpu:
        batches = [[c.contiguous(memory_format=torch.channels_last) for c in b] for b in batches]

    def step(i):
        with torch.autocast(device.type, dtype=torch.bfloat16, enabled=on_gpu):
            _, scores = model(batches[i % 2])
        loss = swav_loss(scores, args.batch, world=world)
        base.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        with torch.no_grad():      # prototypes stay on the unit sphere (swav_loss.py:137-150)
            w = (model.module if world > 1 else model).head.prototypes.weight
            w.copy_(F.normalize(w, dim=1, p=2))
        return loss

    for i in range(args.warmup):
        loss = step(i)
    if world > 1:
        dist.barrier()
    if on_gpu:
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    import time

    t0 = time.perf_counter()
    for i in range(args.steps):
        loss = step(i)
    if on_gpu:
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
    else:
        ms = (time.perf_counter() - t0) * 1e3 / args.steps
    if world > 1:
        t = torch.tensor([ms], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    out = {"metric": "SwAV RN50 images/sec", "impl": "ddp", "value": round(world * args.batch / ms * 1e3, 1),
           "unit": "images/s", "n_gpus": world, "ms_per_step": round(ms, 2), "batch_per_replica": args.batch,
           "crops": f"2x{args.big} + 6x{args.small}", "optimizer": "torch SGD-momentum + LARC (restated), DDP/NCCL",
           "dtype": "bf16 autocast, channels_last", "loss": float(loss.detach()), "data": "synthetic"}
    if world > 1:
        dist.destroy_process_group()
    return out if rank == 0 else {}
