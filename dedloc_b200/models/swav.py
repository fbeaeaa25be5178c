"""SwAV: ResNet trunk + prototypes head + loss with a queue keyed on the COLLABORATION-wide step.

Scope = exactly what the collaborative SwAV recipe executes (SURVEY 2.4 scope decision):
  * trunk `resnet` (torchvision ResNet-50 without the fc layer; VISSL trunks/resnext.py:48-172) with the
    multi-resolution forward of base_ssl_model.py:76-105 (crops of equal resolution share one trunk pass);
  * head `swav_head` (heads/swav_prototypes_head.py:10-112): MLP 2048 -> 2048 -> BN -> ReLU -> 128, L2-normalise,
    bias-free prototype layer(s) 128 -> 3000; parameter names follow VISSL (`projection_head.*`, `prototypes0.weight`);
  * `SwAVLoss` (losses/swav_loss.py:24-380): Sinkhorn-Knopp assignments for the crops in `crops_for_assign`,
    cross-entropy against log_softmax(scores / T) of every other crop, feature queue that switches on when the
    collaboration's optimizer step reaches `queue.start_iter` (DeDLOC edit W3).
Sinkhorn runs in one persistent CUDA kernel (ops/sinkhorn.py); the queue is a ring buffer (no roll copies).
"""
from __future__ import annotations

import logging
import math
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops.sinkhorn import sinkhorn, sinkhorn_reference

logger = logging.getLogger(__name__)


class ResNetTrunk(nn.Module):
    """torchvision ResNe(X)t without the classifier; returns the pooled, flattened `res5avg` feature."""

    def __init__(self, depth: int = 50, width_multiplier: int = 1, zero_init_residual: bool = False):
        super().__init__()
        import torchvision.models as tvm

        ctor = {18: tvm.resnet18, 34: tvm.resnet34, 50: tvm.resnet50, 101: tvm.resnet101, 152: tvm.resnet152}[depth]
        net = ctor(weights=None, zero_init_residual=zero_init_residual, width_per_group=64 * width_multiplier)
        self.conv1, self.bn1, self.relu, self.maxpool = net.conv1, net.bn1, net.relu, net.maxpool
        self.layer1, self.layer2, self.layer3, self.layer4 = net.layer1, net.layer2, net.layer3, net.layer4
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.out_dim = net.fc.in_features

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.maxpool(self.relu(self.bn1(self.conv1(x))))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        return torch.flatten(self.avgpool(x), 1)


class SwAVPrototypesHead(nn.Module):
    def __init__(self, dims: Sequence[int] = (2048, 2048, 128), use_bn: bool = True,
                 num_clusters: Sequence[int] = (3000,), use_bias: bool = True, return_embeddings: bool = True,
                 skip_last_bn: bool = True, normalize_feats: bool = True):
        super().__init__()
        layers: List[nn.Module] = []
        last = dims[0]
        for i, dim in enumerate(dims[1:]):
            layers.append(nn.Linear(last, dim, bias=use_bias))
            if i == len(dims) - 2:
                break
            if use_bn:
                layers.append(nn.BatchNorm1d(dim))
            layers.append(nn.ReLU(inplace=True))
            last = dim
        self.projection_head = nn.Sequential(*layers)
        self.nmb_heads = len(num_clusters)
        self.return_embeddings, self.normalize_feats = return_embeddings, normalize_feats
        for i, k in enumerate(num_clusters):
            self.add_module(f"prototypes{i}", nn.Linear(dims[-1], k, bias=False))

    def prototypes(self, i: int) -> nn.Linear:
        return getattr(self, f"prototypes{i}")

    def _native_ok(self, feats: torch.Tensor) -> bool:
        """The head runs on this repo's kernels when it sits on a GPU, its parameters carry the fused optimizer's fp32
        `main_grad` buffers, and the shapes suit the tcgen05 GEMM (token count and widths multiples of 64; the
        prototype count only needs a multiple of 8: TMA zero-fills the tail of the 64-wide K block)."""
        if not (feats.is_cuda and feats.dim() == 2 and feats.shape[0] % 64 == 0 and self.normalize_feats):
            return False
        lins = [m for m in self.projection_head if isinstance(m, nn.Linear)]
        protos = [self.prototypes(i) for i in range(self.nmb_heads)]
        if len(lins) != 2 or lins[-1].out_features != 128:
            return False
        from .. import lib as native

        if not native.available():
            return False
        for lin in lins + protos:
            if getattr(lin.weight, "main_grad", None) is None or lin.in_features % 64 or lin.out_features % 8:
                return False
            if lin.bias is not None and getattr(lin.bias, "main_grad", None) is None:
                return False
        return all(l.out_features % 64 == 0 for l in lins)

    def _forward_native(self, feats: torch.Tensor) -> List[torch.Tensor]:
        """C13 on the repo's kernels: Linear 2048 -> 2048 (tcgen05 GEMM, bias in the epilogue) -> BatchNorm1d + ReLU
        (batch statistics: PyTorch) -> Linear 2048 -> 128 (GEMM) -> L2-normalise (csrc/swav_ops.cu) -> prototypes
        128 -> 3000 (GEMM). Weight and bias gradients accumulate into the fp32 main_grad buffers (wgrad GEMM)."""
        from ..ops.gemm import linear as native_linear
        from ..ops.swav_ops import l2_normalize128

        def lin16(x, lin):
            w = lin.weight.detach().to(torch.bfloat16)
            b = lin.bias.detach().to(torch.bfloat16) if lin.bias is not None else None
            return native_linear(x, w, b, wgrad_out=lin.weight.main_grad,
                                 bgrad_out=lin.bias.main_grad if lin.bias is not None else None)

        x = feats.to(torch.bfloat16)
        if not x.requires_grad:
            x.requires_grad_(feats.requires_grad)
        for m in self.projection_head:
            if isinstance(m, nn.Linear):
                x = lin16(x, m)
            else:
                x = m(x)                      # BatchNorm1d (bf16 in, fp32 statistics) / ReLU
                if x.dtype != torch.bfloat16:
                    x = x.to(torch.bfloat16)
        x = l2_normalize128(x.contiguous())
        out = [x] if self.return_embeddings else []
        for i in range(self.nmb_heads):
            out.append(lin16(x, self.prototypes(i)))
        return out

    def forward(self, feats: torch.Tensor) -> List[torch.Tensor]:
        if self._native_ok(feats):
            return self._forward_native(feats)
        x = self.projection_head(feats)
        if self.normalize_feats:
            x = F.normalize(x, dim=1, p=2)
        out = [x] if self.return_embeddings else []
        for i in range(self.nmb_heads):
            out.append(self.prototypes(i)(x))
        return out


class SwAVModel(nn.Module):
    """`trunk` + `heads.0` (VISSL parameter prefix `heads.0.prototypes0.weight`, cf. TEMP_FROZEN_PARAMS_ITER_MAP)."""

    def __init__(self, depth: int = 50, head_dims=(2048, 2048, 128), num_clusters=(3000,), use_bn: bool = True,
                 single_pass_every_crop: bool = False):
        super().__init__()
        self.trunk = ResNetTrunk(depth)
        dims = list(head_dims)
        dims[0] = self.trunk.out_dim
        self.heads = nn.ModuleList([SwAVPrototypesHead(dims, use_bn, num_clusters)])
        self.single_pass_every_crop = single_pass_every_crop

    def forward(self, crops: Sequence[torch.Tensor]) -> List[torch.Tensor]:
        """crops: list of [B, 3, H, W] tensors (2 x 224 + 6 x 96 in the reference recipe). Output rows are ordered
        crop-major: rows [c*B, (c+1)*B) belong to crop c."""
        if isinstance(crops, torch.Tensor):
            crops = [crops]
        feats = []
        if self.single_pass_every_crop:
            feats = [self.trunk(c) for c in crops]
        else:
            start = 0
            while start < len(crops):      # consecutive crops of equal resolution share a trunk pass
                end = start + 1
                while end < len(crops) and crops[end].shape[-1] == crops[start].shape[-1]:
                    end += 1
                feats.append(self.trunk(torch.cat(list(crops[start:end]), dim=0)))
                start = end
        return self.heads[0](torch.cat(feats, dim=0))

    @property
    def prototype_weights(self) -> List[nn.Parameter]:
        return [self.heads[0].prototypes(i).weight for i in range(self.heads[0].nmb_heads)]


class SwAVLoss(nn.Module):
    def __init__(self, temperature: float = 0.1, crops_for_assign: Sequence[int] = (0, 1), num_crops: int = 8,
                 num_iters: int = 3, epsilon: float = 0.03, num_prototypes: Sequence[int] = (3000,),
                 embedding_dim: int = 128, queue_length: int = 0, queue_start_iter: int = 0,
                 temp_hard_assignment_iters: int = 0, use_double_precision: bool = False, world_size: int = 1):
        super().__init__()
        self.temperature, self.crops_for_assign, self.num_crops = temperature, list(crops_for_assign), num_crops
        self.num_iters, self.epsilon, self.num_prototypes = num_iters, epsilon, list(num_prototypes)
        self.embedding_dim, self.temp_hard_assignment_iters = embedding_dim, temp_hard_assignment_iters
        self.use_double_precision, self.world_size = use_double_precision, world_size
        self.local_queue_length = queue_length // max(world_size, 1)
        self.queue_start_iter = queue_start_iter
        self.use_queue = False
        self.was_using_queue = False
        self.register_buffer("num_iteration", torch.zeros(1, dtype=torch.long))
        if self.local_queue_length > 0:
            nq = len(self.crops_for_assign)
            self.register_buffer("local_emb_queue", torch.zeros(nq, self.local_queue_length, embedding_dim))
            for i, k in enumerate(self.num_prototypes):
                self.register_buffer(f"local_queue{i}", torch.zeros(nq, self.local_queue_length, k))
            self.register_buffer("queue_head", torch.zeros(1, dtype=torch.long))   # ring-buffer write position

    # ---------------------------------------------------------------- the loss
    def forward(self, output: List[torch.Tensor], target=None, training_iterations: int = 0) -> torch.Tensor:
        """`training_iterations` is the collaboration-wide optimizer step
        (standard_train_step.py:153: `optimizer.optimizer.collaboration_state.optimizer_step`)."""
        self.use_queue = self.local_queue_length > 0 and training_iterations >= self.queue_start_iter
        if not self.was_using_queue and self.use_queue:
            logger.info(f"Using queue now! global niter = {training_iterations}")
            print(f"Using queue now! global niter = {training_iterations}")
        self.was_using_queue = self.use_queue
        loss = 0
        for head_id, scores in enumerate(output[1:]):
            loss = loss + self.criterion(scores, head_id)
        loss = loss / (len(output) - 1)
        self.num_iteration += 1
        if self.use_queue:
            self.update_emb_queue(output[0].detach())
        return loss

    def _ordered_queue(self, q: torch.Tensor) -> torch.Tensor:
        """Newest-first view of a ring buffer (what the reference keeps physically by rolling)."""
        head = int(self.queue_head)
        return q if head == 0 else torch.cat([q[head:], q[:head]], dim=0)

    def assignments(self, scores_this_crop: torch.Tensor, bs: int) -> torch.Tensor:
        fn = sinkhorn if scores_this_crop.is_cuda else sinkhorn_reference
        s = scores_this_crop.double().float() if self.use_double_precision else scores_this_crop
        if not s.is_contiguous():
            s = s.contiguous()
        q = fn(s if s.dtype in (torch.bfloat16, torch.float32) else s.float(), self.epsilon, self.num_iters, bs,
               self.world_size)
        if int(self.num_iteration) < self.temp_hard_assignment_iters:
            idx = q.argmax(dim=1, keepdim=True)
            q = torch.zeros_like(q).scatter_(1, idx, 1.0)
        return q

    def criterion(self, scores: torch.Tensor, head_id: int) -> torch.Tensor:
        assert scores.shape[0] % self.num_crops == 0
        bs = scores.shape[0] // self.num_crops
        from ..ops import swav_ops

        if swav_ops.swav_ce_supported(scores) and not self.use_double_precision:
            # C15: assignments first (Sinkhorn kernel, no gradient), then ONE kernel for log-softmax + all dot products
            with torch.no_grad():
                qs = []
                for i, crop_id in enumerate(self.crops_for_assign):
                    s = scores[bs * crop_id: bs * (crop_id + 1)].detach()
                    if self.use_queue:
                        s = torch.cat((s, getattr(self, f"local_queue{head_id}")[i].to(s.dtype)))
                    qs.append(self.assignments(s, bs).float())
                q_all = torch.stack(qs)
            return swav_ops.swav_cross_entropy(scores, q_all, self.crops_for_assign, self.num_crops, self.temperature)
        logp = F.log_softmax(scores.float() / self.temperature, dim=1).view(self.num_crops, bs, -1)
        total = 0
        for i, crop_id in enumerate(self.crops_for_assign):
            with torch.no_grad():
                s = scores[bs * crop_id: bs * (crop_id + 1)].detach()
                if self.use_queue:
                    queue = getattr(self, f"local_queue{head_id}")[i]
                    s = torch.cat((s, queue.to(s.dtype)))      # order inside the queue does not matter to Sinkhorn
                q = self.assignments(s, bs)                     # [bs, K]
            others = [p for p in range(self.num_crops) if p != crop_id]
            # - mean_b sum_k q[b,k] * logp[p,b,k], averaged over the other crops
            total = total - torch.einsum("bk,pbk->", q, logp[others]) / (bs * len(others))
        return total / len(self.crops_for_assign)

    # ---------------------------------------------------------------- queue
    @torch.no_grad()
    def update_emb_queue(self, emb: torch.Tensor) -> None:
        bs = emb.shape[0] // self.num_crops
        L = self.local_queue_length
        head = (int(self.queue_head) - bs) % L          # newest entries sit right before the previous head
        for i, crop_id in enumerate(self.crops_for_assign):
            new = emb[crop_id * bs:(crop_id + 1) * bs].to(self.local_emb_queue.dtype)
            idx = (head + torch.arange(bs, device=new.device)) % L
            self.local_emb_queue[i].index_copy_(0, idx, new[:L])
        self.queue_head.fill_(head)

    @torch.no_grad()
    def compute_queue_scores(self, head: SwAVPrototypesHead) -> None:
        """Re-score the stored embeddings with the current prototypes (SwAVUpdateQueueScoresHook)."""
        for i in range(len(self.crops_for_assign)):
            for h in range(head.nmb_heads):
                w = head.prototypes(h).weight
                emb, dst = self.local_emb_queue[i], getattr(self, f"local_queue{h}")[i]
                if (w.is_cuda and emb.shape[0] % 64 == 0 and emb.shape[1] % 64 == 0 and w.shape[0] % 8 == 0
                        and dst.dtype == torch.float32 and dst.is_contiguous()):
                    # C16: queue re-scoring on the tcgen05 GEMM, fp32 scores written in place
                    from ..ops.gemm import gemm

                    gemm(emb.to(torch.bfloat16).contiguous(), w.detach().to(torch.bfloat16).contiguous(), dst)
                else:
                    getattr(self, f"local_queue{h}")[i] = (emb.to(w.dtype) @ w.t()).float()


@torch.no_grad()
def normalize_prototypes(model: SwAVModel) -> None:
    """L2-normalise every prototype vector (NormalizePrototypesHook, swav_hooks.py:55-92). With the fused LARS
    kernel this is folded into the optimizer update (DEDLOC_TF_ROWNORM128) and this function is not called."""
    for w in model.prototype_weights:
        w.copy_(F.normalize(w.float(), dim=1, p=2).to(w.dtype))
