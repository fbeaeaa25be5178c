"""Building blocks shared by the ALBERT / sahajBERT trainers and bench.py: one collaborative peer on one GPU.

`build_albert_peer()` wires model -> FusedLamb (two weight-decay groups, albert/run_trainer.py:73-100) -> linear
warm-up schedule -> CollaborativeOptimizer; `PeerTrainer.micro_step()` is one forward/backward + `opt.step()`
(= one HF-Trainer inner-loop iteration of the reference), and `SyntheticMLMBatches` produces
WikiText-shaped synthetic batches (15 % masked tokens, 50 % swapped sentence order) in pinned host memory.
"""
from __future__ import annotations

import collections

import math
from dataclasses import dataclass
from typing import Dict, Iterator, List, Optional

import torch

from ..collab.optimizer import CollaborativeOptimizer
from ..dht import DHT
from ..models.albert import AlbertConfig, AlbertForPreTraining
from ..optim import Lamb
from ..optim.fused import FusedLamb
from ..optim.schedules import get_linear_schedule_with_warmup
from ..utils.compression import CompressionType


def albert_param_groups(model: torch.nn.Module, weight_decay: float) -> List[dict]:
    no_decay = ["bias", "LayerNorm.weight", "layer_norm.weight"]
    named = list(model.named_parameters())
    for n, p in named:
        p._dedloc_name = n
    return [
        {"params": [p for n, p in named if not any(nd in n for nd in no_decay)], "weight_decay": weight_decay},
        {"params": [p for n, p in named if any(nd in n for nd in no_decay)], "weight_decay": 0.0},
    ]


class SyntheticMLMBatches:
    """Random token batches with the shape and label statistics of tokenised WikiText-103 instances
    (tokenize_wikitext103.py: [CLS] A [SEP] B [SEP], max 512 tokens, SOP label; DataCollatorForLanguageModeling:
    15 % of the positions masked -> 80 % [MASK] / 10 % random / 10 % kept)."""

    def __init__(self, vocab_size: int, seq_len: int, batch_size: int, seed: int = 0, mlm_probability: float = 0.15,
                 pin: bool = True, pool: int = 8, mask_token_id: int = 4, cls_id: int = 2, sep_id: int = 3,
                 padded: bool = False, pad_id: int = 0):
        """`padded`: instance lengths are drawn from [seq_len / 2, seq_len] and the tail is padding with
        attention_mask 0, like tokenised WikiText instances (the default packs every sequence to full length)."""
        g = torch.Generator().manual_seed(seed)
        self.batches: List[Dict[str, torch.Tensor]] = []
        for _ in range(pool):
            ids = torch.randint(5, vocab_size, (batch_size, seq_len), generator=g)
            ids[:, 0] = cls_id
            split = torch.randint(seq_len // 4, 3 * seq_len // 4, (batch_size,), generator=g)
            pos = torch.arange(seq_len)[None]
            ids[pos == split[:, None]] = sep_id
            ids[:, -1] = sep_id
            token_type = (pos > split[:, None]).long()
            special = (ids == cls_id) | (ids == sep_id)
            masked = (torch.rand(batch_size, seq_len, generator=g) < mlm_probability) & ~special
            labels = torch.where(masked, ids, torch.full_like(ids, -100))
            r = torch.rand(batch_size, seq_len, generator=g)
            ids = torch.where(masked & (r < 0.8), torch.full_like(ids, mask_token_id), ids)
            rnd = torch.randint(5, vocab_size, (batch_size, seq_len), generator=g)
            ids = torch.where(masked & (r >= 0.8) & (r < 0.9), rnd, ids)
            attn = torch.ones_like(ids)
            if padded:
                length = torch.randint(seq_len // 2, seq_len + 1, (batch_size,), generator=g)
                keep = pos < length[:, None]
                ids = torch.where(keep, ids, torch.full_like(ids, pad_id))
                ids[torch.arange(batch_size), length - 1] = sep_id
                labels = torch.where(keep, labels, torch.full_like(labels, -100))
                token_type = token_type * keep
                attn = keep.long()
            batch = {"input_ids": ids, "token_type_ids": token_type, "labels": labels,
                     "attention_mask": attn,
                     "sentence_order_label": torch.randint(0, 2, (batch_size,), generator=g)}
            if pin and torch.cuda.is_available():
                batch = {k: v.pin_memory() for k, v in batch.items()}
            self.batches.append(batch)
        self.i = 0
        self.bytes_per_batch = sum(v.numel() * v.element_size() for v in self.batches[0].values())

    def next_host(self) -> Dict[str, torch.Tensor]:
        b = self.batches[self.i % len(self.batches)]
        self.i += 1
        return b

    def to_device(self, batch: Dict[str, torch.Tensor], device) -> Dict[str, torch.Tensor]:
        return {k: v.to(device, non_blocking=True) for k, v in batch.items()}


@dataclass
class PeerConfig:
    prefix: str = "albert"
    target_batch_size: int = 4096
    micro_batch: int = 32
    gradient_accumulation_steps: int = 1
    seq_len: int = 512
    learning_rate: float = 0.00176
    warmup_steps: int = 5000
    max_steps: int = 125_000
    adam_epsilon: float = 1e-6
    weight_decay: float = 0.01
    max_grad_norm: float = 1.0
    clamp_value: float = 10000.0
    compression: str = "FLOAT16"
    target_group_size: int = 256
    averaging_expiration: float = 5.0
    averaging_timeout: float = 30.0
    kernel_timeout: Optional[float] = None     # None = CollaborativeOptimizer's default (5 s)
    metadata_expiration: float = 30.0
    min_refresh_period: float = 0.05
    max_refresh_period: float = 1.0
    default_refresh_period: float = 0.2
    client_mode: bool = False
    impl: str = "native"
    # how many launched-but-unfinished micro-steps the host may run ahead of the GPU. The progress this peer
    # reports to the collaboration counts launched samples, so an unbounded queue makes every peer overshoot
    # target_batch_size by the queue depth; 1 keeps the GPU fed while the control plane (progress report, DHT
    # re-fetch, matchmaking) of a global step hides under the micro-step that is still running.
    max_inflight_micro_steps: int = 1
    # launch weight / bias gradient kernels on a side stream (joined before the gradients are read). Measured neutral
    # on B200 (804.0 vs 802.2 samples/s: the persistent GEMMs own the SMs and the step is power-capped), so off.
    overlap_wgrad: bool = False


class PeerTrainer:
    def __init__(self, model, opt, scheduler, collab: CollaborativeOptimizer, cfg: PeerConfig, device):
        self.model, self.opt, self.scheduler, self.collab, self.cfg, self.device = model, opt, scheduler, collab, cfg, device
        self.loss_sum = torch.zeros((), device=device, dtype=torch.float32)
        self.mini_steps = 0
        self.samples_processed = 0
        self._graph = None
        self._static_batch: Optional[Dict[str, torch.Tensor]] = None
        self._graph_launches = 0
        self._inflight: collections.deque = collections.deque()

    # ---------------------------------------------------------------- CUDA-graph capture of the micro-step
    def _forward_backward(self, batch: Dict[str, torch.Tensor]) -> None:
        chunks = self.cfg.gradient_accumulation_steps
        n = batch["input_ids"].shape[0]
        overlap = self.device.type == "cuda" and self.cfg.overlap_wgrad
        if overlap:
            from ..ops import gemm as _gemm

            _gemm.overlap_wgrad(True)          # weight / bias gradients go to a side stream ...
        try:
            for c in range(chunks):
                sub = {k: v[c * n // chunks:(c + 1) * n // chunks] for k, v in batch.items()} if chunks > 1 else batch
                out = self.model(**sub)
                (out.loss / chunks).backward()
                self.loss_sum += out.loss.detach().float() / chunks
        finally:
            if overlap:
                _gemm.overlap_wgrad(False)
                _gemm.join_wgrad_stream(self.device)   # ... and are joined before anybody reads the gradients
        if hasattr(self.model, "flush_fused_grads"):
            self.model.flush_fused_grads()

    def enable_cuda_graph(self, example_batch: Dict[str, torch.Tensor], warmup: int = 3) -> bool:
        """Capture forward + backward (+ fused-gradient flush) of one micro-batch into a CUDA graph: the launch-bound
        inner loop (~600 kernels) then costs one graph launch per micro-batch. Returns False (and stays eager) when
        the capture is not possible."""
        if self.device.type != "cuda" or self._graph is not None:
            return self._graph is not None
        from .. import lib as native

        try:
            self._static_batch = {k: v.to(self.device).clone() for k, v in example_batch.items()}
            side = torch.cuda.Stream(self.device)
            side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(side):
                for _ in range(warmup):
                    self._forward_backward(self._static_batch)
            torch.cuda.current_stream(self.device).wait_stream(side)
            torch.cuda.synchronize(self.device)
            self.collab.zero_grad()          # the warm-up gradients are not part of training
            self.loss_sum.zero_()
            graph = torch.cuda.CUDAGraph()
            before = native.launch_count()
            with torch.cuda.graph(graph):
                self._forward_backward(self._static_batch)
            self._graph_launches = native.launch_count() - before
            self.collab.zero_grad()          # capture does not execute, but keep the buffers clean regardless
            self.loss_sum.zero_()
            self._graph = graph
            return True
        except Exception as e:  # pragma: no cover
            import logging

            logging.getLogger(__name__).warning("CUDA-graph capture failed, staying eager: %r", e)
            self._graph, self._static_batch = None, None
            torch.cuda.synchronize(self.device)
            return False

    def micro_step(self, batch: Dict[str, torch.Tensor]) -> bool:
        """One local step: forward/backward over `gradient_accumulation_steps` x micro_batch samples, then
        CollaborativeOptimizer.step(). Returns True when a collaborative (global) step happened."""
        n = batch["input_ids"].shape[0]
        if self._graph is not None:
            from .. import lib as native

            for k, v in batch.items():       # host (pinned) or device source -> static graph inputs
                self._static_batch[k].copy_(v, non_blocking=True)
            self._graph.replay()
            native.lib().dedloc_add_launch_count(self._graph_launches)
        else:
            self._forward_backward(batch)
        self.mini_steps += 1
        self.samples_processed += n
        if self.device.type == "cuda":
            done = torch.cuda.Event()
            done.record()
            self._inflight.append(done)
            while len(self._inflight) > max(self.cfg.max_inflight_micro_steps, 0):
                self._inflight.popleft().synchronize()
        return bool(self.collab.step(batch_size=n))

    def pop_mean_loss(self) -> float:
        """Device->host read of the mean loss since the last call (the per-step result of the e2e timing)."""
        val = float(self.loss_sum.item()) / max(self.mini_steps, 1)
        if hasattr(self.model, "mlm_dropped_positions"):
            dropped = self.model.mlm_dropped_positions()
            if dropped:
                import logging

                logging.getLogger(__name__).warning(
                    "%d masked positions exceeded the MLM capacity (max_predictions_ratio) and were not scored: raise "
                    "AlbertForPreTraining(max_predictions_ratio=...) for collators that mask more than 25 %%", dropped)
        self.loss_sum.zero_()
        self.mini_steps = 0
        return val


def build_albert_peer(dht: DHT, cfg: PeerConfig, model_cfg: Optional[AlbertConfig] = None, device=None,
                      seed: int = 0, peer_key: Optional[bytes] = None, **collab_kwargs) -> PeerTrainer:
    device = torch.device(device if device is not None else ("cuda" if torch.cuda.is_available() else "cpu"))
    model_cfg = model_cfg or AlbertConfig.albert_large_v2()
    torch.manual_seed(seed)
    on_gpu = device.type == "cuda"
    model = AlbertForPreTraining(model_cfg, impl=cfg.impl if on_gpu else "torch")
    if on_gpu:
        # fp32 initial values become the master copy; the module itself computes in bf16
        for p in model.parameters():
            p._dedloc_master = p.detach().clone().to(device)
        model = model.to(device=device, dtype=torch.bfloat16)
        for p in model.parameters():
            pass
    groups = albert_param_groups(model, cfg.weight_decay)
    if on_gpu:
        opt = FusedLamb(groups, lr=cfg.learning_rate, betas=(0.9, 0.999), eps=cfg.adam_epsilon,
                        weight_decay=cfg.weight_decay, clamp_value=cfg.clamp_value, debias=True)
    else:
        opt = Lamb(groups, lr=cfg.learning_rate, betas=(0.9, 0.999), eps=cfg.adam_epsilon,
                   weight_decay=cfg.weight_decay, clamp_value=cfg.clamp_value, debias=True)
    scheduler = get_linear_schedule_with_warmup(opt, cfg.warmup_steps, cfg.max_steps)
    collab = CollaborativeOptimizer(
        opt, dht=dht, prefix=cfg.prefix, target_batch_size=cfg.target_batch_size,
        batch_size_per_step=cfg.micro_batch, scheduler=scheduler,
        compression_type=CompressionType.Value(cfg.compression), target_group_size=cfg.target_group_size,
        averaging_expiration=cfg.averaging_expiration, averaging_timeout=cfg.averaging_timeout,
        kernel_timeout=cfg.kernel_timeout, metadata_expiration=cfg.metadata_expiration, min_refresh_period=cfg.min_refresh_period,
        max_refresh_period=cfg.max_refresh_period, default_refresh_period=cfg.default_refresh_period,
        client_mode=cfg.client_mode, max_grad_norm=cfg.max_grad_norm, peer_key=peer_key, **collab_kwargs)
    model.train()
    return PeerTrainer(model, opt, scheduler, collab, cfg, device)
