"""Self-supervision task + trainer for collaborative SwAV (SURVEY 3.4 call stack):

train_main -> SelfSupervisionTrainer.train -> task.prepare (dataloader, model = resnet trunk + swav_head, loss,
optimizer = build_optimizer("sgd_collaborative"), param groups) -> for every phase: `standard_train_step` until the
phase's iterations are done (vissl/trainer/train_steps/standard_train_step.py:87-229): read_sample -> forward ->
on_forward hooks -> loss(model_output, target, collaboration_state.optimizer_step) -> on_loss_and_meter hooks ->
zero_grad + backward -> on_backward hooks -> optimizer.step -> on_update hooks. bf16 autocast replaces apex AMP O1
(no loss scaling needed on B200); activation checkpointing and SyncBN are config keys that stay off (180 GB HBM).
"""
from __future__ import annotations

import glob
import logging
import os
from typing import Any, Dict, List, Optional

import torch

from ...models.swav import SwAVLoss, SwAVModel
from ...utils.perf_stats import PerfStats, PerfTimer
from .config import AttrDict
from .data import build_dataloader
from .hooks import ClassyHook, default_hook_generator
from .optim import build_optimizer, get_optimizer_param_groups

logger = logging.getLogger(__name__)
_INIT_LOCK = __import__("threading").Lock()   # seed + parameter init must not interleave between in-process peers


class SelfSupervisionTask:
    def __init__(self, cfg: AttrDict, device: torch.device, dht=None, collab_kwargs: Optional[dict] = None):
        self.cfg, self.device = cfg, device
        self._dht, self._collab_kwargs = dht, collab_kwargs
        self.model: Optional[SwAVModel] = None
        self.loss: Optional[SwAVLoss] = None
        self.optimizer = None
        self.dataloader = None
        self.iteration, self.batches, self.phase_idx = 0, 0, 0
        self.num_phases = int(cfg.OPTIMIZER.num_epochs)
        self.last_loss: Optional[torch.Tensor] = None
        self.last_step_was_global = False
        self.batch_time: List[float] = []
        self.check_nan_every = 1
        self.perf_stats = PerfStats() if cfg.HOOKS.PERF_STATS.MONITOR_PERF_STATS else None
        self.prototypes_normalized_by_kernel = False
        self.amp_dtype = torch.bfloat16 if (cfg.MODEL.AMP_PARAMS.USE_AMP and device.type == "cuda") else None

    # ---------------------------------------------------------------- construction
    def prepare(self, dataloader=None) -> "SelfSupervisionTask":
        cfg = self.cfg
        self.dataloader = dataloader if dataloader is not None else build_dataloader(cfg, self.device)
        head = [p for p in cfg.MODEL.HEAD.PARAMS if p[0] == "swav_head"][0][1]
        with _INIT_LOCK:   # every peer starts from the same SEED_VALUE -> identical initial parameters
            torch.manual_seed(cfg.SEED_VALUE)
            self.model = SwAVModel(depth=cfg.MODEL.TRUNK.TRUNK_PARAMS.RESNETS.DEPTH, head_dims=head["dims"],
                                   num_clusters=head["num_clusters"], use_bn=head.get("use_bn", True),
                                   single_pass_every_crop=cfg.MODEL.SINGLE_PASS_EVERY_CROP and self.device.type != "cuda")
        self.model.to(self.device)
        if self.device.type == "cuda":
            self.model.to(memory_format=torch.channels_last)
        lc = cfg.LOSS.swav_loss
        world = cfg.DISTRIBUTED.NUM_NODES * cfg.DISTRIBUTED.NUM_PROC_PER_NODE
        self.loss = SwAVLoss(temperature=lc.temperature, crops_for_assign=lc.crops_for_assign, num_crops=lc.num_crops,
                             num_iters=lc.num_iters, epsilon=lc.epsilon, num_prototypes=lc.num_prototypes,
                             embedding_dim=lc.embedding_dim, queue_length=lc.queue.queue_length,
                             queue_start_iter=lc.queue.start_iter,
                             temp_hard_assignment_iters=lc.get("temp_hard_assignment_iters", 0),
                             use_double_precision=lc.use_double_precision, world_size=world).to(self.device)
        opt_cfg = dict(cfg.OPTIMIZER)
        if self._dht is not None:
            opt_cfg["dht"] = self._dht
        if self._collab_kwargs:
            opt_cfg["collab_kwargs"] = self._collab_kwargs
        self.optimizer = build_optimizer(opt_cfg)
        groups = get_optimizer_param_groups(self.model, cfg.OPTIMIZER)
        self.optimizer.set_param_groups(groups)
        self._wire_fused_features()
        self._maybe_resume()
        return self

    def _wire_fused_features(self) -> None:
        """When the inner optimizer is the fused LARS kernel: prototypes are re-normalised inside the update."""
        from ... import lib as native
        from ...optim.fused import FusedLARS

        inner = getattr(self.optimizer.optimizer, "opt", None)
        if isinstance(inner, FusedLARS) and self.cfg.LOSS.swav_loss.normalize_last_layer:
            ok = all(w.shape[1] == 128 for w in self.model.prototype_weights)
            if ok:
                for w in self.model.prototype_weights:
                    inner.set_tensor_flag(w, native.TF_ROWNORM128, True)
                self.prototypes_normalized_by_kernel = True
        self._fused_inner = inner if isinstance(inner, FusedLARS) else None

    def set_param_frozen(self, param, frozen: bool) -> None:
        if getattr(self, "_fused_inner", None) is not None:
            from ... import lib as native

            self._fused_inner.set_tensor_flag(param, native.TF_FROZEN, frozen)

    # ---------------------------------------------------------------- checkpoints
    def get_classy_state(self) -> Dict[str, Any]:
        return {"train": True, "base_model": {"model": {"trunk": self.model.trunk.state_dict(),
                                                        "heads": self.model.heads.state_dict()}},
                "meters": [], "optimizer": self.optimizer.get_classy_state(), "phase_idx": self.phase_idx,
                "train_phase_idx": self.phase_idx, "num_updates": self.iteration, "losses": [],
                "hooks": {}, "loss": self.loss.state_dict(), "amp": None}

    def set_classy_state(self, state: Dict[str, Any]) -> None:
        self.model.trunk.load_state_dict(state["base_model"]["model"]["trunk"])
        self.model.heads.load_state_dict(state["base_model"]["model"]["heads"])
        self.loss.load_state_dict(state["loss"])
        self.phase_idx = state.get("train_phase_idx", 0)
        self.iteration = state.get("num_updates", 0)
        # resume in the middle of a phase: skip what was consumed (train_task.py:383-401 does it on the sampler)
        loader = getattr(self, "dataloader", None)
        sampler = getattr(loader, "sampler", None)
        if hasattr(sampler, "set_start_iter") and hasattr(loader, "__len__") and len(loader) > 0:
            sampler.set_epoch(self.phase_idx)
            sampler.set_start_iter(self.iteration % len(loader))

    def _maybe_resume(self) -> None:
        ck = self.cfg.CHECKPOINT
        if not ck.AUTO_RESUME:
            return
        link = os.path.join(ck.DIR, "checkpoint.torch")
        cands = [link] if os.path.exists(link) else sorted(glob.glob(os.path.join(ck.DIR, "model_*.torch")),
                                                           key=os.path.getmtime)
        if cands:
            # the VISSL-shaped checkpoint is tensors and plain containers only: never unpickle arbitrary objects
            state = torch.load(cands[-1], map_location="cpu", weights_only=True)
            self.set_classy_state(state["classy_state_dict"])
            logger.info(f"Resumed from {cands[-1]} at iteration {self.iteration}")

    @property
    def collaboration_step(self) -> int:
        collab = self.optimizer.optimizer
        cs = getattr(collab, "collaboration_state", None)
        return int(cs.optimizer_step) if cs is not None else self.iteration


def standard_train_step(task: SelfSupervisionTask, sample: Dict[str, Any], hooks: List[ClassyHook]) -> SelfSupervisionTask:
    ps = task.perf_stats
    with PerfTimer("train_step_total", ps):
        with PerfTimer("forward", ps):
            inputs = sample["input"]
            if task.device.type == "cuda":
                inputs = [c.to(task.device, non_blocking=True).contiguous(memory_format=torch.channels_last)
                          for c in inputs]
            if task.amp_dtype is not None:
                with torch.autocast("cuda", dtype=task.amp_dtype):
                    output = task.model(inputs)
            else:
                output = task.model(inputs)
        for h in hooks:
            h.on_forward(task)
        with PerfTimer("loss_compute", ps):
            loss = task.loss([o.float() for o in output], sample.get("label"), task.collaboration_step)
        task.last_loss = loss.detach()
        with PerfTimer("loss_all_reduce", ps):
            pass   # world_size == 1 per peer: metrics travel through the DHT, not NCCL
        with PerfTimer("meters_update", ps):
            for h in hooks:
                h.on_loss_and_meter(task)
        with PerfTimer("backward", ps):
            task.optimizer.zero_grad()
            loss.backward()
        for h in hooks:
            h.on_backward(task)
        with PerfTimer("optimizer_step", ps):
            before = task.optimizer.optimizer.local_step if hasattr(task.optimizer.optimizer, "local_step") else None
            task.optimizer.step(where=0.0)
            after = task.optimizer.optimizer.local_step if before is not None else None
            task.last_step_was_global = before is None or after != before
        for h in hooks:
            h.on_update(task)
    for h in hooks:
        h.on_step(task)
    return task


class SelfSupervisionTrainer:
    def __init__(self, cfg: AttrDict, device: Optional[torch.device] = None, hooks: Optional[List[ClassyHook]] = None,
                 dht=None, collab_kwargs: Optional[dict] = None, dataloader=None, max_iterations: int = 0,
                 max_global_steps: int = 0):
        want_gpu = cfg.MACHINE.DEVICE == "gpu" and torch.cuda.is_available()
        self.device = device or torch.device("cuda" if want_gpu else "cpu")
        self.cfg = cfg
        self.task = SelfSupervisionTask(cfg, self.device, dht, collab_kwargs).prepare(dataloader)
        self.hooks = hooks if hooks is not None else default_hook_generator(cfg)
        self.max_iterations, self.max_global_steps = max_iterations, max_global_steps

    def train(self) -> SelfSupervisionTask:
        task, hooks = self.task, self.hooks
        for h in hooks:
            h.on_start(task)
        done = False
        while task.phase_idx < task.num_phases and not done:
            for h in hooks:
                h.on_phase_start(task)
            it = iter(task.dataloader)
            while True:
                with PerfTimer("read_sample", task.perf_stats):
                    try:
                        sample = next(it)
                    except StopIteration:
                        break
                standard_train_step(task, sample, hooks)
                if self.max_iterations and task.iteration >= self.max_iterations:
                    done = True
                    break
                if self.max_global_steps and getattr(task.optimizer.optimizer, "local_step", 0) >= self.max_global_steps:
                    done = True
                    break
            for h in hooks:
                h.on_phase_end(task)
            task.phase_idx += 1
        for h in hooks:
            h.on_end(task)
        return task


def train_main(cfg: AttrDict, **kwargs) -> SelfSupervisionTask:
    logging.basicConfig(level=logging.INFO, format="%(asctime)s [%(levelname)s] %(message)s")
    os.makedirs(cfg.CHECKPOINT.DIR, exist_ok=True)
    return SelfSupervisionTrainer(cfg, **kwargs).train()
