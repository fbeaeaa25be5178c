"""Hooks of the collaborative SwAV run (everything enabled by W4 through vissl/hooks/__init__.py:72-146):

SwAVUpdateQueueScoresHook, NormalizePrototypesHook (swav_hooks.py:11-92), CheckNanLossHook, SetDataSamplerEpochHook,
FreezeParametersHook
(state_update_hooks.py:207-280), UpdateBatchesSeenHook / UpdateTrainIterationNumHook / UpdateTrainBatchTimeHook,
LogLossLrEtaHook, LogPerfTimeMetricsHook, LogGpuStatsHook, LogGpuMemoryHook, LogLossMetricsCheckpointHook
(log_hooks.py), SSLTensorboardHook, SSLModelComplexityHook (state_update_hooks.py:16-63), ProfilerHook (ClassyVision). Hook points follow ClassyHook: on_start, on_forward, on_loss_and_meter,
on_backward, on_update, on_step, on_phase_end, on_end."""
from __future__ import annotations

import json
import logging
import os
import subprocess
import time
from typing import List

import torch

from ...models.swav import normalize_prototypes

logger = logging.getLogger(__name__)


class ClassyHook:
    def on_start(self, task): pass
    def on_phase_start(self, task): pass
    def on_forward(self, task): pass
    def on_loss_and_meter(self, task): pass
    def on_backward(self, task): pass
    def on_update(self, task): pass
    def on_step(self, task): pass
    def on_phase_end(self, task): pass
    def on_end(self, task): pass


class SwAVUpdateQueueScoresHook(ClassyHook):
    """Before the loss: re-score the queued embeddings with the current prototypes when the queue is in use."""

    def on_forward(self, task):
        loss = task.loss
        if loss.local_queue_length > 0 and loss.use_queue:
            loss.compute_queue_scores(task.model.heads[0])


class NormalizePrototypesHook(ClassyHook):
    """After every parameter update: L2-normalise the prototypes. With the fused LARS kernel the normalisation
    happens inside the optimizer update (tensor flag ROWNORM128) and this hook has nothing left to do."""

    def on_start(self, task):
        if not getattr(task, "prototypes_normalized_by_kernel", False):
            normalize_prototypes(task.model)

    def on_update(self, task):
        if not getattr(task, "prototypes_normalized_by_kernel", False) and task.last_step_was_global:
            normalize_prototypes(task.model)


class FreezeParametersHook(ClassyHook):
    """Zero the gradients of the parameters named in MODEL.TEMP_FROZEN_PARAMS_ITER_MAP while the LOCAL iteration is
    below their threshold (prototypes frozen for the first 5005 iterations in the recipe)."""

    def __init__(self, frozen_map):
        self.map = {name.replace("module.", "", 1): int(it) for name, it in frozen_map}
        self._params = None

    def on_start(self, task):
        named = dict(task.model.named_parameters())
        self._params = {n: named[n] for n in self.map if n in named}
        missing = set(self.map) - set(self._params)
        if missing:
            logger.warning("TEMP_FROZEN_PARAMS_ITER_MAP names not found in the model: %s", sorted(missing))
        self._apply(task)

    def _apply(self, task):
        for name, p in self._params.items():
            frozen = task.iteration < self.map[name]
            task.set_param_frozen(p, frozen)

    def on_backward(self, task):
        for name, p in self._params.items():
            if task.iteration < self.map[name]:
                if p.grad is not None:
                    p.grad.zero_()
                mg = getattr(p, "main_grad", None)
                if mg is not None:
                    mg.zero_()
        self._apply(task)


class CheckNanLossHook(ClassyHook):
    def on_loss_and_meter(self, task):
        if task.check_nan_every > 0 and task.iteration % task.check_nan_every == 0:
            if not torch.isfinite(task.last_loss).item():
                raise FloatingPointError(f"Infinite Loss or NaN at iteration={task.iteration}")


class UpdateTrainIterationNumHook(ClassyHook):
    def on_forward(self, task):
        task.iteration += 1


class UpdateBatchesSeenHook(ClassyHook):
    def on_forward(self, task):
        task.batches += 1


class SetDataSamplerEpochHook(ClassyHook):
    """New shuffle at the start of every phase (vissl/hooks/state_update_hooks.py SetDataSamplerEpochHook)."""

    def on_phase_start(self, task):
        sampler = getattr(task.dataloader, "sampler", None)
        if hasattr(sampler, "set_epoch"):
            sampler.set_epoch(task.phase_idx)


class UpdateTestBatchTimeHook(ClassyHook):
    """Batch times of evaluation phases (the collaborative recipe has none; kept for recipes that add TEST_MODEL)."""

    def on_start(self, task):
        task.test_batch_time, self._t = [], time.perf_counter()

    def on_forward(self, task):
        if not getattr(task, "train", True):
            now = time.perf_counter()
            task.test_batch_time.append(now - self._t)
            self._t = now


class UpdateTrainBatchTimeHook(ClassyHook):
    def on_start(self, task):
        task.batch_time, self._t = [], time.perf_counter()

    def on_step(self, task):
        now = time.perf_counter()
        task.batch_time.append(now - self._t)
        self._t = now


class LogLossLrEtaHook(ClassyHook):
    def __init__(self, log_frequency: int):
        self.freq = max(int(log_frequency), 1)

    def on_update(self, task):
        if task.iteration % self.freq == 0 or task.iteration <= 5:
            lr = task.optimizer.optimizer.param_groups[0]["lr"]
            bt = sum(task.batch_time[-self.freq:]) / max(len(task.batch_time[-self.freq:]), 1) if task.batch_time else 0.0
            mem = torch.cuda.max_memory_allocated() / 2 ** 20 if torch.cuda.is_available() else 0
            collab = task.optimizer.optimizer
            gstep = getattr(getattr(collab, "collaboration_state", None), "optimizer_step", None)
            logger.info(f"Rank: 0; [it: {task.iteration}]; lr: {lr:.5g}; loss: {float(task.last_loss):.5f}; "
                        f"btime(ms): {1e3 * bt:.1f}; global step: {gstep}; max_mem: {mem:.0f}M")


class LogPerfTimeMetricsHook(ClassyHook):
    def __init__(self, rolling_freq: int):
        self.freq = rolling_freq

    def on_update(self, task):
        if task.perf_stats is not None and self.freq > 0 and task.iteration % self.freq == 0:
            logger.info("Performance stats:\n" + task.perf_stats.report_str())

    def on_phase_end(self, task):
        if task.perf_stats is not None:
            logger.info("Performance stats (phase):\n" + task.perf_stats.report_str())


class LogGpuStatsHook(ClassyHook):
    def on_start(self, task):
        try:
            out = subprocess.run(["nvidia-smi"], capture_output=True, text=True, timeout=10).stdout
            logger.info("nvidia-smi:\n" + out)
        except Exception:
            pass


class LogGpuMemoryHook(ClassyHook):
    def __init__(self, iteration_num: int):
        self.iteration_num = iteration_num

    def on_step(self, task):
        if torch.cuda.is_available() and task.iteration == self.iteration_num + 1:
            logger.info(torch.cuda.memory_summary())


def model_complexity(model: torch.nn.Module, input_shape, device=None, batch: int = 2) -> dict:
    """#parameters, forward multiply-accumulates and activations (outputs of convolutions / linear layers) PER
    SAMPLE for one crop of `input_shape`, counted with forward hooks on a dummy forward in eval mode
    (classy_vision/generic/profiler.py:89-560 counts the same three; only conv / linear layers carry MACs here)."""
    device = device or next(model.parameters()).device
    macs, acts = [0], [0]

    def conv_hook(m, inp, out):
        kh, kw = m.kernel_size
        macs[0] += out.numel() * (m.in_channels // m.groups) * kh * kw
        acts[0] += out.numel()

    def linear_hook(m, inp, out):
        macs[0] += out.numel() * m.in_features
        acts[0] += out.numel()

    handles = []
    for m in model.modules():
        if isinstance(m, torch.nn.Conv2d):
            handles.append(m.register_forward_hook(conv_hook))
        elif isinstance(m, torch.nn.Linear):
            handles.append(m.register_forward_hook(linear_hook))
    was_training = model.training
    model.eval()
    try:
        with torch.no_grad():
            dtype = next(model.parameters()).dtype
            model([torch.zeros(batch, *input_shape, device=device, dtype=dtype)])
    finally:
        for h in handles:
            h.remove()
        model.train(was_training)
    return {"params": sum(p.numel() for p in model.parameters()), "macs_per_sample": macs[0] // batch,
            "flops_per_sample": 2 * macs[0] // batch, "activations_per_sample": acts[0] // batch}


class SSLModelComplexityHook(ClassyHook):
    """HOOKS.MODEL_COMPLEXITY.COMPUTE_COMPLEXITY: log #params / #flops / #activations once at start
    (vissl/hooks/state_update_hooks.py:16-63)."""

    def __init__(self, input_shape):
        self.input_shape = tuple(input_shape)

    def on_start(self, task):
        try:
            c = model_complexity(task.model, self.input_shape, task.device)
        except Exception as e:       # as in the reference: a model that cannot take the dummy input is not an error
            logger.info(f"model complexity not computed: {e!r}")
            return
        task.model_complexity = c
        logger.info(f"Model complexity for input {list(self.input_shape)}: params {c['params'] / 1e6:.3f} M, "
                    f"forward {c['flops_per_sample'] / 1e9:.3f} GFLOPs ({c['macs_per_sample'] / 1e9:.3f} GMACs), "
                    f"activations {c['activations_per_sample'] / 1e6:.3f} M")


class ProfilerHook(ClassyHook):
    """One profiled dummy forward at start, table sorted by device (or host) time -- classy_vision/hooks/
    profiler_hook.py:15-38. Off unless HOOKS.PROFILER.PROFILE_FORWARD is set; for kernel-level work use ncu / NVTX
    (`DEDLOC_NVTX=1`), this is the quick look the reference offers."""

    def __init__(self, input_shape, row_limit: int = 25):
        self.input_shape, self.row_limit = tuple(input_shape), row_limit

    def on_start(self, task):
        from torch.profiler import ProfilerActivity, profile

        acts = [ProfilerActivity.CPU] + ([ProfilerActivity.CUDA] if task.device.type == "cuda" else [])
        model = task.model
        was_training = model.training
        model.eval()
        try:
            with torch.no_grad(), profile(activities=acts) as prof:
                dtype = next(model.parameters()).dtype
                model([torch.zeros(2, *self.input_shape, device=task.device, dtype=dtype)])
            key = "cuda_time_total" if task.device.type == "cuda" else "cpu_time_total"
            task.profile_table = prof.key_averages().table(sort_by=key, row_limit=self.row_limit)
            logger.info("Forward profile:\n" + task.profile_table)
        except Exception as e:
            logger.info(f"forward profile not taken: {e!r}")
        finally:
            model.train(was_training)


class LogLossMetricsCheckpointHook(ClassyHook):
    """VISSL-shaped checkpoints (log_hooks.py:268-417): `model_iteration{N}.torch` every CHECKPOINT_ITER_FREQUENCY
    iterations, `model_phase{N}.torch` every CHECKPOINT_FREQUENCY phases, `model_final_checkpoint_phase{N}.torch`,
    `checkpoint.torch` symlink to the newest, `metrics.json` lines. The `optimizer` entry is None for
    sgd_collaborative; the SwAV queue rides in `loss`."""

    def __init__(self, cfg):
        self.dir = cfg.CHECKPOINT.DIR
        self.iter_freq = cfg.CHECKPOINT.CHECKPOINT_ITER_FREQUENCY
        self.phase_freq = cfg.CHECKPOINT.CHECKPOINT_FREQUENCY

    def on_step(self, task):
        if self.iter_freq > 0 and task.iteration > 0 and task.iteration % self.iter_freq == 0:
            self._save(task, f"model_iteration{task.iteration}.torch")

    def on_phase_end(self, task):
        final = task.phase_idx + 1 >= task.num_phases
        if final:
            self._save(task, f"model_final_checkpoint_phase{task.phase_idx}.torch")
        elif self.phase_freq > 0 and (task.phase_idx + 1) % self.phase_freq == 0:
            self._save(task, f"model_phase{task.phase_idx}.torch")
        with open(os.path.join(self.dir, "metrics.json"), "a") as f:
            f.write(json.dumps({"iteration": task.iteration, "phase_idx": task.phase_idx,
                                "loss": float(task.last_loss) if task.last_loss is not None else None}) + "\n")

    def _save(self, task, name):
        os.makedirs(self.dir, exist_ok=True)
        state = {"phase_idx": task.phase_idx, "iteration": task.iteration, "iteration_num": task.iteration,
                 "train_phase_idx": task.phase_idx,
                 "loss": float(task.last_loss) if task.last_loss is not None else None,
                 "classy_state_dict": task.get_classy_state()}
        path = os.path.join(self.dir, name)
        torch.save(state, path)
        link = os.path.join(self.dir, "checkpoint.torch")
        if os.path.islink(link) or os.path.exists(link):
            os.remove(link)
        os.symlink(name, link)
        logger.info(f"Saved checkpoint: {path}")


class SSLTensorboardHook(ClassyHook):
    def __init__(self, cfg):
        tb = cfg.HOOKS.TENSORBOARD_SETUP
        self.writer = None
        try:
            from torch.utils.tensorboard import SummaryWriter

            self.writer = SummaryWriter(log_dir=os.path.join(tb.LOG_DIR, tb.EXPERIMENT_LOG_DIR),
                                        flush_secs=60 * tb.FLUSH_EVERY_N_MIN)
        except Exception as e:
            logger.warning("tensorboard unavailable: %r", e)
        self.log_params, self.log_grads = tb.LOG_PARAMS, tb.LOG_PARAMS_GRADIENTS
        self.every = tb.LOG_PARAMS_EVERY_N_ITERS

    def on_update(self, task):
        if self.writer is None:
            return
        it = task.iteration
        self.writer.add_scalar("Training/Loss", float(task.last_loss), it)
        self.writer.add_scalar("Training/Learning_rate", task.optimizer.optimizer.param_groups[0]["lr"], it)
        if self.every > 0 and it % self.every == 0 and self.log_params:
            for name, p in task.model.named_parameters():
                self.writer.add_histogram(f"Parameters/{name}", p.detach().float().cpu(), it)

    def on_end(self, task):
        if self.writer is not None:
            self.writer.close()


def default_hook_generator(cfg) -> List[ClassyHook]:
    hooks: List[ClassyHook] = []
    if cfg.LOSS.name == "swav_loss":
        hooks += [SwAVUpdateQueueScoresHook(), NormalizePrototypesHook()]
    hooks += [CheckNanLossHook(), SetDataSamplerEpochHook(), UpdateTrainIterationNumHook(), UpdateBatchesSeenHook(),
              UpdateTrainBatchTimeHook(), UpdateTestBatchTimeHook()]
    if cfg.MODEL.TEMP_FROZEN_PARAMS_ITER_MAP:
        hooks.append(FreezeParametersHook(cfg.MODEL.TEMP_FROZEN_PARAMS_ITER_MAP))
    if cfg.HOOKS.PERF_STATS.MONITOR_PERF_STATS:
        hooks.append(LogPerfTimeMetricsHook(cfg.HOOKS.PERF_STATS.ROLLING_BTIME_FREQ))
    if cfg.HOOKS.LOG_GPU_STATS:
        hooks.append(LogGpuStatsHook())
    if cfg.HOOKS.MEMORY_SUMMARY.PRINT_MEMORY_SUMMARY:
        hooks.append(LogGpuMemoryHook(cfg.HOOKS.MEMORY_SUMMARY.LOG_ITERATION_NUM))
    if cfg.HOOKS.TENSORBOARD_SETUP.USE_TENSORBOARD:
        hooks.append(SSLTensorboardHook(cfg))
    if cfg.HOOKS.MODEL_COMPLEXITY.COMPUTE_COMPLEXITY:
        hooks.append(SSLModelComplexityHook(cfg.HOOKS.MODEL_COMPLEXITY.INPUT_SHAPE))
    if cfg.HOOKS.PROFILER.PROFILE_FORWARD:
        hooks.append(ProfilerHook(cfg.HOOKS.MODEL_COMPLEXITY.INPUT_SHAPE))
    hooks += [LogLossLrEtaHook(cfg.LOG_FREQUENCY), LogLossMetricsCheckpointHook(cfg)]
    return hooks
