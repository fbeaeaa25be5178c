"""Named-section timers with lazily resolved CUDA events (behaviour of VISSL's PerfTimer / PerfStats,
swav/vissl/vissl/utils/perf_stats.py:12-250, used by standard_train_step.py:110-227). Host time is always
recorded; device time is recorded with a CUDA event pair per section and resolved only when the events have
completed, so timing never forces a synchronisation. Section names of the train step:
read_sample, forward, loss_compute, loss_all_reduce, meters_update, backward, optimizer_step, train_step_total,
plus `exposed_averaging` for the collaborative step. Every section is also an NVTX range (`nvtx_range`),
so an nsys / ncu timeline shows the same names; set DEDLOC_NVTX=0 to switch the ranges off."""
from __future__ import annotations

import time
from collections import defaultdict
from typing import Dict, List, Optional, Tuple

import torch


import os
from contextlib import contextmanager

_NVTX = os.environ.get("DEDLOC_NVTX", "1") != "0"


def nvtx_push(name: str) -> bool:
    """Open an NVTX range when CUDA is up (a no-op on CPU-only machines). Returns whether a range was opened."""
    if _NVTX and torch.cuda.is_available():
        try:
            torch.cuda.nvtx.range_push(name)
            return True
        except Exception:  # pragma: no cover - nvtx missing from the build
            return False
    return False


def nvtx_pop(opened: bool) -> None:
    if opened:
        torch.cuda.nvtx.range_pop()


@contextmanager
def nvtx_range(name: str):
    opened = nvtx_push(name)
    try:
        yield
    finally:
        nvtx_pop(opened)


class PerfMetric:
    EMA_FACTOR = 0.1

    def __init__(self):
        self.n, self.total, self.sq, self.ema, self.last = 0, 0.0, 0.0, None, 0.0

    def update(self, ms: float) -> None:
        self.n += 1
        self.total += ms
        self.sq += ms * ms
        self.last = ms
        self.ema = ms if self.ema is None else (1 - self.EMA_FACTOR) * self.ema + self.EMA_FACTOR * ms

    @property
    def mean(self) -> float:
        return self.total / max(self.n, 1)

    @property
    def std(self) -> float:
        m = self.mean
        return max(self.sq / max(self.n, 1) - m * m, 0.0) ** 0.5


class PerfStats:
    MAX_PENDING = 100

    def __init__(self, use_cuda_events: Optional[bool] = None):
        self.use_cuda = torch.cuda.is_available() if use_cuda_events is None else use_cuda_events
        self.host: Dict[str, PerfMetric] = defaultdict(PerfMetric)
        self.device: Dict[str, PerfMetric] = defaultdict(PerfMetric)
        self._pending: List[Tuple[str, torch.cuda.Event, torch.cuda.Event]] = []

    def update_host(self, name: str, ms: float) -> None:
        self.host[name].update(ms)

    def add_events(self, name: str, start, end) -> None:
        self._pending.append((name, start, end))
        if len(self._pending) > self.MAX_PENDING:
            self.resolve(block=False)

    def resolve(self, block: bool = False) -> None:
        still = []
        for name, s, e in self._pending:
            if block:
                e.synchronize()
            if e.query():
                self.device[name].update(s.elapsed_time(e))
            else:
                still.append((name, s, e))
        self._pending = still

    def report_str(self) -> str:
        self.resolve(block=False)
        lines = ["Timer Name           | host avg ms | host ema ms | device avg ms | calls"]
        for name, m in self.host.items():
            d = self.device.get(name)
            lines.append(f"{name:<20} | {m.mean:11.3f} | {m.ema or 0:11.3f} | "
                         f"{(d.mean if d else float('nan')):13.3f} | {m.n}")
        return "\n".join(lines)

    def as_dict(self) -> Dict[str, Dict[str, float]]:
        self.resolve(block=False)
        return {k: {"host_ms": v.mean, "device_ms": self.device[k].mean if k in self.device else None, "calls": v.n}
                for k, v in self.host.items()}


class PerfTimer:
    """`with PerfTimer("forward", perf_stats): ...`; a None `perf_stats` makes it a no-op."""

    def __init__(self, name: str, perf_stats: Optional[PerfStats]):
        self.name, self.stats = name, perf_stats

    def __enter__(self):
        self._nvtx = nvtx_push(self.name)
        if self.stats is None:
            return self
        self.t0 = time.perf_counter()
        if self.stats.use_cuda:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *exc):
        nvtx_pop(self._nvtx)
        if self.stats is None:
            return False
        if self.stats.use_cuda:
            self.e1.record()
            self.stats.add_events(self.name, self.e0, self.e1)
        self.stats.update_host(self.name, (time.perf_counter() - self.t0) * 1e3)
        return False
