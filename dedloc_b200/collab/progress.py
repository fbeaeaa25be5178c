"""Collaboration progress tracking.

Each peer publishes `TrainingState` under `{prefix}_progress[public key]` after every local step; a background
thread re-reads that key and derives `CollaborationState` (global step = max step over non-client... peers that
report, samples accumulated by peers at that step -- extrapolated by their throughput --, ETA of the next
global step and when to look again). `ready_for_step` <=> the collaboration accumulated `target_batch_size`
samples or the ETA passed. Behavioural spec: hivemind 0.9.9 CollaborativeOptimizer, call sites
/root/reference/albert/run_trainer.py:142-168,251-264.
"""
from __future__ import annotations

import contextlib
import logging
import threading
import time
from dataclasses import dataclass
from typing import Dict, Optional

import numpy as np
from pydantic import BaseModel, StrictBool, StrictFloat, confloat, conint

from ..dht import DHT, get_dht_time

logger = logging.getLogger(__name__)


class TrainingState(BaseModel):
    endpoint: Optional[str]
    step: conint(ge=0, strict=True)
    samples_accumulated: conint(ge=0, strict=True)
    samples_per_second: confloat(ge=0.0, strict=True)
    time: StrictFloat
    client_mode: StrictBool
    auxiliary: StrictBool = False
    # typical seconds between two reports of this peer (0 = unknown); readers use it to tell a peer that stopped
    # reporting from one whose micro-steps are simply long
    report_period: confloat(ge=0.0) = 0.0


@dataclass(frozen=True)
class CollaborationState:
    optimizer_step: int
    samples_accumulated: int
    target_batch_size: int
    num_peers: int
    num_clients: int
    eta_next_step: float
    next_fetch_time: float
    num_aux: int = 0
    num_live_peers: int = -1   # peers whose record is fresh (see ProgressTracker.liveness_window); -1 = not computed
    live_keys: tuple = ()      # their record keys, sorted (every peer that reads the same records gets the same tuple)

    @property
    def live_peers(self) -> int:
        return self.num_peers if self.num_live_peers < 0 else self.num_live_peers

    @property
    def ready_for_step(self) -> bool:
        return self.samples_accumulated >= self.target_batch_size or get_dht_time() >= self.eta_next_step

    def ready_within(self, slack_seconds: float) -> bool:
        """`ready_for_step`, or the collaboration will get there within `slack_seconds`. Peers of one NVSwitch box run
        in lock step, which makes "exactly at a micro-step boundary" the COMMON case for the ETA; without a slack half
        of them start one more micro-step and the other half idle in the all-reduce kernel for its whole duration."""
        return self.ready_for_step or get_dht_time() + max(slack_seconds, 0.0) >= self.eta_next_step

    def register_step(self, local_step: int) -> "CollaborationState":
        return CollaborationState(max(local_step, self.optimizer_step), 0, self.target_batch_size, self.num_peers,
                                  self.num_clients, float("inf"), self.next_fetch_time, self.num_aux,
                                  self.num_live_peers, self.live_keys)


class PerformanceEMA:
    """Exponential moving average of samples per second; paused while averaging so that communication time
    does not count against the peer's compute throughput."""

    def __init__(self, alpha: float = 0.1, eps: float = 1e-20):
        self.alpha, self.eps = alpha, eps
        self.ema_seconds_per_sample, self.samples_per_second = 0.0, eps
        self.timestamp = time.perf_counter()
        self.num_updates = 0
        self.paused = False

    def update(self, num_processed: int) -> float:
        assert not self.paused, "PerformanceEMA is paused"
        assert num_processed > 0
        now = time.perf_counter()
        self.num_updates += 1
        seconds_per_sample = max(0.0, now - self.timestamp) / num_processed
        self.timestamp = now
        self.ema_seconds_per_sample = self.alpha * seconds_per_sample + (1 - self.alpha) * self.ema_seconds_per_sample
        adjusted = self.ema_seconds_per_sample / (1 - (1 - self.alpha) ** self.num_updates)
        self.samples_per_second = 1.0 / max(adjusted, self.eps)
        return self.samples_per_second

    def reset_timer(self) -> None:
        self.timestamp = time.perf_counter()

    @contextlib.contextmanager
    def pause(self):
        self.paused, was_paused = True, self.paused
        try:
            yield
        finally:
            self.timestamp = time.perf_counter()
            self.paused = was_paused


class ProgressTracker:
    def __init__(self, dht: DHT, prefix: str, peer_key: bytes, target_batch_size: int, *, endpoint: Optional[str],
                 client_mode: bool = False, auxiliary: bool = False, metadata_expiration: float = 30.0,
                 min_refresh_period: float = 0.5, max_refresh_period: float = 30.0,
                 default_refresh_period: float = 3.0, expected_drift_peers: float = 3.0,
                 expected_drift_rate: float = 0.2, performance_ema_alpha: float = 0.1, start: bool = True,
                 liveness_window: float = 3.0):
        self.dht, self.prefix, self.peer_key = dht, prefix, peer_key
        self.target_batch_size = target_batch_size
        self.endpoint, self.client_mode, self.auxiliary = endpoint, client_mode, auxiliary
        self.metadata_expiration = metadata_expiration
        self.min_refresh_period, self.max_refresh_period = min_refresh_period, max_refresh_period
        self.default_refresh_period = default_refresh_period
        self.expected_drift_peers, self.expected_drift_rate = expected_drift_peers, expected_drift_rate
        self.performance_ema = PerformanceEMA(alpha=performance_ema_alpha)
        self.local_step = 0
        self.local_samples_accumulated = 0
        self.last_reported_time = 0.0
        # A record stays in the DHT for metadata_expiration (30 s) after its author died. A peer counts as LIVE while
        # its last report is younger than max(liveness_window, 4 x its own report period): matchmaking waits only for
        # live peers, and only they are extrapolated in the ETA.
        self.liveness_window = liveness_window
        self.report_period = 0.0
        self._last_change: Dict[bytes, tuple] = {}     # peer key -> (record timestamp, local clock when it changed)
        self.lock = threading.Lock()
        self.state_updated = threading.Event()
        self.should_refresh = threading.Event()
        self.shutdown_flag = threading.Event()
        self.collaboration_state = CollaborationState(0, 0, target_batch_size, 0, 0, float("inf"), 0.0)
        self._thread: Optional[threading.Thread] = None
        if start:
            self.start()

    @property
    def training_progress_key(self) -> str:
        return f"{self.prefix}_progress"

    def start(self) -> None:
        self.collaboration_state = self.fetch_collaboration_state()
        self._thread = threading.Thread(target=self._updater, name="dedloc-progress", daemon=True)
        self._thread.start()

    def shutdown(self) -> None:
        self.shutdown_flag.set()
        self.should_refresh.set()
        if self._thread is not None:
            self._thread.join(timeout=3)
        # withdraw our record so that the collaboration stops counting on us
        try:
            self.dht.store(self.training_progress_key, None, get_dht_time() + self.metadata_expiration,
                           subkey=self.peer_key)
        except Exception:
            pass

    # ---------------------------------------------------------------- reporting
    def report(self, local_step: int, samples_accumulated: int, update_ema_with: int = 0, wait: bool = False):
        with self.lock:
            self.local_step, self.local_samples_accumulated = local_step, samples_accumulated
            if update_ema_with > 0:
                self.performance_ema.update(update_ema_with)
            now = get_dht_time()
            if self.last_reported_time > 0 and update_ema_with > 0:
                # micro-step reports only: the report that follows a round would add the round's length to the period
                gap = now - self.last_reported_time
                self.report_period = gap if self.report_period == 0 else 0.9 * self.report_period + 0.1 * gap
            self.last_reported_time = now
            state = TrainingState(endpoint=self.endpoint, step=local_step, samples_accumulated=samples_accumulated,
                                  samples_per_second=float(self.performance_ema.samples_per_second), time=now,
                                  client_mode=self.client_mode, auxiliary=self.auxiliary,
                                  report_period=float(self.report_period))
        fut = self.dht.store(self.training_progress_key, state.model_dump(), now + self.metadata_expiration,
                             subkey=self.peer_key, return_future=True)
        if wait:
            fut.result()
        return fut

    def locally_ready(self, slack_seconds: float = 0.0) -> bool:
        """`ready_for_step` of the cached collaboration state, corrected by the samples this peer accumulated
        since that state was fetched (so a fast peer does not wait for the next DHT refresh to notice)."""
        st = self.collaboration_state
        if st.ready_within(slack_seconds):
            return True
        own_delta = self.local_samples_accumulated - getattr(self, "_own_samples_in_state", 0)
        return st.optimizer_step <= self.local_step and st.samples_accumulated + max(own_delta, 0) >= st.target_batch_size

    # ---------------------------------------------------------------- fetching
    def _updater(self) -> None:
        while not self.shutdown_flag.is_set():
            delay = max(0.0, self.collaboration_state.next_fetch_time - get_dht_time())
            self.should_refresh.wait(timeout=delay)
            self.should_refresh.clear()
            if self.shutdown_flag.is_set():
                break
            try:
                state = self.fetch_collaboration_state()
                with self.lock:
                    self.collaboration_state = state
                self.state_updated.set()
            except Exception as e:  # keep the thread alive: the DHT may be briefly unreachable
                logger.debug("failed to fetch collaboration state: %r", e)
                time.sleep(self.min_refresh_period)

    def refresh_now(self, timeout: float = 5.0) -> CollaborationState:
        """Synchronously re-read the DHT (H3 step 4: re-fetch right before deciding to step)."""
        state = self.fetch_collaboration_state()
        with self.lock:
            self.collaboration_state = state
        return state

    def fetch_collaboration_state(self) -> CollaborationState:
        response = self.dht.get(self.training_progress_key, latest=True)
        now = get_dht_time()
        if response is None or not isinstance(response.value, dict):
            return CollaborationState(self.local_step, self.local_samples_accumulated, self.target_batch_size,
                                      0, 0, float("inf"), now + self.default_refresh_period)
        valid: Dict[bytes, TrainingState] = {}
        for key, entry in response.value.items():
            if entry.value is None:
                continue
            try:
                valid[key] = TrainingState.model_validate(entry.value)
            except Exception:
                logger.debug("ignoring malformed progress record from %r", key)
        mine = valid.get(self.peer_key)
        self._own_samples_in_state = mine.samples_accumulated if (mine is not None and mine.step >= self.local_step) else 0
        num_peers = len(valid)
        num_clients = sum(s.client_mode for s in valid.values())
        num_aux = sum(s.auxiliary for s in valid.values())
        global_step = self.local_step
        for s in valid.values():
            if not s.client_mode:
                global_step = max(global_step, s.step)
        total_samples, estimated_samples, total_sps, num_live = 0, 0.0, 0.0, 0
        mono = time.monotonic()
        live_keys: list = []
        for key in [k for k in self._last_change if k not in valid]:
            del self._last_change[key]
        for key, s in valid.items():
            # fresh by its own timestamp, or -- for peers whose clock is off by seconds -- seen changing recently
            seen = self._last_change.get(key)
            if seen is None or seen[0] != s.time:
                first_look = seen is None
                seen = self._last_change[key] = (s.time, mono if not first_look else mono - 1e9)
            window = max(self.liveness_window, 4.0 * s.report_period)
            live = key == self.peer_key or now - s.time <= window or mono - seen[1] <= window
            num_live += live
            if live:
                live_keys.append(bytes(key) if isinstance(key, (bytes, bytearray)) else str(key).encode())
            if live:
                total_sps += s.samples_per_second
            if s.step == global_step:
                total_samples += s.samples_accumulated
                # the ETA (not the sample count) extrapolates what a live peer did since it last reported
                estimated_samples += s.samples_accumulated + (max(0.0, now - s.time) * s.samples_per_second if live else 0)
        total_sps = max(total_sps, 1e-9)
        eta_seconds = max(0.0, self.target_batch_size - estimated_samples) / total_sps
        trainers = max(num_peers - num_aux, 1)
        expected_max_peers = max(trainers + self.expected_drift_peers, trainers * (1 + self.expected_drift_rate))
        time_to_next_fetch = float(np.clip(eta_seconds * trainers / expected_max_peers,
                                           self.min_refresh_period, self.max_refresh_period))
        return CollaborationState(global_step, total_samples, self.target_batch_size, num_peers, num_clients,
                                  eta_next_step=now + eta_seconds, next_fetch_time=now + time_to_next_fetch,
                                  num_aux=num_aux, num_live_peers=num_live, live_keys=tuple(sorted(live_keys)))
