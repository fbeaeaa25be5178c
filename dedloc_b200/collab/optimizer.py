"""CollaborativeOptimizer -- collaborative large-batch SGD (the product of the reference).

Every peer accumulates gradients locally and reports progress to the DHT; when the COLLABORATION has
accumulated `target_batch_size` samples the peers form a group (collab/matchmaking.py), run one weighted
butterfly all-reduce of the accumulated gradients and all apply the same optimizer step. Constructor and
attribute surface follow the `hivemind.CollaborativeOptimizer` call sites of the reference
(albert/run_trainer.py:251-264, run_first_peer.py:97-108, run_aux.py:243-258, sgd_collaborative.py:145-157):
`.step() .step_aux() .zero_grad() .load_state_from_peers() .opt .scheduler .prefix .local_step
.local_samples_accumulated .performance_ema .collaboration_state .is_synchronized .param_groups .state_dict()`.

Two data paths:
  * fused (GPU, `opt` is a FusedLamb / FusedLARS): ONE sm_100a kernel reduce-scatters the encoded gradients
    over NVLink peer memory, applies the sample-count weights, runs LAMB/LARS on the owned shard and all-gathers
    the new parameters (parallel/fused_allreduce.py). No NCCL on this path.
  * generic (any torch optimizer; CPU peers): weighted mean of the gradients through the socket butterfly
    (parallel/cpu_backend.py) or the MEAN mode of the kernel, then `opt.step()` locally.
Averaging failures are logged and swallowed: the peer proceeds with its local gradients (H3 step 6).
"""
from __future__ import annotations

import asyncio
import logging
import os
import threading
import time
from typing import Any, Dict, List, Optional

import torch

from ..dht import DHT, get_dht_time
from ..dht.protocol import RPCClient, RPCServer
from ..optim.fused import FusedOptimizerBase
from ..utils.compression import CHUNK, CompressionType
from ..utils.perf_stats import nvtx_range
from .matchmaking import AllreduceException, GroupInfo, Matchmaking
from .progress import CollaborationState, PerformanceEMA, ProgressTracker

logger = logging.getLogger(__name__)

MODE_NODE, MODE_CLIENT, MODE_AUX = 0, 1, 2


def _host_id() -> str:
    """Identifies the machine: CUDA-IPC peer memory only reaches GPUs of the same host (same boot)."""
    global _HOST_ID
    if _HOST_ID is None:
        import socket

        boot = ""
        try:
            with open("/proc/sys/kernel/random/boot_id") as f:
                boot = f.read().strip()
        except OSError:
            pass
        _HOST_ID = f"{socket.gethostname()}/{boot}"
    return _HOST_ID


_HOST_ID: Optional[str] = None


class CollaborativeOptimizer:
    def __init__(self, opt: torch.optim.Optimizer, *, dht: DHT, prefix: str, target_batch_size: int,
                 batch_size_per_step: Optional[int] = None, scheduler=None, min_refresh_period: float = 0.5,
                 max_refresh_period: float = 30, default_refresh_period: float = 3,
                 expected_drift_peers: float = 3, expected_drift_rate: float = 0.2,
                 performance_ema_alpha: float = 0.1, metadata_expiration: float = 30.0,
                 averaging_timeout: Optional[float] = 30.0, averaging_expiration: float = 5.0,
                 target_group_size: int = 256, step_tolerance: int = 1, client_mode: bool = False,
                 auxiliary: bool = False, allow_state_sharing: bool = True, verbose: bool = False,
                 compression_type: Any = CompressionType.FLOAT16, compression: Any = None,
                 throughput: Optional[float] = None, bandwidth: Optional[float] = None,
                 listen_on: str = "127.0.0.1:*", start: bool = True, max_grad_norm: Optional[float] = None,
                 peer_key: Optional[bytes] = None, expected_group_size_from_tracker: bool = True,
                 replicate_state: bool = False, replicate_state_every: int = 8,
                 step_slack_fraction: float = 0.75, liveness_window: float = 3.0, kernel_timeout: Optional[float] = None,
                 emulated_uplink_mbps: Optional[float] = None, authorizer: Any = None, **unused_kwargs):
        if compression is not None:  # sgd_collaborative.py:150 spells the kwarg `compression=`
            compression_type = compression
        if isinstance(compression_type, str):
            compression_type = CompressionType.Value(compression_type)
        self.opt, self.dht, self.prefix, self.scheduler = opt, dht, prefix, scheduler
        self.step_slack_fraction = step_slack_fraction
        self.target_batch_size, self.batch_size_per_step = target_batch_size, batch_size_per_step
        self.averaging_timeout = averaging_timeout if averaging_timeout is not None else 30.0
        # How long the all-reduce kernel waits for a member's data. averaging_timeout (30 s) bounds a round over
        # sockets; members that share NVLink launch within milliseconds of each other once the group is formed, so a
        # member that is still absent after a few seconds was killed between matchmaking and the launch.
        self.kernel_timeout = min(self.averaging_timeout, 5.0) if kernel_timeout is None else float(kernel_timeout)
        # socket rounds only: shape this peer's sends to a link of that speed (the reference's `tc` setup, AWS_runner.ipynb)
        if not emulated_uplink_mbps and os.environ.get("DEDLOC_EMULATED_UPLINK_MBPS"):
            emulated_uplink_mbps = float(os.environ["DEDLOC_EMULATED_UPLINK_MBPS"])
        self.emulated_uplink_mbps = emulated_uplink_mbps
        self.averaging_expiration, self.target_group_size = averaging_expiration, target_group_size
        self.step_tolerance, self.client_mode, self.auxiliary = step_tolerance, client_mode, auxiliary
        self.allow_state_sharing = allow_state_sharing and not auxiliary
        self.verbose, self.compression_type = verbose, CompressionType(compression_type)
        self.throughput = throughput if throughput is not None else bandwidth
        self.max_grad_norm = max_grad_norm
        self.expected_group_size_from_tracker = expected_group_size_from_tracker
        # Optimizer moments live with the member that owns a chunk in a round; the fp32 master parameters are
        # replicated (bit-identical) at all times. `replicate_state=True` pushes the moments to every member in every
        # round (14 instead of 2 bytes per element leave the owner); the default pushes them every
        # `replicate_state_every`-th collaborative step, which bounds what a vanished owner can take with it, and
        # migrates them between owners whenever the group composition changes (_migrate_moments).
        self.replicate_state = replicate_state
        self.replicate_state_every = int(replicate_state_every)
        self.average_diverged_parameters = bool(unused_kwargs.pop("average_diverged_parameters", True))
        self._pending_param_average = False
        self._agreed_step = 0
        self._moment_owners: Optional[List[bytes]] = None   # handles of last round's owners, in owner order; None = all local
        self._needs_resync = False                          # set when a round left this peer's state torn
        self._state_seq = 0                                 # odd while a fused round may be writing this peer's state
        self.unused_kwargs = unused_kwargs
        assert not (client_mode and auxiliary), "an auxiliary peer must accept inbound connections"

        self.local_step = 0
        self.local_samples_accumulated = 0
        self.local_steps_accumulated = 0
        self.last_step_time: Optional[float] = None
        self.state_lock = threading.RLock()
        # True while this peer sits inside a global step BEFORE any parameter update (it holds state_lock there, but
        # its parameters are stable): state downloads are served without the lock in that window (lineage healing)
        self._params_parked = False
        self.round_counter = 0
        # Parameter lineage: a hash chain over (who averaged with whom, at which step). Peers with equal lineage hold
        # bit-identical parameters (every update is deterministic given the group), so a group whose members disagree
        # -- somebody stepped alone after a failed round -- is healed BEFORE the reduction: the minority copies the
        # state of a majority member. (The reference heals the same situation by averaging parameters every step.)
        self.lineage = "0" * 16
        self.lineage_depth = 0          # optimizer steps behind this lineage: ties go to the longer training history
        self.stats = {"rounds": 0, "failed_rounds": 0, "solo_rounds": 0, "state_downloads": 0, "lineage_resyncs": 0,
                      "matchmaking_s": 0.0, "averaging_host_s": 0.0, "moment_migrations": 0, "torn_rounds": 0}
        self.is_fused = isinstance(opt, FusedOptimizerBase)
        self._params = [p for g in opt.param_groups for p in g["params"]]
        self.on_gpu = bool(self._params) and self._params[0].device.type == "cuda"
        self._cpu_acc: Optional[List[torch.Tensor]] = None
        self._server: Optional[RPCServer] = None
        self._client = RPCClient()
        # With an authorizer (collab/auth.py) the peer-to-peer RPCs of this optimizer -- group formation, socket
        # all-reduce parts and, most of all, state download -- travel in signed envelopes carrying an access token;
        # a peer without a valid token can neither join a group nor read anybody's parameters.
        self.authorizer = authorizer
        if authorizer is not None:
            from .auth import AuthenticatedClient

            self._client = AuthenticatedClient(self._client, authorizer)
        self._matchmaking: Optional[Matchmaking] = None
        self._cpu_averager = None
        self.endpoint: Optional[str] = None
        self.peer_key = peer_key
        self.tracker: Optional[ProgressTracker] = None
        self._tracker_kwargs = dict(
            metadata_expiration=metadata_expiration, min_refresh_period=min_refresh_period,
            max_refresh_period=max_refresh_period, default_refresh_period=default_refresh_period,
            expected_drift_peers=expected_drift_peers, expected_drift_rate=expected_drift_rate,
            performance_ema_alpha=performance_ema_alpha, liveness_window=liveness_window)
        self.metadata_expiration = metadata_expiration
        # mapped buffers of peers nobody met for two record lifetimes are unmapped (parallel/symm.py MappingLeases)
        from ..parallel.symm import MappingLeases

        self._mapping_leases = MappingLeases(ttl=2.0 * metadata_expiration)
        self.listen_on = listen_on
        self._started = False
        if start:
            self.start()

    # ------------------------------------------------------------------------------ lifecycle
    def start(self) -> None:
        if self._started:
            return
        if self.is_fused and self.on_gpu:
            self.opt.materialize()
        loop = self.dht.loop

        async def _start_server():
            if self.client_mode:
                return None, None
            server = RPCServer({})
            host, _, port = self.listen_on.rpartition(":")
            bind = host.strip("[]")
            if bind in ("", "*", "::"):
                bind = "0.0.0.0"
            p = await server.start(bind, 0 if port in ("", "*") else int(port))
            return server, p

        self._server, port = asyncio.run_coroutine_threadsafe(_start_server(), loop).result()
        visible_host = self.dht.get_visible_address().rpartition(":")[0]
        import os
        import uuid
        self.endpoint = f"{visible_host}:{port}" if port else f"client-{os.getpid()}-{uuid.uuid4().hex[:8]}"
        if self.peer_key is None:
            self.peer_key = self.endpoint.encode()
        from ..parallel.cpu_backend import CPUAverager

        self._matchmaking = Matchmaking(self.dht, self._server, self._client, endpoint=self.endpoint,
                                        prefix=self.prefix, target_group_size=self.target_group_size,
                                        averaging_expiration=self.averaging_expiration,
                                        client_mode=self.client_mode)
        self._cpu_averager = CPUAverager(self._server, self._client, self.compression_type,
                                         timeout=self.averaging_timeout, uplink_mbps=self.emulated_uplink_mbps)
        if self._server is not None:
            self._server.handlers["download_state"] = self._rpc_download_state
            if self.authorizer is not None:
                from .auth import protect_handlers

                protect_handlers(self._server, self.authorizer, ("download_state", "join_group", "allreduce_part"))
        self.tracker = ProgressTracker(self.dht, self.prefix, self.peer_key, self.target_batch_size,
                                       endpoint=self.endpoint, client_mode=self.client_mode,
                                       auxiliary=self.auxiliary, **self._tracker_kwargs)
        self._started = True
        self._publish_state_sharing()
        # A peer that never trains (the coordinator builds an optimizer without a batch size, only to pull state) must
        # not show up in the progress records: the trainers would count it as a peer and wait for it in matchmaking.
        if self._reports_progress:
            self.tracker.report(self.local_step, 0)

    @property
    def _reports_progress(self) -> bool:
        return self.batch_size_per_step is not None or self.auxiliary

    def shutdown(self) -> None:
        if not self._started:
            return
        self._started = False
        try:
            self.tracker.shutdown()
        except Exception:
            pass

        async def _stop():
            if self._server is not None:
                await self._server.shutdown()
            await self._client.shutdown()

        try:
            asyncio.run_coroutine_threadsafe(_stop(), self.dht.loop).result(timeout=5)
        except Exception:
            pass

    def __del__(self):
        try:
            self.shutdown()
        except Exception:
            pass

    # ------------------------------------------------------------------------------ torch.optim duck typing
    @property
    def param_groups(self):
        return self.opt.param_groups

    @property
    def state(self):
        return self.opt.state

    @property
    def defaults(self):
        return self.opt.defaults

    def add_param_group(self, group):
        raise ValueError("CollaborativeOptimizer does not support adding parameter groups after creation")

    def state_dict(self) -> dict:
        sd = self.opt.state_dict()
        sd["state"] = dict(sd["state"])
        sd["state"]["local_step"] = self.local_step
        return sd

    def load_state_dict(self, sd: dict) -> None:
        sd = dict(sd)
        state = dict(sd.get("state", {}))
        if "local_step" in state:
            self.local_step = int(state.pop("local_step"))
        sd["state"] = state
        self.opt.load_state_dict(sd)
        self._sync_scheduler()

    def zero_grad(self, *args, **kwargs):
        if self.is_fused and self.on_gpu:
            return self.opt.zero_grad()
        kwargs.setdefault("set_to_none", False)
        return self.opt.zero_grad(*args, **kwargs)

    # ------------------------------------------------------------------------------ reference attributes
    @property
    def performance_ema(self) -> PerformanceEMA:
        return self.tracker.performance_ema

    @property
    def collaboration_state(self) -> CollaborationState:
        return self.tracker.collaboration_state

    @property
    def is_synchronized(self) -> bool:
        if self._needs_resync:
            return False
        return self.local_step >= self.collaboration_state.optimizer_step - self.step_tolerance

    def is_alive(self) -> bool:
        return self._started and self.dht.is_alive()

    # ------------------------------------------------------------------------------ the step
    def step(self, batch_size: Optional[int] = None, **kwargs) -> Optional[bool]:
        """Accumulate the current gradients; run a collaborative optimizer step when the collaboration is ready.
        Returns True when a global step happened."""
        if self.auxiliary:
            raise RuntimeError("auxiliary peers call step_aux()")
        if self.batch_size_per_step is None:
            if batch_size is None:
                raise ValueError("pass batch_size to step() or batch_size_per_step to the constructor")
            self.batch_size_per_step = batch_size
        batch_size = batch_size if batch_size is not None else self.batch_size_per_step

        if not self.is_synchronized:
            logger.warning("%s %s (local step %d, global %d): downloading state", self.endpoint,
                           "must resynchronise after a torn round" if self._needs_resync else "is behind the collaboration",
                           self.local_step, self.collaboration_state.optimizer_step)
            self.load_state_from_peers()
            self.zero_grad()
            return False
        if self.last_step_time is not None and get_dht_time() - self.last_step_time > self.metadata_expiration:
            logger.warning("training step took longer than metadata_expiration; peers may not see this one")

        with self.state_lock, nvtx_range("accumulate_gradients"):
            self._accumulate_local_gradients()
            self.local_samples_accumulated += batch_size
            self.local_steps_accumulated += 1
        report = self.tracker.report(self.local_step, self.local_samples_accumulated, update_ema_with=batch_size)
        self.last_step_time = get_dht_time()
        self._maybe_republish_state_sharing()

        # do not start a micro-step the collaboration will not wait for: ready if the ETA is closer than a fraction of it
        slack = self.step_slack_fraction * batch_size / max(self.performance_ema.samples_per_second, 1e-9)
        slack = min(slack, 0.5 * self.averaging_expiration)
        if not self.tracker.locally_ready(slack):
            return False
        try:
            report.result(timeout=5.0)  # our own latest sample count must be visible to the re-fetch below
        except Exception:
            pass
        state = self.tracker.refresh_now()  # H3 step 4: somebody may already have stepped
        if state.optimizer_step > self.local_step:
            # the collaboration moved on without us: our accumulated samples belong to a finished step
            if not self.is_synchronized:
                return False
        elif not state.ready_within(slack):
            return False
        self._global_step(state)
        return True

    def step_aux(self, **kwargs) -> bool:
        """Auxiliary peers donate bandwidth: they join the averaging round with weight 0 (run_aux.py:260-262)."""
        if not self.auxiliary:
            raise RuntimeError("step_aux() is for auxiliary peers")
        if not self.is_synchronized and self.is_fused and self.on_gpu:
            self.load_state_from_peers()
            if self._needs_resync:
                return False          # an auxiliary peer with a torn state must not own a shard
        # heartbeat: an auxiliary peer has no micro-steps to report, but group leaders wait only for peers whose
        # record is fresh (progress.py liveness_window)
        if self._reports_progress and get_dht_time() - self.tracker.last_reported_time > min(1.0, self.tracker.liveness_window / 3):
            self.tracker.report(self.local_step, 0)
        state = self.tracker.refresh_now()
        if not state.ready_for_step or state.num_peers - state.num_aux < 1:
            return False
        self.local_step = max(self.local_step, state.optimizer_step)
        self._global_step(state)
        return True

    # ------------------------------------------------------------------------------ gradient accumulation
    @torch.no_grad()
    def _accumulate_local_gradients(self) -> None:
        if self.is_fused and self.on_gpu:
            opt = self.opt
            views = opt.layout.views(opt.grad_flat)
            for p, v in zip(opt.flat_params, views):  # somebody may have reset .grad (set_to_none / new tensor)
                if p.grad is None:
                    p.grad = v
                elif p.grad.data_ptr() != v.data_ptr():
                    v.add_(p.grad)
                    p.grad = v
            opt.fold_gradients(self.max_grad_norm)
            return
        if self._cpu_acc is None:
            self._cpu_acc = [torch.zeros_like(p, dtype=torch.float32) for p in self._params]
        grads = [p.grad if p.grad is not None else torch.zeros_like(p) for p in self._params]
        coef = 1.0
        if self.max_grad_norm is not None and self.max_grad_norm > 0:
            total = torch.sqrt(sum((g.float() ** 2).sum() for g in grads))
            coef = float(min(1.0, self.max_grad_norm / (float(total) + 1e-6)))
        for a, g in zip(self._cpu_acc, grads):
            a.add_(g.float(), alpha=coef)

    # ------------------------------------------------------------------------------ global step
    def _my_blob(self) -> dict:
        mode = MODE_AUX if self.auxiliary else (MODE_CLIENT if self.client_mode else MODE_NODE)
        blob = {"weight": 0.0 if self.auxiliary else float(self.local_samples_accumulated), "mode": mode,
                "throughput": None if self.client_mode else float(self.throughput or 1.0),
                "step": self.local_step, "handle": None, "total": None, "device": None, "lineage": self.lineage,
                "lineage_depth": self.lineage_depth, "host": _host_id(),
                "n_avg_tensors": 0 if (self.is_fused and self.on_gpu) else len(self._averaged_tensors())}
        if self.is_fused and self.on_gpu:
            handle = self.opt.buffers.handle
            # an all-zero handle means cudaIpcGetMemHandle failed on this peer: it cannot be mapped by anybody
            blob.update(handle=handle if any(handle) else None, total=self.opt.layout.total,
                        device=self.opt.buffers.device)
        return blob

    def _global_step(self, state: CollaborationState) -> None:
        t0 = time.perf_counter()
        with self.performance_ema.pause(), self.state_lock:
            self._params_parked = True
            group: Optional[GroupInfo] = None
            alone = self.expected_group_size_from_tracker and 0 <= state.num_live_peers <= 1
            if state.num_peers > 1 and not alone:  # num_peers counts this peer's own progress record too
                # more live peers than target_group_size: balanced, per-round reshuffled buckets from the records that
                # everybody reads (falls back to the bits derived from the last group when we are not in them yet)
                self._balanced_bits = (state.live_keys and self._matchmaking.assign_group_bits(
                    state.live_keys, bytes(self.peer_key), self.local_step) is not None)
                # wait for the peers that are still reporting, not for the records dead peers left behind
                # ... and, when the collaboration is split into buckets, for this bucket's share of them
                buckets = 1 << len(self._matchmaking.group_bits)
                share = lambda live: max(-(-live // buckets), 2)  # noqa: E731
                expected = share(state.live_peers) if self.expected_group_size_from_tracker else None
                # the tracker thread keeps re-reading the records while we wait for the group
                # (a failed read shows up as "no records": keep the earlier expectation then)
                def live_now():
                    live = self.tracker.collaboration_state.live_peers
                    return share(live) if live >= 1 else None

                if not self.expected_group_size_from_tracker:
                    live_now = None
                # the others finished this step while we were looking for them: nobody is left to form a group with
                # ... or every other record belongs to a peer that stopped reporting (the last survivor of a fleet
                # would otherwise wait averaging_expiration in every round until those records expire)
                def left_behind():
                    cs = self.tracker.collaboration_state
                    return cs.optimizer_step > self.local_step or (self.expected_group_size_from_tracker
                                                                   and 0 <= cs.num_live_peers <= 1)
                try:
                    with nvtx_range("matchmaking"):
                        group = self.dht.run_coroutine(
                            lambda dht, node: self._matchmaking.look_for_group(
                                self._my_blob(), timeout=self.averaging_timeout, expected_size=expected,
                                expected_size_fn=live_now, give_up_fn=left_behind))
                except Exception as e:
                    logger.warning("matchmaking failed: %r", e)
            t1 = time.perf_counter()
            self.stats["matchmaking_s"] += t1 - t0
            self.stats["rounds"] += 1
            self._last_round = {"expected": state.live_peers, "records": state.num_peers,
                                "group": group.group_size if group is not None else 0, "matchmaking_s": t1 - t0}
            # next round's bucket: members of one group scatter, so large collaborations reshuffle every round
            # (counted over live peers: the record of a dead peer must not push the collaboration over target_group_size)
            if not getattr(self, "_balanced_bits", False):
                self._matchmaking.rederive_group_bits(group, state.live_peers)
            averaged = False
            if group is not None and group.group_size > 1:
                try:
                    with nvtx_range("averaging_round"):
                        self._run_group_round(group)
                    averaged = True
                except AllreduceException as e:
                    self.stats["failed_rounds"] += 1
                    logger.warning("Skipped averaging: averaging round failed with %r", e)
                except Exception as e:
                    # hivemind logs and swallows ANY failure of the averaging step (H3 step 6): a round must never take
                    # the trainer down. What the others did in the meantime shows up as a lineage difference in the
                    # next group and is healed there.
                    self.stats["failed_rounds"] += 1
                    logger.exception("Skipped averaging: averaging round failed with an unexpected %r", e)
            if self._needs_resync:
                # the round died after it had begun to change this peer's state: neither a local step nor a progress
                # report makes sense on torn parameters. step() downloads a consistent state before anything else.
                self.stats["torn_rounds"] += 1
                self._params_parked = False
                self.stats["averaging_host_s"] += time.perf_counter() - t1
                return
            if not averaged and not self.auxiliary:
                self.stats["solo_rounds"] += 1
                self._run_solo_round()
                self.lineage = self._mix(self.lineage, "S", self.local_step, self.endpoint)
                self.lineage_depth += 1
            elif averaged:
                self.lineage = self._round_lineage
                self.lineage_depth = self._agreed_depth + 1
            self._last_round["averaging_s"] = time.perf_counter() - t1
            self._finish_global_step()
            self._close_stale_mappings()
            self.stats["averaging_host_s"] += time.perf_counter() - t1

    def _close_stale_mappings(self) -> None:
        if not (self.is_fused and self.on_gpu):
            return
        keep = list(self._moment_owners or ())          # their moment chunks may still have to be pulled
        for handle in self._mapping_leases.expired(time.monotonic(), keep=keep):
            if handle != self.opt.buffers.handle:
                try:
                    self.opt.buffers.close_peer(handle)
                    self.stats["mappings_closed"] = self.stats.get("mappings_closed", 0) + 1
                except Exception as e:   # the peer's process is gone: nothing to salvage
                    logger.debug("closing a stale peer mapping failed: %r", e)

    def _finish_global_step(self) -> None:
        self._params_parked = False
        if self.is_fused and self.on_gpu:
            self.opt.buffers.acc.zero_()
        elif self._cpu_acc is not None:
            for a in self._cpu_acc:
                a.zero_()
        self.local_step += 1
        self.local_samples_accumulated = 0
        self.local_steps_accumulated = 0
        self._sync_scheduler()
        self.tracker.collaboration_state = self.tracker.collaboration_state.register_step(self.local_step)
        self.tracker.report(self.local_step, 0)
        self.tracker.should_refresh.set()
        self._publish_state_sharing()
        if self.verbose:
            r = getattr(self, "_last_round", None) or {}
            logger.info("%s: collaborative step %d done (group of %d, %d live of %d records; matchmaking %.2f s, "
                        "averaging %.2f s)", self.endpoint, self.local_step, r.get("group", 0), r.get("expected", 0),
                        r.get("records", 0), r.get("matchmaking_s", 0.0), r.get("averaging_s", 0.0))

    def _sync_scheduler(self) -> None:
        if self.scheduler is not None:
            # the update itself ran inside the fused all-reduce kernel (or the generic step above), not through
            # torch's patched Optimizer.step: tell the scheduler's order check that the optimizer did step
            sched_opt = getattr(self.scheduler, "optimizer", None)
            if sched_opt is not None:
                sched_opt._opt_called = True
            while getattr(self.scheduler, "_step_count", 0) <= self.local_step:
                self.scheduler.step()

    # ---- fused (GPU) rounds --------------------------------------------------------------------------
    def _next_tag(self, group_id: Optional[bytes] = None) -> int:
        self.round_counter += 1
        if group_id is not None:
            tag = int.from_bytes(group_id[:4], "little") & 0x3FFFFFFF
        else:
            tag = (0x20000000 | (self.round_counter & 0x1FFFFFFF))
        return tag or 1

    def _encode_local(self, codec: CompressionType, in_kernel: bool = False) -> float:
        """Encode the accumulated gradient (mean over the local micro-steps) into the send buffer; with `in_kernel`
        the push-based kernel does that itself from the accumulators. Returns the scale."""
        from ..ops import flat_ops

        buf = self.opt.buffers
        # device-side bracket of the averaging work (encode + fused kernel): the "exposed averaging time"
        self._avg_ev0 = torch.cuda.Event(enable_timing=True)
        self._avg_ev1 = torch.cuda.Event(enable_timing=True)
        self._avg_ev0.record(torch.cuda.current_stream(buf.device))
        scale = 1.0 / max(self.local_steps_accumulated, 1)
        if not in_kernel:
            flat_ops.encode(codec, buf.acc, scale, buf.send, buf.aux)
        return scale

    def _launch_fused(self, group_spec, codec, replicate: bool = True, push_kernel: bool = False,
                      scale: float = 1.0) -> dict:
        # GPU peers serve state downloads straight from the mapped buffer; `_state_seq` is odd from before the launch
        # until the kernel has completed, so a copy that overlaps a round is rejected by the requester (seqlock)
        from ..parallel.fused_allreduce import fused_allreduce, fused_allreduce_v2

        opt = self.opt
        opt.opt_step += 1
        hyper = opt.hyper()
        hyper.replicate_state = replicate
        self._state_seq += 1
        try:
            if push_kernel:
                fused_allreduce_v2(opt.buffers, group_spec, codec, hyper, src_f32=True, scale=scale)
            else:
                fused_allreduce(opt.buffers, group_spec, codec, hyper, xgpu_barriers=len(group_spec.bases) > 1)
            self._avg_ev1.record(torch.cuda.current_stream(opt.buffers.device))
            status = opt.buffers.read_status()
        finally:
            self._state_seq += 1
        self.stats["averaging_device_ms"] = self.stats.get("averaging_device_ms", 0.0) + self._avg_ev0.elapsed_time(self._avg_ev1)
        self.stats["fused_rounds"] = self.stats.get("fused_rounds", 0) + 1
        if status["code"] in (1, 2):
            opt.opt_step -= 1
        return status

    def _run_solo_round(self) -> None:
        if self.is_fused and self.on_gpu:
            from ..parallel.fused_allreduce import GroupSpec

            buf = self.opt.buffers
            self._encode_local(CompressionType.NONE)
            spec = GroupSpec(bases=[buf.symm.ptr], me=0, weights=[1.0], shards=[(0, self.opt.layout.n_chunks)],
                             slots=[0], tag=self._next_tag())
            status = self._launch_fused(spec, CompressionType.NONE)
            self._moment_owners = None      # this peer updated every chunk itself: all of its moments are current
            if status["code"] == 4:
                logger.warning("non-finite gradients: optimizer update skipped at step %d", self.local_step)
            return
        steps = max(self.local_steps_accumulated, 1)
        with torch.no_grad():
            for p, a in zip(self._params, self._cpu_acc or []):
                p.grad = (a / steps).to(p.dtype)
        self._generic_opt_step()

    def _generic_opt_step(self) -> None:
        self._params_parked = False
        finite = all(torch.isfinite(p.grad).all() for p in self._params if p.grad is not None)
        if finite:
            self.opt.step()
        else:
            logger.warning("non-finite gradients: optimizer update skipped at step %d", self.local_step)

    @staticmethod
    def _mix(prev: str, *parts) -> str:
        import hashlib

        h = hashlib.blake2b(prev.encode(), digest_size=8)
        for part in parts:
            h.update(b"|" + str(part).encode())
        return h.hexdigest()

    def _heal_lineage(self, group: GroupInfo, blobs: List[dict]) -> str:
        """Agree on ONE lineage for the group. Returns it; sets `_pending_param_average` when the members have to
        average their parameters first.

        * Members that trained in different groups of the previous round (target_group_size < number of peers: the
          reshuffle of contributor_notebook.ipynb cell 3) or stepped alone after a failed round are at the SAME step
          with different lineages: nobody's step is thrown away -- the group averages parameters (and, off the GPU,
          optimizer statistics) before it averages gradients, which is what hivemind's TrainingAverager does in
          every round.
        * Members at different steps: the ones behind copy the state of a member of the most advanced lineage
          (most members, then longest history). Donors are members of this very round: nothing in it modifies their
          state before every member has launched, so their parameters do not change while we copy."""
        lins = [str(b.get("lineage", "")) for b in blobs]
        depths = [int(b.get("lineage_depth", 0) or 0) for b in blobs]
        steps = [int(b.get("step") or 0) for b in blobs]
        # auxiliary peers take whatever the trainers agree on
        voters = [i for i, b in enumerate(blobs) if b.get("mode") != MODE_AUX] or list(range(len(blobs)))
        distinct = sorted({lins[i] for i in voters})
        self._pending_param_average = False
        vsteps = [steps[i] for i in voters]
        if len(distinct) > 1 and self.average_diverged_parameters and max(vsteps) - min(vsteps) <= self.step_tolerance:
            # (a group that finished its previous round a moment earlier may already be one step ahead)
            self._pending_param_average = True
            self._agreed_depth = max(depths[i] for i in voters)
            self._agreed_step = max(vsteps)
            return self._mix("A", *distinct)
        votes: Dict[str, List[str]] = {}
        depth: Dict[str, int] = {}
        top = max(steps[i] for i in voters)
        for i in voters:
            if steps[i] < top and len(distinct) > 1:
                continue                                   # a member that is behind does not get to vote
            votes.setdefault(lins[i], []).append(group.endpoints[i])
            depth[lins[i]] = max(depth.get(lins[i], 0), depths[i])
        # most members; then the longest training history (a freshly started peer never wins a tie); then endpoint
        agreed = min(votes, key=lambda l: (-len(votes[l]), -depth[l], min(votes[l])))
        self._agreed_depth = depth[agreed]
        # auxiliary GPU peers own a share in optimizer rounds, so a stale one would push a stale shard to everybody:
        # they are healed like everyone else; a CPU auxiliary peer only relays and keeps no parameters worth healing
        heal_me = not self.auxiliary or (self.is_fused and self.on_gpu)
        if self.lineage != agreed and heal_me and (len(distinct) > 1 or self.auxiliary):
            donor = min(votes[agreed])
            logger.warning("%s: parameters diverged from the group (lineage %s vs %s): copying the state of %s",
                           self.endpoint, self.lineage, agreed, donor)
            ok = False
            try:
                ok = self._download_from(donor, bool(self.is_fused and self.on_gpu), in_round=True)
            except Exception as e:
                logger.warning("lineage resync from %s failed: %r", donor, e)
            if not ok:
                raise AllreduceException("could not resynchronise parameters with the group")
            self.stats["lineage_resyncs"] += 1
        return agreed

    @staticmethod
    def _check_blobs(blobs: List[Any]) -> None:
        """Every member's blob came over the network through the group leader. One malformed blob must cost a failed
        round (the peer steps on its own gradients), not an exception out of step()."""
        import math

        for i, b in enumerate(blobs):
            ok = isinstance(b, dict) and b.get("mode") in (MODE_NODE, MODE_CLIENT, MODE_AUX)
            if ok:
                w, tp, h = b.get("weight"), b.get("throughput"), b.get("handle")
                ok = (isinstance(w, (int, float)) and not isinstance(w, bool) and math.isfinite(w) and w >= 0
                      and (tp is None or (isinstance(tp, (int, float)) and not isinstance(tp, bool)
                                          and math.isfinite(tp) and tp >= 0))
                      and isinstance(b.get("step"), int) and b["step"] >= 0
                      and (h is None or (isinstance(h, (bytes, bytearray)) and len(h) <= 256))
                      and (b.get("total") is None or (isinstance(b["total"], int) and b["total"] > 0))
                      and isinstance(b.get("lineage", ""), str) and len(b.get("lineage", "")) <= 128
                      and isinstance(b.get("lineage_depth", 0) or 0, int)
                      and isinstance(b.get("n_avg_tensors", 0) or 0, int))
            if not ok:
                raise AllreduceException(f"member {i} of the group sent a malformed description of itself")

    def _run_group_round(self, group: GroupInfo) -> None:
        blobs = list(group.gathered)
        self._check_blobs(blobs)
        agreed = self._heal_lineage(group, blobs)
        self._round_lineage = self._mix(agreed, "G", blobs[0].get("step"), *sorted(group.endpoints))
        from .. import lib as native

        handles = [b.get("handle") for b in blobs]
        all_gpu = (all(h and any(bytes(h)) for h in handles) and len({b.get("total") for b in blobs}) == 1
                   and len({b.get("host") for b in blobs}) == 1      # CUDA IPC reaches GPUs of one host only
                   and len(blobs) <= native.MAX_PEERS)
        bases = None
        if self.is_fused and self.on_gpu and all_gpu:
            try:
                bases = [self.opt.buffers.peer_base(bytes(h)) for h in handles]
                self._mapping_leases.touch(handles, time.monotonic())
            except Exception as e:
                logger.warning("cannot map peer memory (%r): falling back to the socket butterfly", e)
                bases = None
        if self._pending_param_average:
            self._average_parameters(group, blobs, bases)
            self.lineage, self.lineage_depth = agreed, self._agreed_depth
            if self._agreed_step > self.local_step and not self.auxiliary:
                self.local_step = self._agreed_step            # merged with a group that was one step ahead
                self._sync_scheduler()
            self.stats["parameter_averagings"] = self.stats.get("parameter_averagings", 0) + 1
        if bases is not None:
            return self._run_fused_group_round(group, blobs, bases)
        return self._run_socket_group_round(group, blobs)

    # ---- parameter averaging (H4): merges the work of groups that stepped side by side -----------------------------
    def _averaged_tensors(self) -> List[torch.Tensor]:
        """Parameters, then every floating-point optimizer statistic, in a fixed order (generic / CPU path)."""
        out = [p.data for p in self._params]
        for p in self._params:
            st = self.opt.state.get(p, {}) if hasattr(self.opt, "state") else {}
            for key in sorted(st):
                v = st[key]
                if isinstance(v, torch.Tensor) and v.is_floating_point() and v.shape == p.shape:
                    out.append(v)
        return out

    def _average_parameters(self, group: GroupInfo, blobs: List[dict], bases: Optional[List[int]]) -> None:
        weights = [0.0 if b.get("mode") == MODE_AUX else 1.0 for b in blobs]
        if bases is not None:
            from ..parallel.fused_allreduce import GroupSpec, fused_average_parameters, v2_supported

            opt, buf = self.opt, self.opt.buffers
            shares = [0.0 if b["mode"] == MODE_CLIENT else 1.0 for b in blobs]
            base_tag = (int.from_bytes(group.group_id[4:8], "little") & 0x3FFFFFF0) or 16
            spec = GroupSpec(bases=bases, me=group.my_index, weights=weights, shards=opt.layout.shard_ranges(shares),
                             slots=list(range(len(blobs))), tag=base_tag, timeout_s=self.kernel_timeout)
            if not v2_supported(opt.layout, buf.regions, spec, CompressionType.NONE):
                raise AllreduceException("parameter averaging needs the push-based kernel")
            self._state_seq += 1
            try:
                fused_average_parameters(buf, spec)
                status = buf.read_status()
            finally:
                self._state_seq += 1
            if status["code"] != 0:
                if status["code"] == 3:
                    self._needs_resync = True
                raise AllreduceException(f"parameter averaging aborted: {status['name']}")
            return
        from ..parallel.flat import FlatLayout

        if self.is_fused and self.on_gpu:
            raise AllreduceException("parameter averaging between GPU and CPU peers is not supported")
        tensors = self._averaged_tensors()
        counts = {int(b.get("n_avg_tensors") or 0) for b in blobs if b.get("mode") != MODE_AUX}
        if len(counts) > 1:
            tensors = [p.data for p in self._params]        # somebody has no optimizer statistics yet: parameters only
        layout = FlatLayout([t.shape for t in tensors])
        flat = layout.pack([t.float() for t in tensors], dtype=torch.float32, device="cpu")
        throughputs = [b["throughput"] if b["mode"] != MODE_CLIENT else None for b in blobs]
        try:
            averaged = self.dht.run_coroutine(
                lambda dht, node: self._cpu_averager.allreduce(group, flat, weights, throughputs, phase="params"))
        except AllreduceException:
            raise
        except Exception as e:
            raise AllreduceException(repr(e))
        if self.auxiliary:
            return
        with torch.no_grad():
            for t, src in zip(tensors, layout.views(averaged)):
                t.copy_(src.to(t.dtype))

    def _run_fused_group_round(self, group: GroupInfo, blobs: List[dict], bases: List[int]) -> None:
        import os

        from ..parallel.fused_allreduce import _FLY_CODECS, GroupSpec, v2_supported

        opt, buf = self.opt, self.opt.buffers
        if blobs[0]["total"] != opt.layout.total:
            raise AllreduceException("peers disagree on the flat layout size")
        shares = [0.0 if b["mode"] == MODE_CLIENT else 1.0 for b in blobs]
        weights = [float(b["weight"]) for b in blobs]
        if sum(weights) <= 0:
            raise AllreduceException("group has no gradient contributors")
        if sum(shares) <= 0:
            raise AllreduceException("group has no member that can own a shard")
        handles = [bytes(b["handle"]) for b in blobs]
        owners = [h for h, s in zip(handles, shares) if s > 0]
        spec = GroupSpec(bases=bases, me=group.my_index, weights=weights,
                         shards=opt.layout.shard_ranges(shares), slots=list(range(len(blobs))),
                         tag=self._next_tag(group.group_id), timeout_s=self.kernel_timeout)
        push = (os.environ.get("DEDLOC_AR_KERNEL", "push") != "pull"
                and v2_supported(opt.layout, buf.regions, spec, self.compression_type))
        # every member derives the same answer from the agreed step, so the whole group replicates in the same rounds
        step = int(blobs[0].get("step") or 0)
        replicate = self.replicate_state or (self.replicate_state_every > 0 and step % self.replicate_state_every == 0)
        if not push:
            replicate = True      # the pull kernel's contiguous shards follow the group size: keep every replica whole
        self._migrate_moments(handles, bases, None if replicate else owners)
        scale = self._encode_local(self.compression_type, in_kernel=push and self.compression_type in _FLY_CODECS)
        status = self._launch_fused(spec, self.compression_type, replicate=replicate, push_kernel=push, scale=scale)
        code = status["code"]
        if code in (0, 4):
            self._moment_owners = None if replicate else owners
            if code == 4:
                logger.warning("non-finite averaged gradients: the update of %d tensor(s) was skipped by the whole "
                               "group", status.get("nonfinite_chunks", 0))
            return
        if code == 3:
            # the round died between the first and the last update of this peer's parameters
            self._needs_resync = True
            raise AllreduceException("peer vanished while updates were in flight; state will be re-downloaded")
        raise AllreduceException(f"fused all-reduce aborted: {status['name']}")

    def _migrate_moments(self, handles: List[bytes], bases: List[int], new_owners: Optional[List[bytes]] = None) -> None:
        """Moments are current only on the member that owned a chunk in the last round (interleaved: chunk c belongs to
        owner c mod K). Before a round whose owners differ, pull every previous owner's chunks out of its mapped buffer,
        as far as that owner is a member of the new group. Nobody writes moments in the new round before every owner
        has launched its kernel (UPDATE waits for every owner's norm words), and every owner launches after its own
        pulls, so the copies read a quiescent state."""
        prev = self._moment_owners
        buf = self.opt.buffers
        if prev is None or not buf.regions.m or (new_owners is not None and new_owners == prev):
            return        # everything is local, or the same owners in the same order keep their chunks
        from .. import lib as native
        from ..parallel.flat import CHUNK

        mine = buf.handle
        r, K, C = buf.regions, len(prev), buf.layout.n_chunks
        stream = torch.cuda.current_stream(buf.device)
        pulled = 0
        with torch.cuda.device(buf.device):
            for k, h in enumerate(prev):
                if h == mine or h not in handles:
                    continue            # my own chunks are current; an absent owner's chunks stay as stale as they are
                base = bases[handles.index(h)]
                rows = (C - k + K - 1) // K
                if rows <= 0:
                    continue
                for off in (r.m, r.v):
                    native.check(native.lib().dedloc_memcpy2d_async(
                        buf.symm.ptr + off + k * CHUNK * 4, K * CHUNK * 4, base + off + k * CHUNK * 4, K * CHUNK * 4,
                        CHUNK * 4, rows, stream.cuda_stream), "moment migration")
                pulled += 1
        if pulled:
            self.stats["moment_migrations"] += 1

    def _run_socket_group_round(self, group: GroupInfo, blobs: List[dict]) -> None:
        """Socket butterfly (CPU peers, or a mixed group): average the gradients, then step locally."""
        from ..parallel.flat import FlatLayout

        steps = max(self.local_steps_accumulated, 1)
        if self.is_fused and self.on_gpu:
            layout = self.opt.layout
            flat = (self.opt.buffers.acc / steps).cpu()
        else:
            layout = FlatLayout([p.shape for p in self._params])
            acc = self._cpu_acc or [torch.zeros_like(p, dtype=torch.float32) for p in self._params]
            flat = layout.pack([a / steps for a in acc], dtype=torch.float32, device="cpu")
        weights = [float(b["weight"]) for b in blobs]
        throughputs = [b["throughput"] if b["mode"] != MODE_CLIENT else None for b in blobs]
        if sum(weights) <= 0:
            raise AllreduceException("group has no gradient contributors")
        try:
            averaged = self.dht.run_coroutine(
                lambda dht, node: self._cpu_averager.allreduce(group, flat, weights, throughputs))
        except AllreduceException:
            raise
        except Exception as e:
            raise AllreduceException(repr(e))
        if self.auxiliary:
            return
        if self.is_fused and self.on_gpu:
            from ..parallel.fused_allreduce import GroupSpec

            buf = self.opt.buffers
            buf.acc.copy_(averaged.to(buf.acc.device))
            self.local_steps_accumulated = 1
            self._encode_local(CompressionType.NONE)
            spec = GroupSpec(bases=[buf.symm.ptr], me=0, weights=[1.0], shards=[(0, layout.n_chunks)],
                             slots=[0], tag=self._next_tag())
            self._launch_fused(spec, CompressionType.NONE)
            return
        with torch.no_grad():
            for p, g in zip(self._params, layout.views(averaged)):
                p.grad = g.to(device=p.device, dtype=p.dtype).clone()
        self._generic_opt_step()

    # ------------------------------------------------------------------------------ state sharing (X2, H4)
    @property
    def _state_sharing_key(self) -> str:
        return f"{self.prefix}_averaging.all_averagers"

    def _publish_state_sharing(self) -> None:
        self._last_state_publish = get_dht_time()
        if not self.allow_state_sharing or self.client_mode or self._server is None:
            return
        self.dht.store(self._state_sharing_key, {"step": self.local_step}, get_dht_time() + self.metadata_expiration,
                       subkey=self.endpoint, return_future=True)

    def _maybe_republish_state_sharing(self) -> None:
        if get_dht_time() - getattr(self, "_last_state_publish", 0.0) > self.metadata_expiration / 2:
            self._publish_state_sharing()

    def _scheduler_state(self):
        try:
            return None if self.scheduler is None else {k: v for k, v in self.scheduler.state_dict().items()
                                                        if isinstance(v, (int, float, list, str, bool))}
        except Exception:
            return None

    async def _rpc_download_state(self, peer_host: str, payload) -> dict:
        """Donor side. `meta_only` answers are used by the requester as a seqlock around its NVLink copy."""
        if not self.allow_state_sharing:
            raise RuntimeError("this peer does not share its state")
        meta = {"local_step": self.local_step, "fused": bool(self.is_fused and self.on_gpu), "handle": None,
                "total": None, "opt_step": getattr(self.opt, "opt_step", None), "lineage": self.lineage,
                "lineage_depth": self.lineage_depth, "seq": self._state_seq, "host": _host_id(),
                "torn": bool(self._needs_resync),
                "moment_owners": None if self._moment_owners is None else list(self._moment_owners)}
        if self._needs_resync:
            raise RuntimeError("this peer's own state is torn; ask somebody else")
        if meta["fused"]:
            meta.update(handle=self.opt.buffers.handle, total=self.opt.layout.total)
        if payload.get("meta_only") or (meta["fused"] and payload.get("can_map")):
            return meta
        # host path: parameters + optimizer state as raw tensor bytes (never pickled: the receiver may be anybody)
        from ..utils.serialization import pack_state

        def snapshot():
            while True:
                locked = self.state_lock.acquire(timeout=0.05)
                if locked or self._params_parked:
                    break
            try:
                step_before = self.local_step
                blob = pack_state({"params": [p.detach().float().cpu().clone() for p in self._params],
                                   "opt": self.opt.state_dict()})
                if not locked and (not self._params_parked or self.local_step != step_before):
                    raise RuntimeError("state changed while it was being copied; retry")
                return blob
            finally:
                if locked:
                    self.state_lock.release()

        meta["blob"] = await asyncio.get_event_loop().run_in_executor(None, snapshot)
        return meta

    def load_state_from_peers(self, **kwargs) -> bool:
        """Download parameters + optimizer statistics + step from the most advanced peer that shares state
        (albert/run_trainer.py:124-128; run_first_peer.py:119-121). Returns False when nobody could serve us."""
        found = self.dht.get(self._state_sharing_key, latest=True)
        donors = []
        if found is not None and isinstance(found.value, dict):
            for ep, entry in found.value.items():
                if ep != self.endpoint and isinstance(entry.value, dict):
                    donors.append((int(entry.value.get("step", 0)), ep))
        donors.sort(reverse=True)
        can_map = bool(self.is_fused and self.on_gpu)
        for _, ep in donors:
            try:
                ok = self._download_from(ep, can_map)
            except Exception as e:
                logger.warning("state download from %s failed: %r", ep, e)
                ok = False
            if ok:
                self.stats["state_downloads"] += 1
                self._sync_scheduler()
                self.tracker.local_step = self.local_step
                if self._reports_progress:
                    self.tracker.report(self.local_step, self.local_samples_accumulated)
                return True
        if donors:
            logger.warning("could not load state from any of %d peers", len(donors))
        else:
            logger.info("no peers share state yet; keeping the local state")
        # nobody to learn from: adopt the collaboration's step so that we are not "behind" forever. A torn state that
        # nobody can replace is kept (its parameters are finite and close to every other replica); training goes on
        # and the lineage check heals the difference as soon as a consistent peer is in reach.
        self.local_step = max(self.local_step, self.collaboration_state.optimizer_step)
        if self._needs_resync:
            self._needs_resync = False
            self.lineage = self._mix(self.lineage, "T", self.local_step, self.endpoint)
        return False

    def _call(self, endpoint: str, method: str, payload: dict, timeout: float):
        return asyncio.run_coroutine_threadsafe(self._client.call(endpoint, method, payload, timeout),
                                                self.dht.loop).result()

    def _download_from(self, endpoint: str, can_map: bool, in_round: bool = False) -> bool:
        """`in_round`: the donor is a member of the round this peer is about to join (lineage healing). Its kernel may
        already be waiting for us -- nothing in that round modifies anybody's state before every member has launched
        (every member either sends a gradient or owns a share) -- so an odd sequence number is fine as long as it
        does not move."""
        # with can_map the first answer is metadata only: a donor that does not give it within seconds is frozen or dead
        meta = self._call(endpoint, "download_state", {"can_map": can_map},
                          self.kernel_timeout if can_map else self.averaging_timeout)
        handle = meta.get("handle")
        mappable = bool(handle) and any(bytes(handle)) and meta.get("host") == _host_id()
        if meta.get("fused") and can_map and mappable and meta["total"] == self.opt.layout.total:
            # NVLink path: map the donor's symmetric buffer and copy master params, bf16 mirror and moments
            buf = self.opt.buffers
            r = buf.regions
            for _attempt in range(5):
                base = buf.peer_base(bytes(meta["handle"]))
                self._mapping_leases.touch([meta["handle"]], time.monotonic())
                from .. import lib as native

                stream = torch.cuda.current_stream(buf.device)
                with torch.cuda.device(buf.device):
                    native.check(native.lib().dedloc_memcpy_async(buf.symm.ptr + r.p32, base + r.p32,
                                                                  r.state_end - r.p32, stream.cuda_stream),
                                 "state copy")
                    self._overlay_moments(meta, buf, stream)
                stream.synchronize()
                after = self._call(endpoint, "download_state", {"meta_only": True}, self.averaging_timeout)
                # seqlock: the donor's sequence number is odd while one of its rounds may be writing, and moves on
                # with every round; both reads must show the same even value (and the same step)
                quiet = (after.get("seq") == meta.get("seq") and after.get("opt_step") == meta.get("opt_step")
                         and (in_round or int(meta.get("seq") or 0) % 2 == 0))
                if after["local_step"] == meta["local_step"] and quiet:
                    self.local_step = int(meta["local_step"])
                    if meta.get("opt_step") is not None:
                        self.opt.opt_step = int(meta["opt_step"])
                    self.lineage = str(meta.get("lineage", self.lineage))
                    self.lineage_depth = int(meta.get("lineage_depth", self.lineage_depth) or 0)
                    self._moment_owners = None
                    self._needs_resync = False
                    return True
                meta = after
            return False
        if "blob" not in meta:
            meta = self._call(endpoint, "download_state", {"can_map": False}, self.averaging_timeout * 4)
        from ..utils.serialization import unpack_state

        blob = unpack_state(meta["blob"])
        if not isinstance(blob, dict) or not isinstance(blob.get("params"), list):
            raise ValueError("malformed state message")
        if len(blob["params"]) != len(self._params) or any(
                not isinstance(src, torch.Tensor) or tuple(src.shape) != tuple(p.shape) or not src.is_floating_point()
                for src, p in zip(blob["params"], self._params)):
            raise ValueError("the donor's parameters do not match this model")
        with self.state_lock, torch.no_grad():
            if self.is_fused and self.on_gpu:
                views32 = self.opt.layout.views(self.opt.buffers.p32)
                views16 = self.opt.layout.views(self.opt.buffers.p16)
                for v32, v16, src in zip(views32, views16, blob["params"]):
                    v32.copy_(src)
                    v16.copy_(src)
            else:
                for p, src in zip(self._params, blob["params"]):
                    p.copy_(src.to(p.dtype))
            try:
                self.opt.load_state_dict(blob["opt"])
            except Exception as e:  # e.g. the donor has not created optimizer state yet
                logger.debug("optimizer state not loaded: %r", e)
            self.local_step = int(meta["local_step"])
            self.lineage = str(meta.get("lineage", self.lineage))
            self.lineage_depth = int(meta.get("lineage_depth", self.lineage_depth) or 0)
            self._moment_owners = None
            self._needs_resync = False
        return True

    def _overlay_moments(self, meta: dict, buf, stream) -> None:
        """The donor's moments are current only for the chunks it owned in its last round; the other owners of that
        round hold the rest. Pull their chunks too where their buffers can be mapped (same host, still alive)."""
        owners = meta.get("moment_owners")
        if not owners or not buf.regions.m:
            return
        from .. import lib as native
        from ..parallel.flat import CHUNK

        r, K, C = buf.regions, len(owners), buf.layout.n_chunks
        donor = bytes(meta["handle"])
        for k, h in enumerate(owners):
            h = bytes(h)
            if h == donor:
                continue
            rows = (C - k + K - 1) // K
            if rows <= 0:
                continue
            src_is_me = h == buf.handle
            try:
                base = buf.symm.ptr if src_is_me else buf.peer_base(h)
                self._mapping_leases.touch([h], time.monotonic())
            except Exception:
                continue          # that owner is gone: its chunks keep the donor's (older) values
            if src_is_me:
                continue          # my own stale copy is no better than the donor's
            for off in (r.m, r.v):
                native.check(native.lib().dedloc_memcpy2d_async(
                    buf.symm.ptr + off + k * CHUNK * 4, K * CHUNK * 4, base + off + k * CHUNK * 4, K * CHUNK * 4,
                    CHUNK * 4, rows, stream.cuda_stream), "moment overlay")
