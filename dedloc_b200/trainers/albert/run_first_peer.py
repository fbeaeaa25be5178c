#!/usr/bin/env python
"""Coordinator / first peer (A2; parity with /root/reference/albert/run_first_peer.py:1-218).

Starts the root DHT node and prints its address (`--initial_peers` for everyone else), polls
`{prefix}_metrics` every `refresh_period` seconds, aggregates loss / throughput / alive peers (to wandb when it is
installed and `--wandb_project` is given, always to the log), and every `save_checkpoint_step_interval`
collaboration steps pulls the newest state from the peers (`load_state_from_peers`) and -- every
`upload_interval` seconds -- writes an HF-compatible checkpoint (`save_pretrained` + `optimizer_state.pt`) into
`repo_path`, committing and pushing it when that directory is a git repository.
"""
from __future__ import annotations

import logging
import subprocess
import time
from dataclasses import asdict
from typing import Optional

import torch

from ...collab.optimizer import CollaborativeOptimizer
from ...dht import DHT
from ...models.albert import AlbertForPreTraining
from ...optim import Lamb
from ...utils.argparse_dataclass import DataclassArgumentParser
from ...utils.compression import CompressionType
from ..pretrain import albert_param_groups
from . import metrics_utils
from .arguments import AveragerArguments, CollaborativeOptimizerArguments, CoordinatorArguments
from .run_trainer import load_config

logger = logging.getLogger(__name__)


def get_local_address() -> str:
    """The reference asks Google DNS for the public IP (whatsmyip, run_first_peer.py:153-155); without a network
    the routable local address is the best answer."""
    import socket

    try:
        s = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
        s.connect(("10.255.255.255", 1))
        addr = s.getsockname()[0]
        s.close()
        return addr
    except Exception:
        return "127.0.0.1"


class CheckpointHandler:
    """CPU copy of the model + LAMB wrapped in a CollaborativeOptimizer that never trains (no batch size); its only
    use is `load_state_from_peers()` followed by saving (run_first_peer.py:59-146)."""

    def __init__(self, coordinator_args: CoordinatorArguments, collab_optimizer_args: CollaborativeOptimizerArguments,
                 averager_args: AveragerArguments, dht: DHT):
        self.save_checkpoint_step_interval = coordinator_args.save_checkpoint_step_interval
        self.repo_path = coordinator_args.repo_path
        self.upload_interval = coordinator_args.upload_interval
        self.previous_step = -1
        self.previous_timestamp = time.time()
        config = load_config(coordinator_args.model_config_path)
        self.model = AlbertForPreTraining(config)
        opt = Lamb(albert_param_groups(self.model, 0.01), lr=0.00176, weight_decay=0.01, clamp_value=10000.0,
                   debias=True)
        adjusted_target_batch_size = collab_optimizer_args.target_batch_size - collab_optimizer_args.batch_size_lead
        self.collaborative_optimizer = CollaborativeOptimizer(
            opt=opt, dht=dht, prefix=coordinator_args.experiment_prefix,
            compression_type=CompressionType.Value(collab_optimizer_args.compression),
            throughput=collab_optimizer_args.bandwidth, target_batch_size=adjusted_target_batch_size,
            client_mode=collab_optimizer_args.client_mode, verbose=True, start=True, allow_state_sharing=False,
            **asdict(averager_args))
        self.previous_timestamp = time.time()

    def is_time_to_save_state(self, cur_step: int) -> bool:
        if self.save_checkpoint_step_interval is None:
            return False
        return cur_step - self.previous_step >= self.save_checkpoint_step_interval

    def save_state(self, cur_step: int) -> bool:
        ok = self.collaborative_optimizer.load_state_from_peers()
        self.previous_step = cur_step
        return ok

    def is_time_to_upload(self) -> bool:
        if self.repo_path is None:
            return False
        return self.upload_interval is None or time.time() - self.previous_timestamp >= self.upload_interval

    def upload_checkpoint(self, current_loss: float) -> None:
        self.model.save_pretrained(self.repo_path)
        torch.save(self.collaborative_optimizer.opt.state_dict(), f"{self.repo_path}/optimizer_state.pt")
        self.previous_timestamp = time.time()
        try:
            if subprocess.run("git rev-parse --is-inside-work-tree", shell=True, cwd=self.repo_path,
                              capture_output=True).returncode == 0:
                subprocess.run("git add --all", shell=True, check=True, cwd=self.repo_path)
                msg = f"Step {self.collaborative_optimizer.local_step}, loss {current_loss:.3f}"
                subprocess.run(f"git commit -m '{msg}'", shell=True, check=True, cwd=self.repo_path)
                subprocess.run("git push", shell=True, check=False, cwd=self.repo_path)
        except subprocess.CalledProcessError as e:
            logger.warning("error while uploading model: %s", e)


def aggregate_metrics(metrics_dict) -> Optional[dict]:
    """metrics_dict: {public key: ValueWithExpiration(LocalMetrics dict)} -> the numbers the reference logs."""
    metrics = []
    for entry in metrics_dict.values():
        try:
            metrics.append(metrics_utils.LocalMetrics.parse_obj(entry.value))
        except Exception:
            continue
    if not metrics:
        return None
    latest_step = max(m.step for m in metrics)
    sum_loss = sum(m.loss for m in metrics)
    sum_mini_steps = sum(m.mini_steps for m in metrics)
    return {"step": latest_step, "alive peers": len(metrics), "samples": sum(m.samples_accumulated for m in metrics),
            "performance": sum(m.samples_per_second for m in metrics),
            "loss": sum_loss / max(sum_mini_steps, 1)}


def main(argv=None):
    parser = DataclassArgumentParser((CoordinatorArguments, CollaborativeOptimizerArguments, AveragerArguments))
    coordinator_args, collab_optimizer_args, averager_args = parser.parse_args_into_dataclasses(argv)
    logging.basicConfig(level=logging.INFO, format="%(asctime)s [%(levelname)s] %(message)s")
    if coordinator_args.address is None:
        logger.warning("No address specified. Attempting to infer address from DNS.")
        coordinator_args.address = get_local_address()

    experiment_prefix = coordinator_args.experiment_prefix
    validators, local_public_key = metrics_utils.make_validators(experiment_prefix)
    dht = DHT(start=True, listen_on=coordinator_args.dht_listen_on, endpoint=f"{coordinator_args.address}:*",
              initial_peers=coordinator_args.initial_peers, record_validators=validators)
    logger.info(f"Running DHT root at {coordinator_args.address}:{dht.port}")

    wandb = None
    if coordinator_args.wandb_project is not None:
        try:
            import wandb as _wandb

            _wandb.init(project=coordinator_args.wandb_project, mode="offline")
            wandb = _wandb
        except Exception as e:
            logger.warning("wandb unavailable (%r): metrics go to the log only", e)

    current_step = 0
    checkpoint_handler = CheckpointHandler(coordinator_args, collab_optimizer_args, averager_args, dht)
    iterations = 0
    while True:
        metrics_entry = dht.get(experiment_prefix + "_metrics", latest=True)
        if metrics_entry is not None and isinstance(metrics_entry.value, dict) and len(metrics_entry.value) > 0:
            agg = aggregate_metrics(metrics_entry.value)
            if agg is not None and agg["step"] != current_step:
                current_step = agg["step"]
                logger.info(f"Step #{current_step}\tloss = {agg['loss']:.5f}\talive peers = {agg['alive peers']}"
                            f"\tsamples = {agg['samples']}\tperformance = {agg['performance']:.1f} samples/s")
                if wandb is not None:
                    wandb.log(agg)
                if checkpoint_handler.is_time_to_save_state(current_step):
                    checkpoint_handler.save_state(current_step)
                    if checkpoint_handler.is_time_to_upload():
                        checkpoint_handler.upload_checkpoint(agg["loss"])
        logger.debug("Peer is still alive...")
        iterations += 1
        if coordinator_args.max_iterations and iterations >= coordinator_args.max_iterations:
            break
        time.sleep(coordinator_args.refresh_period)
    checkpoint_handler.collaborative_optimizer.shutdown()
    dht.shutdown()


if __name__ == "__main__":
    main()
