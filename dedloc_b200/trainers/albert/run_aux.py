#!/usr/bin/env python
"""Auxiliary peer (A3; parity with /root/reference/albert/run_aux.py:1-266): same construction as a trainer but
no data and no training loop -- `CollaborativeOptimizer(auxiliary=True, allow_state_sharing=False)` and a loop of
`sleep(0.5); step_aux()` that donates this machine's bandwidth (a B200's NVLink shard, or a CPU host's sockets)
to the averaging rounds with weight 0."""
from __future__ import annotations

import logging
import time
from dataclasses import asdict

import torch

from ...collab.optimizer import CollaborativeOptimizer
from ...dht import DHT
from ...utils.argparse_dataclass import DataclassArgumentParser
from ...utils.compression import CompressionType
from . import metrics_utils
from .arguments import AlbertTrainingArguments, CollaborationArguments, DatasetArguments
from .run_trainer import get_model, get_optimizer_and_scheduler, load_config, setup_logging

logger = logging.getLogger(__name__)


def main(argv=None):
    parser = DataclassArgumentParser((AlbertTrainingArguments, DatasetArguments, CollaborationArguments))
    training_args, dataset_args, collaboration_args = parser.parse_args_into_dataclasses(argv)
    logger.info(f"Found {len(collaboration_args.initial_peers)} initial peers: {collaboration_args.initial_peers}")
    if len(collaboration_args.initial_peers) == 0:
        raise ValueError("Please specify at least one network endpoint in initial peers.")
    collaboration_args_dict = asdict(collaboration_args)
    setup_logging(training_args)
    torch.manual_seed(training_args.seed)
    device = torch.device(training_args.device or "cpu")   # the reference runs aux peers on CPU-only machines
    on_gpu = device.type == "cuda"

    config = load_config(dataset_args.config_path)
    model = get_model(training_args, config, None, impl="native" if on_gpu else "torch")
    if on_gpu:
        for p in model.parameters():
            p._dedloc_master = p.detach().clone().to(device)
        model = model.to(device=device, dtype=torch.bfloat16)
    opt, scheduler = get_optimizer_and_scheduler(training_args, model, fused=on_gpu)

    validators, local_public_key = metrics_utils.make_validators(collaboration_args_dict["experiment_prefix"])
    dht = DHT(start=True, initial_peers=collaboration_args_dict.pop("initial_peers"),
              listen=not collaboration_args_dict["client_mode"], listen_on=collaboration_args_dict.pop("dht_listen_on"),
              endpoint=collaboration_args_dict.pop("endpoint"), record_validators=validators)

    total_batch_size_per_step = training_args.per_device_train_batch_size * training_args.gradient_accumulation_steps
    collaboration_args_dict.pop("statistics_expiration")
    adjusted_target_batch_size = collaboration_args_dict.pop("target_batch_size") - collaboration_args_dict.pop(
        "batch_size_lead")
    collaborative_optimizer = CollaborativeOptimizer(
        opt=opt, dht=dht, scheduler=scheduler, prefix=collaboration_args_dict.pop("experiment_prefix"),
        compression_type=CompressionType.Value(collaboration_args_dict.pop("compression")),
        batch_size_per_step=total_batch_size_per_step, throughput=collaboration_args_dict.pop("bandwidth"),
        target_batch_size=adjusted_target_batch_size, client_mode=collaboration_args_dict.pop("client_mode"),
        verbose=True, start=True, auxiliary=True, allow_state_sharing=False, peer_key=local_public_key,
        **collaboration_args_dict)

    rounds = 0
    try:
        while True:
            time.sleep(0.5)
            if collaborative_optimizer.step_aux():
                rounds += 1
            if training_args.max_local_steps and rounds >= training_args.max_local_steps:
                break
    finally:
        collaborative_optimizer.shutdown()
        dht.shutdown()


if __name__ == "__main__":
    main()
