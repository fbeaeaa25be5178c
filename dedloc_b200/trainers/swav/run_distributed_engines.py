#!/usr/bin/env python
"""CLI of the SwAV trainer -- same command line as the reference (swav/README.md:17-32,
vissl/tools/run_distributed_engines.py:21-58):

    python -m dedloc_b200.trainers.swav.run_distributed_engines hydra.verbose=true \
        config=pretrain/swav/swav_1node_resnet_submit config.DATA.TRAIN.BATCHSIZE_PER_REPLICA=64 \
        +config.OPTIMIZER.batch_size_for_tracking=64 +config.OPTIMIZER.lr=2.4 +config.OPTIMIZER.warmup_start_lr=0.3 \
        +config.OPTIMIZER.warmup_epochs=500 +config.OPTIMIZER.max_epochs=5000 +config.OPTIMIZER.eta_min=0.0048 \
        +config.OPTIMIZER.exp_prefix=my_swav +config.OPTIMIZER.target_group_size=4 \
        config.OPTIMIZER.dht_listen_on_port=1124 config.OPTIMIZER.averager_listen_on_port=1125 \
        +config.OPTIMIZER.dht_initial_peers=["host:port"] +config.DATA.TRAIN.DATA_PATHS=["/imagenet/train"]

One invocation = one peer = one GPU (NUM_NODES = NUM_PROC_PER_NODE = 1); `max_iterations=N` (extra key) bounds
smoke runs."""
import sys

from .config import compose
from .train import train_main


def hydra_main(overrides):
    if any(ov in ("-h", "--help", "help") for ov in overrides):
        print(__doc__)
        return None
    max_iterations = 0
    rest = []
    for ov in overrides:
        if ov.startswith("max_iterations="):
            max_iterations = int(ov.split("=", 1)[1])
        else:
            rest.append(ov)
    cfg = compose(rest)
    return train_main(cfg, max_iterations=max_iterations)


if __name__ == "__main__":
    hydra_main(sys.argv[1:])
