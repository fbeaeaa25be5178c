#!/usr/bin/env python
"""sahajBERT coordinator (S2; behaviour of /root/reference/sahajbert/run_first_peer.py:1-217): the ALBERT
coordinator with `save_checkpoint_step_interval` optional (None = never pull state) and the checkpoint model's
embeddings resized to the Bengali tokenizer's vocabulary (32 000)."""
from __future__ import annotations

import logging
import os
import time

from ...dht import DHT
from ...utils.argparse_dataclass import DataclassArgumentParser
from ..albert import metrics_utils
from ..albert.run_first_peer import CheckpointHandler, aggregate_metrics, get_local_address
from .arguments import AveragerArguments, CollaborativeOptimizerArguments, CoordinatorArguments

logger = logging.getLogger(__name__)
SAHAJBERT_VOCAB_SIZE = 32000


def tokenizer_vocab_size(path: str = "tokenizer/data") -> int:
    tok = os.path.join(path, "tokenizer.json") if os.path.isdir(path) else path
    if os.path.exists(tok):
        from .tokenization_albert_bengali_fast import AlbertBengaliTokenizerFast

        return len(AlbertBengaliTokenizerFast(tokenizer_file=tok))
    return SAHAJBERT_VOCAB_SIZE


def main(argv=None):
    parser = DataclassArgumentParser((CoordinatorArguments, CollaborativeOptimizerArguments, AveragerArguments))
    coordinator_args, collab_optimizer_args, averager_args = parser.parse_args_into_dataclasses(argv)
    logging.basicConfig(level=logging.INFO, format="%(asctime)s [%(levelname)s] %(message)s")
    if coordinator_args.address is None:
        logger.warning("No address specified. Attempting to infer address from DNS.")
        coordinator_args.address = get_local_address()
    prefix = coordinator_args.experiment_prefix
    validators, _ = metrics_utils.make_validators(prefix)
    dht = DHT(start=True, listen_on=coordinator_args.dht_listen_on, endpoint=f"{coordinator_args.address}:*",
              initial_peers=coordinator_args.initial_peers, record_validators=validators)
    logger.info(f"Running DHT root at {coordinator_args.address}:{dht.port}")
    handler = CheckpointHandler(coordinator_args, collab_optimizer_args, averager_args, dht)
    handler.model.resize_token_embeddings(tokenizer_vocab_size())
    current_step, iterations = 0, 0
    while True:
        entry = dht.get(prefix + "_metrics", latest=True)
        if entry is not None and isinstance(entry.value, dict) and entry.value:
            agg = aggregate_metrics(entry.value)
            if agg is not None and agg["step"] != current_step:
                current_step = agg["step"]
                logger.info(f"Step #{current_step}\tloss = {agg['loss']:.5f}\talive peers = {agg['alive peers']}")
                if handler.is_time_to_save_state(current_step):
                    handler.save_state(current_step)
                    if handler.is_time_to_upload():
                        handler.upload_checkpoint(agg["loss"])
        iterations += 1
        if coordinator_args.max_iterations and iterations >= coordinator_args.max_iterations:
            break
        time.sleep(coordinator_args.refresh_period)
    handler.collaborative_optimizer.shutdown()
    dht.shutdown()


if __name__ == "__main__":
    main()
