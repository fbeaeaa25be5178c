#!/usr/bin/env python
"""sahajBERT trainer peer (S1; behaviour of /root/reference/sahajbert/run_trainer.py:1-304): the ALBERT trainer
with (a) the Bengali fast tokenizer, (b) the lazy wiki+OSCAR streaming dataset shuffled with a per-peer seed
`hash(public key) % 2**31`, (c) `dataloader_num_workers == 0` and no evaluation (asserted), (d) metrics published
only while the peer is synchronised with the collaboration, (e) no mandatory `--initial_peers` (a volunteer may be
the first one up). Volunteers typically run with `--client_mode --batch_size_lead 400
--averaging_expiration 10 --statistics_expiration 120` (contributor_notebook.ipynb cell 1)."""
from __future__ import annotations

import logging
import os
from dataclasses import asdict
from typing import Dict, Iterator

import torch

from ...collab.optimizer import CollaborativeOptimizer
from ...dht import DHT
from ...utils.argparse_dataclass import DataclassArgumentParser
from ...utils.compression import CompressionType
from ..albert import metrics_utils
from ..albert.data import MLMCollator
from ..albert.run_trainer import (CollaborativeCallback, get_model, get_optimizer_and_scheduler,
                                  independent_shuffling_seed, load_config, save_checkpoint, setup_logging, train_loop)
from ..pretrain import SyntheticMLMBatches
from .arguments import AlbertTrainingArguments, CollaborationArguments, DatasetArguments

logger = logging.getLogger(__name__)


def load_tokenizer(dataset_args):
    from .tokenization_albert_bengali_fast import AlbertBengaliTokenizerFast

    path = dataset_args.tokenizer_path
    if not os.path.exists(path):
        # the default "tokenizer/data" is relative to the sahajbert directory (the reference is run from there)
        packaged = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tokenizer", "data")
        if os.path.isdir(packaged):
            logger.info("tokenizer path %r not found: using the packaged assets in %s", path, packaged)
            path = packaged
    if os.path.isdir(path):
        path = os.path.join(path, "tokenizer.json")
    return AlbertBengaliTokenizerFast(tokenizer_file=path)


def streaming_batches(tokenizer, training_args, dataset_args, seed: int) -> Iterator[Dict[str, torch.Tensor]]:
    from .dataset_streaming import make_lazy_wikioscar_dataset, text_file_stream

    sources = None
    if dataset_args.local_text_files:
        sources = [text_file_stream(p) for p in dataset_args.local_text_files.split(",")]
    dataset = make_lazy_wikioscar_dataset(tokenizer, shuffle_seed=seed, sources=sources,
                                          max_seq_length=training_args.seq_length,
                                          shuffle_buffer_size=10 ** 4 if sources is None else 256,
                                          preprocessing_batch_size=256 if sources is None else 8)
    collator = MLMCollator(len(tokenizer), tokenizer.mask_token_id, tokenizer.pad_token_id, training_args.seq_length,
                           generator=torch.Generator().manual_seed(seed))
    bs = training_args.per_device_train_batch_size * training_args.gradient_accumulation_steps
    buf = []
    for sample in dataset:
        buf.append(sample)
        if len(buf) == bs:
            yield collator(buf)
            buf = []


def main(argv=None):
    parser = DataclassArgumentParser((AlbertTrainingArguments, DatasetArguments, CollaborationArguments))
    training_args, dataset_args, collaboration_args = parser.parse_args_into_dataclasses(argv)
    logger.info(f"Found {len(collaboration_args.initial_peers)} initial peers: {collaboration_args.initial_peers}")
    collaboration_args_dict = asdict(collaboration_args)
    setup_logging(training_args)
    torch.manual_seed(training_args.seed)
    assert training_args.dataloader_num_workers == 0, "the streaming dataset must be consumed in the main process"
    assert not training_args.do_eval, "evaluation is not supported with the lazy streaming dataset"

    device = torch.device(training_args.device or ("cuda" if torch.cuda.is_available() else "cpu"))
    on_gpu = device.type == "cuda"
    config = load_config(dataset_args.config_path)
    tokenizer = None if dataset_args.synthetic_data else load_tokenizer(dataset_args)
    vocab = len(tokenizer) if tokenizer is not None else 32000   # sahajBERT vocabulary: 31 995 pieces + 5 specials
    model = get_model(training_args, config, vocab, impl="native" if on_gpu else "torch")
    if on_gpu:
        for p in model.parameters():
            p._dedloc_master = p.detach().clone().to(device)
        model = model.to(device=device, dtype=torch.bfloat16)
    model.train()
    opt, scheduler = get_optimizer_and_scheduler(training_args, model, fused=on_gpu)

    validators, local_public_key = metrics_utils.make_validators(collaboration_args_dict["experiment_prefix"])
    dht = DHT(start=True, initial_peers=collaboration_args_dict.pop("initial_peers"),
              listen=not collaboration_args_dict["client_mode"], listen_on=collaboration_args_dict.pop("dht_listen_on"),
              endpoint=collaboration_args_dict.pop("endpoint"), record_validators=validators)
    total_batch_size_per_step = training_args.per_device_train_batch_size * training_args.gradient_accumulation_steps
    statistics_expiration = collaboration_args_dict.pop("statistics_expiration")
    adjusted_target_batch_size = collaboration_args_dict.pop("target_batch_size") - collaboration_args_dict.pop(
        "batch_size_lead")
    authorizer = None
    if getattr(training_args, "authorize", False):
        from .huggingface_auth import authorize_with_huggingface

        authorizer = authorize_with_huggingface()     # contributor_notebook.ipynb cell 3
    collaborative_optimizer = CollaborativeOptimizer(
        opt=opt, dht=dht, scheduler=scheduler, prefix=collaboration_args_dict.pop("experiment_prefix"),
        authorizer=authorizer,
        compression_type=CompressionType.Value(collaboration_args_dict.pop("compression")),
        batch_size_per_step=total_batch_size_per_step, throughput=collaboration_args_dict.pop("bandwidth"),
        target_batch_size=adjusted_target_batch_size, client_mode=collaboration_args_dict.pop("client_mode"),
        verbose=True, start=True, max_grad_norm=training_args.max_grad_norm, peer_key=local_public_key,
        **collaboration_args_dict)
    callback = CollaborativeCallback(dht, collaborative_optimizer, model, local_public_key, statistics_expiration)
    if not training_args.do_train:
        return

    seed = independent_shuffling_seed(local_public_key)
    if tokenizer is None:
        synth = SyntheticMLMBatches(model.config.vocab_size, min(training_args.seq_length, model.config.max_position_embeddings),
                                    total_batch_size_per_step, seed=seed, pool=32)
        batches = iter(lambda: synth.next_host(), None)
    else:
        batches = streaming_batches(tokenizer, training_args, dataset_args, seed)
    callback.on_train_begin()
    # the same loop as the ALBERT trainer (PeerTrainer: CUDA-graph micro-step, one host read per collaborative step)
    train_loop(training_args, model, opt, scheduler, collaborative_optimizer, callback, batches, device,
               total_batch_size_per_step, entry="dedloc_b200.trainers.sahajbert.run_trainer")
    collaborative_optimizer.shutdown()
    dht.shutdown()


if __name__ == "__main__":
    main()
