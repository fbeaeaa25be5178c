#!/usr/bin/env python
"""ALBERT trainer peer (A1; parity with /root/reference/albert/run_trainer.py:1-297).

    python -m dedloc_b200.trainers.albert.run_trainer --experiment_prefix my_exp \
        --initial_peers 127.0.0.1:31337 --per_device_train_batch_size 32 [--synthetic_data]

Same flags, same life cycle: parse three argument groups -> model (newest `checkpoint*` in output_dir, else fresh +
resize_token_embeddings) -> LAMB with two weight-decay groups + linear warm-up -> DHT with schema + signature
validators -> CollaborativeOptimizer -> training loop with an independent per-peer shuffling seed ->
`CollaborativeCallback`: state download on start, non-finite guard with rollback, metrics published to
`{prefix}_metrics[public key]` whenever the collaboration step advances, checkpoint every `save_steps`.
The loop is `trainers/pretrain.py:PeerTrainer` -- the very code path bench.py times: micro-step captured in a CUDA
graph, at most one micro-step of host run-ahead, loss accumulated on the device and read once per collaborative step,
no per-step host synchronisation (the non-finite guard lives in the all-reduce kernel, which skips the update of a
tensor whose averaged gradient is not finite). `--use_hf_trainer` drives the same optimizer through
transformers.Trainer exactly like the reference does.
"""
from __future__ import annotations

import logging
import os
import shutil
import time
from dataclasses import asdict
from pathlib import Path
from typing import Any, Dict, Iterator, Optional

import torch

from ...collab.optimizer import CollaborativeOptimizer
from ...dht import DHT, get_dht_time
from ...models.albert import AlbertConfig, AlbertForPreTraining
from ...optim import Lamb, NoOpScheduler, get_linear_schedule_with_warmup
from ...optim.fused import FusedLamb
from ...utils.argparse_dataclass import DataclassArgumentParser
from ...utils.compression import CompressionType
from ..pretrain import SyntheticMLMBatches, albert_param_groups
from . import metrics_utils
from .arguments import AlbertTrainingArguments, CollaborationArguments, DatasetArguments

logger = logging.getLogger(__name__)


def setup_logging(training_args) -> None:
    logging.basicConfig(format="%(asctime)s - %(levelname)s - %(name)s -   %(message)s", datefmt="%m/%d/%Y %H:%M:%S",
                        level=logging.INFO if training_args.local_rank in (-1, 0) else logging.WARN)
    logger.info("Training/evaluation parameters %s", training_args)


def load_config(path: str) -> AlbertConfig:
    if path and os.path.exists(path):
        return AlbertConfig.from_pretrained(path)
    if path and path.startswith("http"):
        logger.warning("no network access: using the built-in albert-large-v2 dimensions instead of %s", path)
    return AlbertConfig.albert_large_v2()


def latest_checkpoint(output_dir: str) -> Optional[Path]:
    return max(Path(output_dir).glob("checkpoint*"), default=None, key=os.path.getctime)


def get_model(training_args, config: AlbertConfig, vocab_size: Optional[int], impl: str) -> AlbertForPreTraining:
    latest = latest_checkpoint(training_args.output_dir)
    logger.info(f"Checkpoint dir {training_args.output_dir}, latest {latest}")
    if latest is not None:
        logger.info(f"Loading model from {latest}")
        return AlbertForPreTraining.from_pretrained(str(latest), impl=impl)
    logger.info("Training from scratch")
    model = AlbertForPreTraining(config, impl=impl)
    if vocab_size is not None:
        model.resize_token_embeddings(vocab_size)
    return model


def get_optimizer_and_scheduler(training_args, model, fused: bool):
    groups = albert_param_groups(model, training_args.weight_decay)
    cls = FusedLamb if fused else Lamb
    opt = cls(groups, lr=training_args.learning_rate, betas=(training_args.adam_beta1, training_args.adam_beta2),
              eps=training_args.adam_epsilon, weight_decay=training_args.weight_decay,
              clamp_value=training_args.clamp_value, debias=True)
    scheduler = get_linear_schedule_with_warmup(opt, num_warmup_steps=training_args.warmup_steps,
                                                num_training_steps=training_args.max_steps)
    return opt, scheduler


class CollaborativeCallback:
    """State sync, divergence guard and metric publishing around every local step (run_trainer.py:103-186).
    Also usable as a `transformers.TrainerCallback` (same method names and signatures)."""

    def __init__(self, dht: DHT, optimizer: CollaborativeOptimizer, model: torch.nn.Module,
                 local_public_key: bytes, statistics_expiration: float):
        self.model, self.dht, self.collaborative_optimizer = model, dht, optimizer
        self.local_public_key, self.statistics_expiration = local_public_key, statistics_expiration
        self.last_reported_collaboration_step = -1
        self.previous_state = self.get_current_state()
        self.samples = 0
        self.steps = 0
        self.loss = 0.0
        self.total_samples_processed = 0

    def on_train_begin(self, args=None, state=None, control=None, **kwargs):
        logger.warning("Loading state from peers")
        self.collaborative_optimizer.load_state_from_peers()

    def on_step_end(self, args=None, state=None, control=None, loss: Optional[float] = None, mini_steps: int = 1,
                    samples: Optional[int] = None, **kwargs):
        """`loss` is the SUM of the losses of `mini_steps` local steps (the reference calls this once per local step;
        the PeerTrainer loop calls it once per collaborative step with everything that happened since)."""
        if control is not None:
            control.should_log = True
        if samples is not None:
            self.samples = samples
        if not self.params_are_finite():
            logger.warning("non-finite parameters detected: rolling back to the previous state")
            self.load_from_state(self.previous_state)
            return control
        if self.collaborative_optimizer.local_step != self.last_reported_collaboration_step:
            # snapshot only when the parameters changed (the reference snapshots references on every step)
            self.previous_state = self.get_current_state()
        if loss is None and state is not None and getattr(state, "log_history", None):
            loss = state.log_history[-1].get("loss")
        if loss is not None:
            self.loss += float(loss)
            self.steps += mini_steps
            if self.collaborative_optimizer.local_step != self.last_reported_collaboration_step:
                self.last_reported_collaboration_step = self.collaborative_optimizer.local_step
                self.total_samples_processed += self.samples
                statistics = metrics_utils.LocalMetrics(
                    step=self.collaborative_optimizer.local_step,
                    samples_per_second=float(self.collaborative_optimizer.performance_ema.samples_per_second),
                    samples_accumulated=self.samples, loss=self.loss, mini_steps=self.steps)
                logger.info(f"Step {self.collaborative_optimizer.local_step}")
                logger.info(f"Your current contribution: {self.total_samples_processed} samples")
                if self.steps:
                    logger.info(f"Loss of your model: {self.loss / self.steps}")
                self.loss = 0.0
                self.steps = 0
                if self.collaborative_optimizer.is_synchronized:   # sahajbert/run_trainer.py:164-171
                    self.dht.store(key=self.collaborative_optimizer.prefix + "_metrics",
                                   subkey=self.local_public_key, value=statistics.dict(),
                                   expiration_time=get_dht_time() + self.statistics_expiration, return_future=True)
        if samples is None:
            self.samples = self.collaborative_optimizer.local_samples_accumulated
        return control

    @torch.no_grad()
    def get_current_state(self) -> Dict[str, Any]:
        opt = self.collaborative_optimizer.opt
        if isinstance(opt, FusedLamb) and opt.is_materialized:
            # no 3 x 71 MB clone per step: the fused kernel never commits a non-finite update (it skips the tensor),
            # so there is nothing to roll back to; a peer whose state is damaged resynchronises from its peers
            return {"fused": None}
        return {"model": {k: v.clone() for k, v in self.model.state_dict().items()},
                "opt": _clone_state(opt.state_dict())}

    @torch.no_grad()
    def load_from_state(self, state) -> None:
        opt = self.collaborative_optimizer.opt
        if "fused" in state:
            if state["fused"] is None:
                self.collaborative_optimizer.load_state_from_peers()
                return
            p32, m, v, step = state["fused"]
            b = opt.buffers
            b.p32.copy_(p32)
            b.p16.copy_(p32)
            b.m.copy_(m)
            b.v.copy_(v)
            opt.opt_step = step
            return
        self.model.load_state_dict(state["model"])
        opt.load_state_dict(state["opt"])

    @torch.no_grad()
    def params_are_finite(self) -> bool:
        opt = self.collaborative_optimizer.opt
        if isinstance(opt, FusedLamb) and opt.is_materialized:
            return bool(torch.isfinite(opt.buffers.p32).all())   # one flat reduction instead of 32 syncing calls
        return all(bool(torch.isfinite(p).all()) for p in self.model.parameters())


def _clone_state(sd):
    if isinstance(sd, torch.Tensor):
        return sd.clone()
    if isinstance(sd, dict):
        return {k: _clone_state(v) for k, v in sd.items()}
    if isinstance(sd, (list, tuple)):
        return type(sd)(_clone_state(v) for v in sd)
    return sd


def independent_shuffling_seed(local_public_key: bytes) -> int:
    """Shuffle data independently for each peer to avoid duplicating batches (run_trainer.py:266-270)."""
    return hash(local_public_key) % (2 ** 31)


def make_batches(training_args, dataset_args, config: AlbertConfig, seed: int) -> Iterator[Dict[str, torch.Tensor]]:
    bs = training_args.per_device_train_batch_size * training_args.gradient_accumulation_steps
    seq = min(training_args.seq_length, config.max_position_embeddings)
    if dataset_args.synthetic_data or not (dataset_args.dataset_path and os.path.exists(dataset_args.dataset_path)):
        if not dataset_args.synthetic_data:
            logger.warning("dataset %s not found: falling back to synthetic tokens", dataset_args.dataset_path)
        data = SyntheticMLMBatches(config.vocab_size, seq, bs, seed=seed, pool=32)
        while True:
            yield data.next_host()
    from .data import tokenized_dataset_batches

    yield from tokenized_dataset_batches(dataset_args.dataset_path, dataset_args.tokenizer_path, bs, seq, seed,
                                         num_workers=training_args.dataloader_num_workers)


def save_checkpoint(training_args, model, collab: CollaborativeOptimizer, step: int) -> None:
    out = Path(training_args.output_dir) / f"checkpoint-{step}"
    model.save_pretrained(str(out))
    torch.save(collab.state_dict(), out / "optimizer.pt")          # = CollaborativeOptimizer.state_dict()
    torch.save({}, out / "scheduler.pt")                           # NoOpScheduler.state_dict()
    import json

    (out / "trainer_state.json").write_text(json.dumps({"global_step": step, "local_step": collab.local_step}))
    ckpts = sorted(Path(training_args.output_dir).glob("checkpoint-*"), key=os.path.getctime)
    for old in ckpts[:-training_args.save_total_limit] if training_args.save_total_limit > 0 else []:
        shutil.rmtree(old, ignore_errors=True)


def main(argv=None):
    parser = DataclassArgumentParser((AlbertTrainingArguments, DatasetArguments, CollaborationArguments))
    training_args, dataset_args, collaboration_args = parser.parse_args_into_dataclasses(argv)
    logger.info(f"Found {len(collaboration_args.initial_peers)} initial peers: {collaboration_args.initial_peers}")
    if len(collaboration_args.initial_peers) == 0:
        raise ValueError("Please specify at least one network endpoint in initial peers.")
    collaboration_args_dict = asdict(collaboration_args)
    setup_logging(training_args)
    torch.manual_seed(training_args.seed)

    device = torch.device(training_args.device or ("cuda" if torch.cuda.is_available() else "cpu"))
    on_gpu = device.type == "cuda"
    config = load_config(dataset_args.config_path)
    vocab_size = None
    if dataset_args.tokenizer_path and os.path.exists(dataset_args.tokenizer_path) and not dataset_args.synthetic_data:
        from transformers import AlbertTokenizerFast

        vocab_size = len(AlbertTokenizerFast.from_pretrained(dataset_args.tokenizer_path, cache_dir=dataset_args.cache_dir))
    model = get_model(training_args, config, vocab_size, impl="native" if on_gpu else "torch")
    if on_gpu:
        for p in model.parameters():
            p._dedloc_master = p.detach().clone().to(device)
        model = model.to(device=device, dtype=torch.bfloat16)
    model.train()
    opt, scheduler = get_optimizer_and_scheduler(training_args, model, fused=on_gpu)

    validators, local_public_key = metrics_utils.make_validators(collaboration_args_dict["experiment_prefix"])
    dht = DHT(start=True, initial_peers=collaboration_args_dict.pop("initial_peers"),
              listen=not collaboration_args_dict["client_mode"],
              listen_on=collaboration_args_dict.pop("dht_listen_on"),
              endpoint=collaboration_args_dict.pop("endpoint"), record_validators=validators)

    total_batch_size_per_step = training_args.per_device_train_batch_size * training_args.gradient_accumulation_steps
    statistics_expiration = collaboration_args_dict.pop("statistics_expiration")
    adjusted_target_batch_size = collaboration_args_dict.pop("target_batch_size") - collaboration_args_dict.pop(
        "batch_size_lead")
    collaborative_optimizer = CollaborativeOptimizer(
        opt=opt, dht=dht, scheduler=scheduler, prefix=collaboration_args_dict.pop("experiment_prefix"),
        compression_type=CompressionType.Value(collaboration_args_dict.pop("compression")),
        batch_size_per_step=total_batch_size_per_step, throughput=collaboration_args_dict.pop("bandwidth"),
        target_batch_size=adjusted_target_batch_size, client_mode=collaboration_args_dict.pop("client_mode"),
        verbose=True, start=True, max_grad_norm=training_args.max_grad_norm, peer_key=local_public_key,
        **collaboration_args_dict)
    callback = CollaborativeCallback(dht, collaborative_optimizer, model, local_public_key, statistics_expiration)

    if training_args.use_hf_trainer:
        return _train_with_hf_trainer(training_args, dataset_args, model, collaborative_optimizer, callback)

    if not training_args.do_train:
        return
    callback.on_train_begin()
    batches = make_batches(training_args, dataset_args, model.config, independent_shuffling_seed(local_public_key))
    train_loop(training_args, model, opt, scheduler, collaborative_optimizer, callback, batches, device,
               total_batch_size_per_step, entry="dedloc_b200.trainers.albert.run_trainer")
    collaborative_optimizer.shutdown()
    dht.shutdown()


def train_loop(training_args, model, opt, scheduler, collaborative_optimizer, callback, batches, device,
               total_batch_size_per_step: int, entry: str) -> None:
    """The training loop of the ALBERT and sahajBERT trainers: `PeerTrainer` (the code path bench.py times) plus the
    reference's per-step bookkeeping (CollaborativeCallback, checkpoints every `save_steps`), done once per
    collaborative step."""
    from ..pretrain import PeerConfig, PeerTrainer

    on_gpu = device.type == "cuda"
    peer_cfg = PeerConfig(micro_batch=total_batch_size_per_step,
                          gradient_accumulation_steps=training_args.gradient_accumulation_steps,
                          seq_len=training_args.seq_length)
    trainer = PeerTrainer(model, opt, scheduler, collaborative_optimizer, peer_cfg, device)
    first = next(batches)
    if on_gpu and not training_args.no_cuda_graph:
        trainer.enable_cuda_graph(first)
    local_steps = 0
    last_saved = collaborative_optimizer.local_step
    pending: Optional[Dict[str, torch.Tensor]] = first
    samples_at_last_step = 0
    marks = []          # (wall clock, samples processed) right after every collaborative step's loss read
    t_log = time.time()
    while collaborative_optimizer.local_step < training_args.max_steps:
        host = pending if pending is not None else next(batches)
        pending = None
        mini_before = trainer.mini_steps
        stepped = trainer.micro_step(host)                # forward/backward (graph replay) + CollaborativeOptimizer.step()
        local_steps += 1
        if stepped:
            # one device->host read per collaborative step: everything the reference reports after every local step
            n_mini = max(trainer.mini_steps, mini_before + 1)
            mean_loss = trainer.pop_mean_loss()
            callback.on_step_end(loss=mean_loss * n_mini, mini_steps=n_mini,
                                 samples=trainer.samples_processed - samples_at_last_step)
            samples_at_last_step = trainer.samples_processed
            marks.append((time.time(), trainer.samples_processed))
            step = collaborative_optimizer.local_step
            if training_args.logging_steps and step % max(training_args.logging_steps, 1) == 0:
                dt = time.time() - t_log
                logger.info(f"collaborative step {step}: {trainer.samples_processed / max(dt, 1e-9):.1f} samples/s locally")
            if step != last_saved and step % training_args.save_steps == 0:
                save_checkpoint(training_args, model, collaborative_optimizer, step)
                last_saved = step
        if training_args.max_local_steps and local_steps >= training_args.max_local_steps:
            break
    if on_gpu:
        torch.cuda.synchronize(device)
    if training_args.report_json:
        import json

        dt = time.time() - t_log
        steady = None
        if len(marks) >= 5:       # from the end of the 3rd collaborative step to the end of the last one
            (ta, sa), (tb, sb) = marks[2], marks[-1]
            steady = round((sb - sa) / max(tb - ta, 1e-9), 2)
        print(json.dumps({"entry": entry, "local_steps": local_steps, "samples_per_second_steady": steady,
                          "collaborative_steps": collaborative_optimizer.local_step,
                          "samples": trainer.samples_processed, "seconds": round(dt, 3),
                          "samples_per_second_incl_startup": round(trainer.samples_processed / max(dt, 1e-9), 2),
                          "cuda_graph": trainer._graph is not None,
                          "collab": {k: v for k, v in collaborative_optimizer.stats.items() if isinstance(v, (int, float))}}),
              flush=True)


def _train_with_hf_trainer(training_args, dataset_args, model, collaborative_optimizer, callback):
    """The reference's exact driving code: transformers.Trainer(optimizers=(collab_opt, NoOpScheduler))."""
    import transformers
    from datasets import load_from_disk

    tokenizer = transformers.AlbertTokenizerFast.from_pretrained(dataset_args.tokenizer_path)
    tokenized = load_from_disk(dataset_args.dataset_path)
    hf_args = transformers.TrainingArguments(
        output_dir=training_args.output_dir, per_device_train_batch_size=training_args.per_device_train_batch_size,
        gradient_accumulation_steps=training_args.gradient_accumulation_steps, max_steps=training_args.max_steps,
        max_grad_norm=0.0, logging_steps=1, save_steps=training_args.save_steps,
        save_total_limit=training_args.save_total_limit, dataloader_num_workers=training_args.dataloader_num_workers,
        bf16=training_args.bf16 and torch.cuda.is_available(), report_to=[])

    class HFCallback(transformers.TrainerCallback):
        def on_train_begin(self, args, state, control, **kw):
            callback.on_train_begin(args, state, control)

        def on_step_end(self, args, state, control, **kw):
            return callback.on_step_end(args, state, control)

    class TrainerWithIndependentShuffling(transformers.Trainer):
        def get_train_dataloader(self):
            torch.manual_seed(independent_shuffling_seed(callback.local_public_key))
            return super().get_train_dataloader()

    trainer = TrainerWithIndependentShuffling(
        model=model, args=hf_args, processing_class=tokenizer,
        data_collator=transformers.DataCollatorForLanguageModeling(tokenizer=tokenizer),
        train_dataset=tokenized["train"], optimizers=(collaborative_optimizer, NoOpScheduler(collaborative_optimizer)),
        callbacks=[HFCallback()])
    trainer.remove_callback(transformers.trainer_callback.PrinterCallback)
    trainer.remove_callback(transformers.trainer_callback.ProgressCallback)
    trainer.train(resume_from_checkpoint=latest_checkpoint(training_args.output_dir))


if __name__ == "__main__":
    main()
