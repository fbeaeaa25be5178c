"""Shared fine-tuning loop of the downstream evaluations (S7): AdamW + linear warm-up, evaluation every epoch,
early stopping on the monitored metric, best checkpoint kept. No collaboration involved (plain single-GPU/CPU)."""
from __future__ import annotations

import copy
import logging
from dataclasses import dataclass
from typing import Callable, Dict, Iterable, List, Optional, Sequence

import torch

from ...optim.schedules import get_linear_schedule_with_warmup

logger = logging.getLogger(__name__)


@dataclass
class FinetuneArguments:
    model_path: Optional[str] = None          # directory with an HF-compatible sahajBERT / ALBERT checkpoint
    tokenizer_path: str = "tokenizer/data"
    output_dir: str = "finetune_outputs"
    max_seq_length: int = 128                 # sahajbert/README.md:83
    per_device_train_batch_size: int = 32
    per_device_eval_batch_size: int = 64
    learning_rate: float = 1e-5
    weight_decay: float = 0.01
    warmup_ratio: float = 0.1
    num_train_epochs: int = 20
    early_stopping_patience: int = 3
    early_stopping_threshold: float = 0.0
    seed: int = 42
    device: Optional[str] = None
    synthetic: bool = False                   # tiny synthetic dataset (offline smoke runs)
    max_train_samples: int = 0


def entity_f1(true_tags: Sequence[Sequence[str]], pred_tags: Sequence[Sequence[str]]) -> Dict[str, float]:
    """Entity-level precision / recall / F1 over IOB2 tag sequences (what seqeval's `f1_score` computes)."""

    def spans(tags):
        out, start, kind = set(), None, None
        for i, t in enumerate(list(tags) + ["O"]):
            if t == "O" or t.startswith("B-") or (t.startswith("I-") and t[2:] != kind):
                if kind is not None:
                    out.add((start, i, kind))
                start, kind = (i, t[2:]) if t != "O" else (None, None)
        return out

    tp = fp = fn = 0
    for t, p in zip(true_tags, pred_tags):
        st, sp = spans(t), spans(p)
        tp += len(st & sp)
        fp += len(sp - st)
        fn += len(st - sp)
    prec = tp / (tp + fp) if tp + fp else 0.0
    rec = tp / (tp + fn) if tp + fn else 0.0
    f1 = 2 * prec * rec / (prec + rec) if prec + rec else 0.0
    return {"precision": prec, "recall": rec, "f1": f1}


def run_finetuning(model: torch.nn.Module, train_batches: Callable[[], Iterable[dict]],
                   eval_fn: Callable[[torch.nn.Module], Dict[str, float]], args: FinetuneArguments,
                   metric_for_best_model: str, steps_per_epoch: int) -> Dict[str, float]:
    device = torch.device(args.device or ("cuda" if torch.cuda.is_available() else "cpu"))
    model.to(device)
    no_decay = ("bias", "LayerNorm.weight", "layer_norm.weight")
    groups = [{"params": [p for n, p in model.named_parameters() if not any(k in n for k in no_decay)],
               "weight_decay": args.weight_decay},
              {"params": [p for n, p in model.named_parameters() if any(k in n for k in no_decay)],
               "weight_decay": 0.0}]
    opt = torch.optim.AdamW(groups, lr=args.learning_rate)
    total = max(1, steps_per_epoch * args.num_train_epochs)
    sched = get_linear_schedule_with_warmup(opt, int(args.warmup_ratio * total), total)
    best, best_state, bad_epochs = None, None, 0
    for epoch in range(args.num_train_epochs):
        model.train()
        for batch in train_batches():
            batch = {k: v.to(device) for k, v in batch.items()}
            loss = model(**batch).loss
            loss.backward()
            torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
            opt.step()
            sched.step()
            opt.zero_grad(set_to_none=True)
        model.eval()
        with torch.no_grad():
            metrics = eval_fn(model)
        score = metrics[metric_for_best_model]
        logger.info(f"epoch {epoch}: {metrics}")
        if best is None or score > best + args.early_stopping_threshold:
            best, best_state, bad_epochs = score, copy.deepcopy(model.state_dict()), 0
        else:
            bad_epochs += 1
            if bad_epochs >= args.early_stopping_patience:
                logger.info("early stopping")
                break
    if best_state is not None:
        model.load_state_dict(best_state)
    return {metric_for_best_model: best if best is not None else 0.0}
