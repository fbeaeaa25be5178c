#!/usr/bin/env python
"""News-category classification on IndicGLUE `sna.bn` with accuracy (S7; behaviour of
/root/reference/sahajbert/train_ncc.py:1-253): sequence classification on the pooled [CLS] representation,
early stopping on eval accuracy."""
from __future__ import annotations

import logging

import torch

from ...models.albert import AlbertConfig, AlbertForPreTraining, AlbertForSequenceClassification
from ...utils.argparse_dataclass import DataclassArgumentParser
from .finetune import FinetuneArguments, run_finetuning

logger = logging.getLogger(__name__)
NUM_CLASSES = 6  # kolkata, state, national, sports, entertainment, international


def _synthetic(cfg: AlbertConfig, n: int, seq: int, seed: int):
    g = torch.Generator().manual_seed(seed)
    labels = torch.randint(0, NUM_CLASSES, (n,), generator=g)
    ids = torch.randint(5 + NUM_CLASSES, cfg.vocab_size, (n, seq), generator=g)
    ids[:, 1] = 5 + labels                          # learnable: one token reveals the class
    return {"input_ids": ids, "attention_mask": torch.ones_like(ids), "labels": labels}


def main(argv=None):
    (args,) = DataclassArgumentParser((FinetuneArguments,)).parse_args_into_dataclasses(argv)
    logging.basicConfig(level=logging.INFO)
    torch.manual_seed(args.seed)
    if args.model_path:
        pre = AlbertForPreTraining.from_pretrained(args.model_path)
        cfg = pre.config
    else:
        pre, cfg = None, AlbertConfig.tiny(max_position_embeddings=args.max_seq_length)
    model = AlbertForSequenceClassification(cfg, num_labels=NUM_CLASSES)
    if pre is not None:
        model.load_encoder_from(pre)
    if args.synthetic:
        train, dev = _synthetic(cfg, 256, 32, 0), _synthetic(cfg, 64, 32, 1)
    else:
        from datasets import load_dataset

        from .run_trainer import load_tokenizer

        tok = load_tokenizer(args)
        raw = load_dataset("indic_glue", "sna.bn")

        def prep(split):
            enc = tok(raw[split]["text"], truncation=True, max_length=args.max_seq_length, padding="max_length")
            return {"input_ids": torch.tensor(enc["input_ids"]), "attention_mask": torch.tensor(enc["attention_mask"]),
                    "labels": torch.tensor(raw[split]["label"])}

        train, dev = prep("train"), prep("validation")
    if args.max_train_samples:
        train = {k: v[:args.max_train_samples] for k, v in train.items()}
    bs, n = args.per_device_train_batch_size, train["input_ids"].shape[0]

    def train_batches():
        perm = torch.randperm(n)
        for i in range(0, n - bs + 1, bs):
            idx = perm[i:i + bs]
            yield {k: v[idx] for k, v in train.items()}

    def evaluate(m):
        dev_ = next(m.parameters()).device
        correct = total = 0
        for i in range(0, dev["input_ids"].shape[0], args.per_device_eval_batch_size):
            b = {k: v[i:i + args.per_device_eval_batch_size].to(dev_) for k, v in dev.items()}
            pred = m(input_ids=b["input_ids"], attention_mask=b["attention_mask"]).prediction_logits.argmax(-1)
            correct += int((pred == b["labels"]).sum())
            total += pred.numel()
        return {"accuracy": correct / max(total, 1)}

    result = run_finetuning(model, train_batches, evaluate, args, "accuracy", max(1, n // bs))
    logger.info(f"best eval accuracy: {result['accuracy']:.4f}")
    return result


if __name__ == "__main__":
    main()
