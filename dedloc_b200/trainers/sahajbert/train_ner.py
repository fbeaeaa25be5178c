#!/usr/bin/env python
"""Named-entity recognition on WikiANN-bn with entity-level F1 (S7; behaviour of
/root/reference/sahajbert/train_ner.py:1-299): word-level IOB2 tags are aligned to the first sub-token of every
word (-100 elsewhere), fine-tuning with early stopping on eval F1."""
from __future__ import annotations

import logging
from typing import Dict, List

import torch

from ...models.albert import AlbertConfig, AlbertForPreTraining, AlbertForTokenClassification
from ...utils.argparse_dataclass import DataclassArgumentParser
from .finetune import FinetuneArguments, entity_f1, run_finetuning

logger = logging.getLogger(__name__)
WIKIANN_LABELS = ["O", "B-PER", "I-PER", "B-ORG", "I-ORG", "B-LOC", "I-LOC"]


def tokenize_and_align_labels(tokenizer, words: List[List[str]], tags: List[List[int]], max_len: int) -> Dict[str, list]:
    enc = tokenizer(words, is_split_into_words=True, truncation=True, max_length=max_len, padding="max_length")
    labels = []
    for i, word_tags in enumerate(tags):
        prev, row = None, []
        for wid in enc.word_ids(batch_index=i):
            row.append(-100 if wid is None or wid == prev else word_tags[wid])
            prev = wid
        labels.append(row)
    enc["labels"] = labels
    return enc


def _synthetic(cfg: AlbertConfig, n: int, seq: int, seed: int):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(5, cfg.vocab_size, (n, seq), generator=g)
    labels = (ids % len(WIKIANN_LABELS)).clone()   # learnable: the tag is a function of the token id
    labels[:, 0] = -100
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
    model = AlbertForTokenClassification(cfg, num_labels=len(WIKIANN_LABELS))
    if pre is not None:
        model.load_encoder_from(pre)
    if args.synthetic:
        train, dev = _synthetic(cfg, 256, 32, 0), _synthetic(cfg, 64, 32, 1)
    else:
        from datasets import load_dataset

        from .run_trainer import load_tokenizer

        tok = load_tokenizer(args)
        raw = load_dataset("wikiann", "bn")

        def prep(split):
            enc = tokenize_and_align_labels(tok, raw[split]["tokens"], raw[split]["ner_tags"], args.max_seq_length)
            return {k: torch.tensor(enc[k]) for k in ("input_ids", "attention_mask", "labels")}

        train, dev = prep("train"), prep("validation")
    if args.max_train_samples:
        train = {k: v[:args.max_train_samples] for k, v in train.items()}
    bs = args.per_device_train_batch_size
    n = train["input_ids"].shape[0]

    def train_batches():
        perm = torch.randperm(n)
        for i in range(0, n - bs + 1, bs):
            idx = perm[i:i + bs]
            yield {k: v[idx] for k, v in train.items()}

    def evaluate(m):
        dev_ = next(m.parameters()).device
        true, pred = [], []
        for i in range(0, dev["input_ids"].shape[0], args.per_device_eval_batch_size):
            b = {k: v[i:i + args.per_device_eval_batch_size].to(dev_) for k, v in dev.items()}
            logits = m(input_ids=b["input_ids"], attention_mask=b["attention_mask"]).prediction_logits
            p = logits.argmax(-1).cpu()
            for row_p, row_t in zip(p, b["labels"].cpu()):
                keep = row_t != -100
                true.append([WIKIANN_LABELS[t] for t in row_t[keep].tolist()])
                pred.append([WIKIANN_LABELS[t] for t in row_p[keep].tolist()])
        return entity_f1(true, pred)

    result = run_finetuning(model, train_batches, evaluate, args, "f1", max(1, n // bs))
    logger.info(f"best eval F1: {result['f1']:.4f}")
    return result


if __name__ == "__main__":
    main()
