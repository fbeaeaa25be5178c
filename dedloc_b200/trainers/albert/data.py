"""Data path of the ALBERT trainer: pre-tokenised dataset from disk + MLM masking collator
(`load_from_disk` + `DataCollatorForLanguageModeling`, /root/reference/albert/run_trainer.py:229-231), yielding
pinned host batches that the training loop copies to the GPU asynchronously."""
from __future__ import annotations

from typing import Dict, Iterator, List, Optional

import torch


class MLMCollator:
    """15 % of the non-special positions become labels: 80 % -> [MASK], 10 % -> random token, 10 % unchanged
    (DataCollatorForLanguageModeling semantics); sequences are padded to `seq_len`."""

    def __init__(self, vocab_size: int, mask_token_id: int, pad_token_id: int = 0, seq_len: int = 512,
                 mlm_probability: float = 0.15, generator: Optional[torch.Generator] = None):
        self.vocab_size, self.mask_token_id, self.pad_token_id = vocab_size, mask_token_id, pad_token_id
        self.seq_len, self.p, self.g = seq_len, mlm_probability, generator

    def __call__(self, examples: List[Dict]) -> Dict[str, torch.Tensor]:
        B, S = len(examples), self.seq_len
        ids = torch.full((B, S), self.pad_token_id, dtype=torch.long)
        tt = torch.zeros((B, S), dtype=torch.long)
        special = torch.ones((B, S), dtype=torch.bool)
        att = torch.zeros((B, S), dtype=torch.long)
        sop = torch.zeros((B,), dtype=torch.long)
        for i, ex in enumerate(examples):
            n = min(len(ex["input_ids"]), S)
            ids[i, :n] = torch.as_tensor(ex["input_ids"][:n])
            tt[i, :n] = torch.as_tensor(ex.get("token_type_ids", [0] * n)[:n])
            special[i, :n] = torch.as_tensor(ex.get("special_tokens_mask", [0] * n)[:n]).bool()
            att[i, :n] = 1
            sop[i] = int(ex.get("sentence_order_label", 0))
        prob = torch.full((B, S), self.p)
        prob.masked_fill_(special, 0.0)
        masked = torch.bernoulli(prob, generator=self.g).bool()
        labels = torch.where(masked, ids, torch.full_like(ids, -100))
        r = torch.rand((B, S), generator=self.g)
        ids = torch.where(masked & (r < 0.8), torch.full_like(ids, self.mask_token_id), ids)
        rnd = torch.randint(0, self.vocab_size, (B, S), generator=self.g)
        ids = torch.where(masked & (r >= 0.8) & (r < 0.9), rnd, ids)
        return {"input_ids": ids, "token_type_ids": tt, "attention_mask": att, "labels": labels,
                "sentence_order_label": sop}


def tokenized_dataset_batches(dataset_path: str, tokenizer_path: str, batch_size: int, seq_len: int, seed: int,
                              num_workers: int = 0) -> Iterator[Dict[str, torch.Tensor]]:
    from datasets import load_from_disk
    from torch.utils.data import DataLoader
    from transformers import AlbertTokenizerFast

    tokenizer = AlbertTokenizerFast.from_pretrained(tokenizer_path)
    train = load_from_disk(dataset_path)["train"]
    g = torch.Generator().manual_seed(seed)
    collator = MLMCollator(len(tokenizer), tokenizer.mask_token_id, tokenizer.pad_token_id or 0, seq_len, generator=None)
    while True:
        loader = DataLoader(train, batch_size=batch_size, shuffle=True, generator=g, collate_fn=collator,
                            num_workers=num_workers, drop_last=True, pin_memory=torch.cuda.is_available())
        for batch in loader:
            yield batch
