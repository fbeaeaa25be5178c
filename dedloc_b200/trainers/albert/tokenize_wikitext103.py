#!/usr/bin/env python
"""Tokenise WikiText-103 into ALBERT pre-training instances (A6; parity with
/root/reference/albert/tokenize_wikitext103.py:1-104): split documents into sentences, pack them into segment
A / segment B pairs of at most 512 tokens, swap the segments with probability 0.5 (`sentence_order_label`), keep
`special_tokens_mask`, and save dataset + tokenizer under ./data/. Sentence splitting uses nltk when it is
installed and a punctuation regex otherwise (nltk is not in this image)."""
from __future__ import annotations

import random
import re
from collections import defaultdict
from functools import partial
from multiprocessing import cpu_count
from typing import Dict, List

COLUMN_NAMES = ("attention_mask", "input_ids", "sentence_order_label", "special_tokens_mask", "token_type_ids")
_SENT_RE = re.compile(r"(?<=[.!?])\s+(?=[A-Z0-9\"'(\[])")


def sent_tokenize(text: str) -> List[str]:
    try:
        import nltk

        return nltk.sent_tokenize(text)
    except Exception:
        return [s for s in _SENT_RE.split(text.strip()) if s]


def create_instances_from_document(tokenizer, document: str, max_seq_length: int, rng=random) -> List[Dict]:
    """ALBERT SOP instances from one document (tokenize_wikitext103.py:13-72)."""
    instances = []
    current_chunk: List[List[int]] = []
    current_length = 0
    segmented = [s for s in sent_tokenize(document) if s.strip()]
    tokenized = [tokenizer(s, add_special_tokens=False, return_attention_mask=False)["input_ids"] for s in segmented]
    max_tokens = max_seq_length - 3  # [CLS] A [SEP] B [SEP]
    for i, sent in enumerate(tokenized):
        current_chunk.append(sent)
        current_length += len(sent)
        if i == len(tokenized) - 1 or current_length >= max_tokens:
            if len(current_chunk) > 1:
                a_end = rng.randint(1, len(current_chunk) - 1)  # how many sentences go into segment A
                tokens_a = [t for s in current_chunk[:a_end] for t in s]
                tokens_b = [t for s in current_chunk[a_end:] for t in s]
                if rng.random() < 0.5:
                    is_swapped = True
                    tokens_a, tokens_b = tokens_b, tokens_a
                else:
                    is_swapped = False
                # truncate the longer segment from a random side until the pair fits
                while len(tokens_a) + len(tokens_b) > max_tokens:
                    longer = tokens_a if len(tokens_a) > len(tokens_b) else tokens_b
                    if rng.random() < 0.5:
                        del longer[0]
                    else:
                        longer.pop()
                if tokens_a and tokens_b:
                    inst = tokenizer.prepare_for_model(tokens_a, tokens_b, add_special_tokens=True,
                                                       return_token_type_ids=True, return_attention_mask=False,
                                                       return_special_tokens_mask=True, truncation=False)
                    assert len(inst["input_ids"]) <= max_seq_length
                    inst["sentence_order_label"] = 1 if is_swapped else 0
                    instances.append(dict(inst))
            current_chunk, current_length = [], 0
    return instances


def tokenize_function(tokenizer, examples, max_seq_length: int = 512):
    texts = [t for t in examples["text"] if len(t) > 0 and not t.isspace()]
    new_examples = defaultdict(list)
    for text in texts:
        for inst in create_instances_from_document(tokenizer, text, max_seq_length=max_seq_length):
            for key, value in inst.items():
                new_examples[key].append(value)
    return new_examples


def main():
    from datasets import load_dataset
    from transformers import AlbertTokenizerFast

    random.seed(0)
    tokenizer = AlbertTokenizerFast.from_pretrained("albert-large-v2")
    wikitext = load_dataset("wikitext", "wikitext-103-v1", cache_dir="./data/cache")
    tokenized = wikitext.map(partial(tokenize_function, tokenizer), batched=True, num_proc=cpu_count(),
                             remove_columns=["text"])
    tokenized.save_to_disk("./data/albert_tokenized_wikitext")
    tokenizer.save_pretrained("./data/tokenizer")


if __name__ == "__main__":
    main()
