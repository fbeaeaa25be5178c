"""Lazy Bengali pre-training dataset (S3; behaviour of /root/reference/sahajbert/dataset_streaming.py:1-139).

Two text streams (Wikipedia-bn and OSCAR-bn) are merged with probabilities (0.23, 0.77), passed through a
shuffle buffer of 10 000 documents seeded PER PEER, split into sentences with a danda-aware splitter, and turned
on the fly into ALBERT sentence-order-prediction instances of at most 512 tokens; iteration never ends.
With network access the streams are Hugging Face streaming datasets; offline, any iterables of
`{"text": ...}` records (e.g. local files) can be passed as `sources`.
"""
from __future__ import annotations

import logging
import random
import re
from collections import defaultdict
from typing import Dict, Iterable, Iterator, List, Optional, Sequence

import torch

logger = logging.getLogger(__name__)

_BN_SENT_END = re.compile(r"(?<=[।॥?!.])\s+")


def bengali_sentence_tokenize(document: str) -> List[str]:
    """Danda-aware sentence splitter (the reference pads every danda with spaces and asks bnlp's NLTK tokenizer;
    bnlp is not installable here, so the split is a regex over danda / double danda / . ? !)."""
    text = document.replace("।", " । ")
    parts = _BN_SENT_END.split(text)
    return [p.strip() for p in parts if p and not p.isspace()]


def create_instances_from_document(tokenizer, document: str, max_seq_length: int, rng=random) -> List[Dict]:
    """Segments A / B are built from whole sentences (never an arbitrary split point, which would make sentence
    ordering trivial); with probability 0.5 they are swapped and labelled 1."""
    instances, chunk, length = [], [], 0
    sentences = bengali_sentence_tokenize(document)
    for i, sent in enumerate(sentences):
        chunk.append(sent)
        length += len(tokenizer.tokenize(sent))
        if i == len(sentences) - 1 or length >= max_seq_length:
            if len(chunk) > 1:
                a_end = rng.randint(1, len(chunk) - 1)
                first, second = chunk[:a_end], chunk[a_end:]
                swapped = rng.random() < 0.5
                if swapped:
                    first, second = second, first
                inst = tokenizer(" ".join(first), " ".join(second), truncation="longest_first",
                                 max_length=max_seq_length, return_special_tokens_mask=True)
                assert len(inst["input_ids"]) <= max_seq_length
                inst = dict(inst)
                inst["sentence_order_label"] = 1 if swapped else 0
                instances.append(inst)
            chunk, length = [], 0
    return instances


def tokenize_function(tokenizer, examples: Dict[str, List[str]], max_seq_length: int = 512, rng=random):
    out = defaultdict(list)
    for text in examples["text"]:
        if not text or text.isspace():
            continue
        for inst in create_instances_from_document(tokenizer, text, max_seq_length, rng):
            for k, v in inst.items():
                out[k].append(v)
    return out


def merge_streams(streams: Sequence[Iterable[dict]], probabilities: Sequence[float], rng: random.Random) -> Iterator[dict]:
    """Sample the next record from stream i with probability p_i; exhausted streams restart (infinite corpus)."""
    iters = [iter(s) for s in streams]
    while True:
        i = rng.choices(range(len(streams)), weights=probabilities, k=1)[0]
        try:
            yield next(iters[i])
        except StopIteration:
            iters[i] = iter(streams[i])
            try:
                yield next(iters[i])
            except StopIteration:
                raise RuntimeError(f"stream {i} is empty")


def shuffle_buffer(stream: Iterator[dict], buffer_size: int, rng: random.Random) -> Iterator[dict]:
    buf: List[dict] = []
    for item in stream:
        if len(buf) < buffer_size:
            buf.append(item)
            continue
        j = rng.randrange(buffer_size)
        yield buf[j]
        buf[j] = item
    rng.shuffle(buf)
    yield from buf


class WrappedIterableDataset(torch.utils.data.IterableDataset):
    """Infinite torch IterableDataset over tokenised instances (restarts the underlying iterable when it ends)."""

    def __init__(self, make_iterable, verbose: bool = True):
        self.make_iterable, self.verbose = make_iterable, verbose

    def __iter__(self):
        started = False
        logger.info("Pre-fetching training samples...")
        while True:
            for sample in self.make_iterable():
                if not started and self.verbose:
                    logger.info("Began iterating minibatches!")
                    started = True
                yield sample


def _hf_streams():
    from datasets import load_dataset

    wiki = load_dataset("lhoestq/wikipedia_bn", split="train", streaming=True)
    oscar = load_dataset("oscar", "unshuffled_deduplicated_bn", split="train", streaming=True)
    return [wiki, oscar]


def make_lazy_wikioscar_dataset(tokenizer, probs: Sequence[float] = (0.23, 0.77), shuffle_buffer_size: int = 10 ** 4,
                                shuffle_seed: Optional[int] = None, preprocessing_batch_size: int = 256,
                                sources: Optional[Sequence[Iterable[dict]]] = None, max_seq_length: int = 512):
    streams = list(sources) if sources is not None else _hf_streams()
    assert len(streams) == len(probs)

    def make_iterable():
        rng = random.Random(shuffle_seed)
        docs = shuffle_buffer(merge_streams(streams, list(probs), rng), shuffle_buffer_size, rng)
        batch: List[str] = []
        for rec in docs:
            batch.append(rec["text"])
            if len(batch) >= preprocessing_batch_size:
                cols = tokenize_function(tokenizer, {"text": batch}, max_seq_length, rng)
                batch = []
                n = len(cols.get("input_ids", []))
                for i in range(n):
                    yield {k: v[i] for k, v in cols.items()}

    return WrappedIterableDataset(make_iterable)


def text_file_stream(path: str) -> Iterable[dict]:
    class _Lines:
        def __iter__(self):
            with open(path, encoding="utf-8") as f:
                for line in f:
                    if line.strip():
                        yield {"text": line.rstrip("\n")}

    return _Lines()
