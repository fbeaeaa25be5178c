#!/usr/bin/env python
"""Train the sahajBERT tokenizer (S5; /root/reference/sahajbert/tokenizer/tokenizer_training_custom.py:1-31):
Unigram with 31 995 pieces + the 5 specials <pad> <unk> [CLS] [SEP] [MASK] = 32 000, on OSCAR-bn (or any local
text file with --text_file when there is no network). Writes data/tokenizer.json."""
import argparse
import os

from .tokenizer_model import SentencePieceUnigramCustomizedTokenizer

VOCAB_SIZE = 31_995
SPECIAL_TOKENS = ["<pad>", "<unk>", "[CLS]", "[SEP]", "[MASK]"]


def batch_iterator(dataset, input_sentence_size=None, batch_length: int = 100):
    n = len(dataset) if input_sentence_size is None else input_sentence_size
    for i in range(0, n, batch_length):
        yield dataset[i:i + batch_length]["text"]


def train_tokenizer(texts=None, vocab_size: int = VOCAB_SIZE, out_path: str = "data/tokenizer.json",
                    show_progress: bool = True) -> SentencePieceUnigramCustomizedTokenizer:
    tokenizer = SentencePieceUnigramCustomizedTokenizer()
    if texts is None:
        import datasets

        dataset = datasets.load_dataset("oscar", name="unshuffled_deduplicated_bn", split="train")
        iterator = batch_iterator(dataset)
    else:
        iterator = ([t] for t in texts)
    # `vocab_size` counts the learned pieces; the trainer's budget includes the special tokens (tokenizers >= 0.10)
    tokenizer.train_from_iterator(iterator=iterator, vocab_size=vocab_size + len(SPECIAL_TOKENS),
                                  show_progress=show_progress, special_tokens=SPECIAL_TOKENS)
    if out_path:
        os.makedirs(os.path.dirname(out_path) or ".", exist_ok=True)
        tokenizer.save(out_path)
    return tokenizer


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--text_file", default=None, help="local corpus, one document per line (offline training)")
    ap.add_argument("--vocab_size", type=int, default=VOCAB_SIZE)
    ap.add_argument("--out", default="data/tokenizer.json")
    a = ap.parse_args()
    lines = open(a.text_file, encoding="utf-8").read().splitlines() if a.text_file else None
    train_tokenizer(lines, a.vocab_size, a.out)
