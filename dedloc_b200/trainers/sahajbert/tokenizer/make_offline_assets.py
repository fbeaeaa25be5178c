#!/usr/bin/env python
"""Build `tokenizer/data/` without network access (S5 assets).

The reference ships `sahajbert/tokenizer/data/{tokenizer.json, special_tokens_map.json, tokenizer_config.json}`,
trained on OSCAR-bn (/root/reference/sahajbert/tokenizer/tokenizer_training_custom.py:6-31). OSCAR is not reachable
from an offline box, so this script trains the SAME tokenizer pipeline (tokenizer_model.py: NMT/NFKC/Bengali
punctuation normalisers, Metaspace + Digits + Punctuation, Unigram, BERT template) with the SAME vocabulary budget
(31 995 pieces + 5 specials = 32 000, <unk> id 1) on a deterministic synthetic corpus of Bengali-script pseudo-words:
syllables assembled from the Unicode Bengali block (consonant (+ hasanta + consonant) + vowel sign / independent
vowel, optional anusvara / visarga / candrabindu), word frequencies Zipf-distributed, sentences closed by a danda.

What the assets are good for: every code path that needs a 32 000-entry sahajBERT vocabulary -- the tokenizer class,
`resize_token_embeddings`, the data collator, the streaming dataset, fine-tuning scripts -- runs end to end offline,
and any Bengali text round-trips (the Unigram alphabet covers the whole block, unknown characters map to <unk>).
What they are NOT: a linguistically meaningful segmentation. Re-run `tokenizer_training_custom.py` on real text
(`--text_file`, or OSCAR-bn when online) before pre-training a model somebody will use.

    python -m dedloc_b200.trainers.sahajbert.tokenizer.make_offline_assets [--out DIR] [--words N] [--sentences N]
"""
import argparse
import json
import os
import random

from .tokenizer_training_custom import SPECIAL_TOKENS, VOCAB_SIZE, train_tokenizer

CONSONANTS = [chr(c) for c in range(0x0995, 0x09BA) if chr(c) not in "঩঱঳঴঵"] + ["ড়", "ঢ়", "য়"]
INDEPENDENT_VOWELS = list("অআইঈউঊঋএঐওঔ")
VOWEL_SIGNS = ["", "া", "ি", "ী", "ু", "ূ", "ৃ", "ে", "ৈ", "ো", "ৌ"]
FINALS = ["", "", "", "", "ং", "ঃ", "ঁ"]
HASANTA = "্"
DIGITS = "০১২৩৪৫৬৭৮৯"
PUNCTUATION = [",", ";", "?", "!", "-", "(", ")", "\"", "'", ":"]


def _syllable(rng: random.Random) -> str:
    r = rng.random()
    if r < 0.08:
        return rng.choice(INDEPENDENT_VOWELS) + rng.choice(FINALS)
    onset = rng.choice(CONSONANTS)
    if r < 0.25:   # conjunct
        onset += HASANTA + rng.choice(CONSONANTS)
    return onset + rng.choice(VOWEL_SIGNS) + rng.choice(FINALS)


def synthetic_corpus(n_words: int = 80_000, n_sentences: int = 100_000, seed: int = 0):
    """Yields sentences. The lexicon has `n_words` pseudo-words of 1-5 syllables; word ranks follow a Zipf law."""
    rng = random.Random(seed)
    lexicon = set()
    while len(lexicon) < n_words:
        lexicon.add("".join(_syllable(rng) for _ in range(rng.choice((1, 2, 2, 3, 3, 3, 4, 4, 5)))))
    lexicon = sorted(lexicon)
    rng.shuffle(lexicon)
    weights = [1.0 / (rank + 10) for rank in range(len(lexicon))]
    for _ in range(n_sentences):
        words = rng.choices(lexicon, weights=weights, k=rng.randint(4, 24))
        if rng.random() < 0.15:
            words.insert(rng.randrange(len(words)), "".join(rng.choice(DIGITS) for _ in range(rng.randint(1, 4))))
        if rng.random() < 0.3:
            i = rng.randrange(len(words))
            words[i] = words[i] + rng.choice(PUNCTUATION)
        yield " ".join(words) + (" ।" if rng.random() < 0.9 else " ॥")


def write_assets(out_dir: str, n_words: int = 80_000, n_sentences: int = 100_000, vocab_size: int = VOCAB_SIZE,
                 show_progress: bool = False) -> str:
    os.makedirs(out_dir, exist_ok=True)
    tokenizer = train_tokenizer(synthetic_corpus(n_words, n_sentences), vocab_size=vocab_size,
                                out_path=os.path.join(out_dir, "tokenizer.json"), show_progress=show_progress)
    specials = {"bos_token": "[CLS]", "eos_token": "[SEP]", "unk_token": "<unk>", "sep_token": "[SEP]",
                "pad_token": "<pad>", "cls_token": "[CLS]",
                # lstrip: the space in front of [MASK] belongs to it (the shipped reference assets set it the same way)
                "mask_token": {"content": "[MASK]", "single_word": False, "lstrip": True, "rstrip": False,
                               "normalized": False}}
    with open(os.path.join(out_dir, "special_tokens_map.json"), "w", encoding="utf-8") as f:
        json.dump(specials, f, ensure_ascii=False, indent=1)
    config = {"tokenizer_class": "AlbertBengaliTokenizerFast", "model_max_length": 512, "padding_side": "right",
              "trained_on": "synthetic Bengali-script corpus (make_offline_assets.py); re-train on real text",
              **{k: (v if isinstance(v, str) else v["content"]) for k, v in specials.items()}}
    with open(os.path.join(out_dir, "tokenizer_config.json"), "w", encoding="utf-8") as f:
        json.dump(config, f, ensure_ascii=False, indent=1)
    assert tokenizer.get_vocab_size() <= vocab_size + len(SPECIAL_TOKENS)
    return out_dir


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(os.path.dirname(os.path.abspath(__file__)), "data"))
    ap.add_argument("--words", type=int, default=80_000)
    ap.add_argument("--sentences", type=int, default=100_000)
    a = ap.parse_args()
    print(write_assets(a.out, a.words, a.sentences, show_progress=True))
