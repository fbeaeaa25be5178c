"""SentencePiece-style Unigram tokenizer for Bengali (S5; parity with
/root/reference/sahajbert/tokenizer/tokenizer_model.py:9-87): NMT + NFKC normalisation, runs of spaces collapsed,
look-alike danda / double-danda / colon characters mapped to their Bengali code points, lower-casing; Metaspace +
per-digit + punctuation pre-tokenisation; BERT-style template with [CLS]=2 and [SEP]=3."""
from typing import Iterator, List, Union

from tokenizers import AddedToken, Regex, Tokenizer, decoders, normalizers, pre_tokenizers, trainers
from tokenizers.implementations.base_tokenizer import BaseTokenizer
from tokenizers.models import Unigram
from tokenizers.processors import TemplateProcessing

DANDA, DOUBLE_DANDA, VISARGA = "।", "॥", "ঃ"
BENGALI_NORMALISATION = [
    ("৤", DANDA), ("৥", DOUBLE_DANDA),       # Bengali-block look-alikes of the Devanagari dandas
    ("|", DANDA), ("৷", DANDA),              # ASCII pipe and the currency numerator four used as danda
]


def _metaspace(replacement: str, add_prefix_space: bool, cls):
    try:
        return cls(replacement=replacement, prepend_scheme="always" if add_prefix_space else "never")
    except TypeError:  # tokenizers < 0.15
        return cls(replacement=replacement, add_prefix_space=add_prefix_space)


class SentencePieceUnigramCustomizedTokenizer(BaseTokenizer):
    def __init__(self, replacement: str = "▁", add_prefix_space: bool = True):
        tokenizer = Tokenizer(Unigram())
        steps = [normalizers.Nmt(), normalizers.NFKC(), normalizers.Replace(Regex(" {2,}"), " ")]
        steps += [normalizers.Replace(src, dst) for src, dst in BENGALI_NORMALISATION]
        steps += [normalizers.Replace(Regex(r"(?<=[ঀ-৿]):"), VISARGA), normalizers.Lowercase()]
        tokenizer.normalizer = normalizers.Sequence(steps)
        tokenizer.pre_tokenizer = pre_tokenizers.Sequence([
            _metaspace(replacement, add_prefix_space, pre_tokenizers.Metaspace),
            pre_tokenizers.Digits(individual_digits=True),
            pre_tokenizers.Punctuation(),
        ])
        tokenizer.decoder = _metaspace(replacement, add_prefix_space, decoders.Metaspace)
        tokenizer.post_processor = TemplateProcessing(
            single="[CLS] $A [SEP]", pair="[CLS] $A [SEP] $B:1 [SEP]:1",
            special_tokens=[("[CLS]", 2), ("[SEP]", 3)])
        super().__init__(tokenizer, {"model": "SentencePieceUnigram", "replacement": replacement,
                                     "add_prefix_space": add_prefix_space})

    def train(self, files: Union[str, List[str]], vocab_size: int = 8000, show_progress: bool = True,
              special_tokens: List[Union[str, AddedToken]] = ()):
        trainer = trainers.UnigramTrainer(vocab_size=vocab_size, special_tokens=list(special_tokens),
                                          show_progress=show_progress, unk_token="<unk>")
        self._tokenizer.train([files] if isinstance(files, str) else files, trainer=trainer)

    def train_from_iterator(self, iterator: Union[Iterator[str], Iterator[Iterator[str]]], vocab_size: int = 8000,
                            show_progress: bool = True, special_tokens: List[Union[str, AddedToken]] = ()):
        trainer = trainers.UnigramTrainer(vocab_size=vocab_size, special_tokens=list(special_tokens),
                                          show_progress=show_progress, unk_token="<unk>")
        self._tokenizer.train_from_iterator(iterator, trainer=trainer)
