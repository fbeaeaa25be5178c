"""`AlbertBengaliTokenizerFast` (S4): the fast tokenizer class of sahajBERT, parity with
/root/reference/sahajbert/tokenization_albert_bengali_fast.py:19-103. A thin PreTrainedTokenizerFast over a
`tokenizer.json` (Unigram, vocab 32 000) whose special tokens are [CLS] [SEP] <unk> <pad> [MASK]; sequences are
capped at 512 tokens. Build the json with tokenizer/tokenizer_training_custom.py."""
import os
from typing import Optional, Tuple

from transformers import AddedToken, PreTrainedTokenizerFast

VOCAB_FILES_NAMES = {"tokenizer_file": "tokenizer.json"}
MODEL_MAX_LENGTH = 512


class AlbertBengaliTokenizerFast(PreTrainedTokenizerFast):
    vocab_files_names = VOCAB_FILES_NAMES
    model_input_names = ["input_ids", "token_type_ids", "attention_mask"]

    def __init__(self, tokenizer_file=None, bos_token="[CLS]", eos_token="[SEP]", unk_token="<unk>",
                 sep_token="[SEP]", pad_token="<pad>", cls_token="[CLS]", mask_token="[MASK]",
                 padding_side="right", model_max_length=MODEL_MAX_LENGTH, **kwargs):
        if tokenizer_file is None and kwargs.get("tokenizer_object") is None:
            # transformers >= 5 hands a class with its own __init__ only the vocabulary (and expects it to rebuild the
            # normalizer / pre-tokenizer); the pipeline of this tokenizer lives in tokenizer.json, so load that file
            path = kwargs.get("name_or_path")
            candidate = os.path.join(path, VOCAB_FILES_NAMES["tokenizer_file"]) if path and os.path.isdir(path) else None
            if candidate is not None and os.path.isfile(candidate):
                tokenizer_file = candidate
                kwargs.pop("vocab", None)
            elif kwargs.get("vocab") is not None:
                from .tokenizer.tokenizer_model import SentencePieceUnigramCustomizedTokenizer
                from tokenizers.models import Unigram

                backend = SentencePieceUnigramCustomizedTokenizer()._tokenizer
                backend.model = Unigram(vocab=[tuple(v) for v in kwargs.pop("vocab")], unk_id=1)
                kwargs["tokenizer_object"] = backend
        if isinstance(mask_token, str):
            # the mask token behaves like an ordinary word: the space in front of it is kept
            mask_token = AddedToken(mask_token, lstrip=False, rstrip=False, normalized=False)
        super().__init__(tokenizer_file=tokenizer_file, bos_token=bos_token, eos_token=eos_token,
                         unk_token=unk_token, sep_token=sep_token, cls_token=cls_token, pad_token=pad_token,
                         mask_token=mask_token, padding_side=padding_side, model_max_length=model_max_length,
                         **kwargs)

    def save_vocabulary(self, save_directory: str, filename_prefix: Optional[str] = None) -> Tuple[str]:
        files = self._tokenizer.model.save(save_directory, filename_prefix)
        return tuple(files)
