"""sahajBERT CLI contract: the ALBERT dataclasses with the three defaults that /root/reference/sahajbert/arguments.py
changes (tokenizer under `tokenizer/data`, cache in `./cache-data`, no dataloader workers because the dataset is a
python generator fed by streaming downloads) and the coordinator's optional checkpoint interval
(/root/reference/sahajbert/run_first_peer.py:45-47)."""
from dataclasses import dataclass, field
from typing import Optional

from ..albert.arguments import (AlbertTrainingArguments as _AlbertTrainingArguments, AveragerArguments,  # noqa: F401
                                BaseTrainingArguments, CollaborationArguments, CollaborativeOptimizerArguments,
                                CoordinatorArguments as _CoordinatorArguments)


@dataclass
class DatasetArguments:
    tokenizer_path: Optional[str] = field(default="tokenizer/data", metadata={"help": "Path to the tokenizer"})
    config_path: Optional[str] = field(
        default="https://s3.amazonaws.com/models.huggingface.co/bert/albert-large-v2-config.json",
        metadata={"help": "Path to the model config"})
    cache_dir: Optional[str] = field(default="./cache-data", metadata={"help": "Path to the cache"})
    synthetic_data: bool = field(default=False, metadata={"help": "Use synthetic tokens instead of wiki+OSCAR-bn"})
    local_text_files: Optional[str] = field(
        default=None, metadata={"help": "Comma-separated 'wiki.txt,oscar.txt' local corpora (no network needed)"})


@dataclass
class AlbertTrainingArguments(_AlbertTrainingArguments):
    dataloader_num_workers: int = 0
    authorize: bool = field(
        default=False,
        metadata={"help": "Log in to the Hugging Face hub (HF_USERNAME / HF_PASSWORD or a prompt), obtain an access "
                          "token for the experiment and sign every averaging / state-sharing RPC with it; peers "
                          "without a token of the same experiment are refused (what the contributor notebook does)"})


@dataclass
class CoordinatorArguments(_CoordinatorArguments):
    save_checkpoint_step_interval: Optional[int] = field(
        default=None, metadata={"help": "Coordinator will load and save state from peers once every that many steps"})
