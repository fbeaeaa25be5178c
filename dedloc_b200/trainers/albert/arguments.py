"""CLI contract of the ALBERT experiment: the dataclasses of /root/reference/albert/arguments.py:7-128 and
run_first_peer.py:24-56 with identical field names and defaults. `AlbertTrainingArguments` lists the
HF TrainingArguments fields the reference relies on instead of subclassing transformers' class (its field set
drifted; removed flags such as --fp16_opt_level / --overwrite_output_dir are accepted and ignored)."""
from dataclasses import dataclass, field
from typing import List, Optional


@dataclass
class BaseTrainingArguments:
    experiment_prefix: str = field(
        metadata={"help": "Name of the run; every DHT key of the collaboration is prefixed with it"})
    initial_peers: List[str] = field(
        default_factory=list,
        metadata={"help": "host:port of one or more peers that are already in the collaboration (bootstrap nodes)"})
    dht_listen_on: str = field(
        default="[::]:*", metadata={"help": "host:port the DHT node binds to ('*' picks a free port)"})


@dataclass
class AveragerArguments:
    averaging_expiration: float = field(
        default=5.0, metadata={"help": "How long (s) a forming group stays open for late members before the round starts"})
    averaging_timeout: float = field(
        default=30.0, metadata={"help": "A round that has not finished after this many seconds is abandoned; the peer then steps on its own gradients"})
    listen_on: str = field(
        default="[::]:*",
        metadata={"help": "host:port the averaging RPC server binds to ('*' picks a free port)"})
    min_refresh_period: float = field(
        default=0.5, metadata={"help": "Lower bound (s) on the interval between two reads of the collaboration's progress records"})
    max_refresh_period: float = field(
        default=30, metadata={"help": "Upper bound (s) on that interval"})
    default_refresh_period: float = field(
        default=3, metadata={"help": "Interval (s) used while no progress record has been found yet"})
    expected_drift_peers: float = field(
        default=3, metadata={"help": "Head-room, in peers, for newcomers when the time of the next step is extrapolated"})
    expected_drift_rate: float = field(
        default=0.2, metadata={"help": "The same head-room as a fraction of the current number of peers"})
    performance_ema_alpha: float = field(
        default=0.1, metadata={"help": "Smoothing factor of the samples-per-second estimate this peer reports"})
    target_group_size: int = field(default=256, metadata={"help": "Largest number of peers that average together in one round; more peers split into several groups"})
    metadata_expiration: float = field(
        default=30, metadata={"help": "Lifetime (s) of a progress record: a peer that stops reporting drops out of the count after this"})
    liveness_window: float = field(
        default=3.0, metadata={"help": "A peer whose last progress report is older than this (and older than 4x its usual reporting interval) is no longer waited for when a group forms"})
    emulated_uplink_mbps: float = field(
        default=0.0, metadata={"help": "Socket rounds only: pace everything this peer sends as a link of this many Mbit/s would (0 = as fast as the machine allows); for reproducing heterogeneous-bandwidth experiments on one machine"})
    kernel_timeout: float = field(
        default=5.0, metadata={"help": "How long (s) the NVLink all-reduce kernel waits for a group member that never shows up before the round is abandoned"})


@dataclass
class CollaborativeOptimizerArguments:
    target_batch_size: int = field(
        default=4096,
        metadata={"help": "Samples the whole collaboration gathers between two optimizer steps"})
    client_mode: bool = field(
        default=False,
        metadata={"help": "Never accept inbound connections (peers behind NAT / firewalls); such a peer sends gradients but owns no shard"})
    batch_size_lead: int = field(
        default=0,
        metadata={"help": "Start matchmaking this many samples early so that group formation overlaps the last micro-batches"})
    bandwidth: float = field(
        default=100.0,
        metadata={"help": "Link speed of this peer in Mbit/s; shard sizes of the socket all-reduce are proportional to it"})
    compression: str = field(
        default="FLOAT16", metadata={"help": "Wire codec of the averaged gradients: NONE, FLOAT16, BFLOAT16, MEANSTD_16BIT, UNIFORM_8BIT, QUANTILE_8BIT, BLOCKWISE_FP8"})


@dataclass
class CollaborationArguments(AveragerArguments, CollaborativeOptimizerArguments, BaseTrainingArguments):
    statistics_expiration: float = field(
        default=600, metadata={"help": "Lifetime (s) of the metrics record the coordinator aggregates"})
    endpoint: Optional[str] = field(
        default=None,
        metadata={"help": "Address other peers should dial when it differs from the bound one (proxy / port forwarding)"})


@dataclass
class DatasetArguments:
    dataset_path: Optional[str] = field(
        default="data/albert_tokenized_wikitext", metadata={"help": "Directory written by tokenize_wikitext103.py (datasets.save_to_disk)"})
    tokenizer_path: Optional[str] = field(default="data/tokenizer", metadata={"help": "Directory with the ALBERT tokenizer files"})
    config_path: Optional[str] = field(
        default="https://s3.amazonaws.com/models.huggingface.co/bert/albert-large-v2-config.json",
        metadata={"help": "config.json of the model to train (file, directory or URL)"})
    cache_dir: Optional[str] = field(default="data", metadata={"help": "Download / cache directory"})
    synthetic_data: bool = field(
        default=False, metadata={"help": "Train on synthetic WikiText-shaped tokens (no dataset / tokenizer needed)"})


@dataclass
class AlbertTrainingArguments:
    dataloader_num_workers: int = 4
    per_device_train_batch_size: int = 4
    per_device_eval_batch_size: int = 4
    gradient_accumulation_steps: int = 2
    seq_length: int = 512

    max_steps: int = 1_000_000  # Albert is actually ready after 125000 steps
    learning_rate: float = 0.00176
    warmup_steps: int = 5000
    adam_epsilon: float = 1e-6
    weight_decay: float = 0.01
    max_grad_norm: float = 1.0
    clamp_value: float = 10000.0
    adam_beta1: float = 0.9
    adam_beta2: float = 0.999

    fp16: bool = True          # kept for CLI parity; the B200 build computes in bf16 (no loss scaling needed)
    fp16_opt_level: str = "O2"
    bf16: bool = True
    do_train: bool = True
    do_eval: bool = False

    logging_steps: int = 100
    save_total_limit: int = 2
    save_steps: int = 500

    output_dir: str = "outputs"
    overwrite_output_dir: bool = False
    seed: int = 42
    local_rank: int = -1
    device: Optional[str] = None
    max_local_steps: int = 0   # 0 = unbounded; handy for smoke runs
    use_hf_trainer: bool = False
    no_cuda_graph: bool = False   # run the micro-step eagerly instead of replaying a captured CUDA graph
    report_json: bool = False     # print one JSON line with the throughput of this process when the loop ends


@dataclass
class CoordinatorArguments(BaseTrainingArguments):
    """The coordinator is only one of possibly many bootstrap nodes: pass --initial_peers here too and list several live
    peers when starting trainers, so that joining does not depend on this process staying up."""
    address: Optional[str] = field(
        default=None,
        metadata={"help": "Address under which this coordinator is reachable (public for internet-wide runs, "
                          "127.0.0.1 for one box)"})
    refresh_period: float = field(
        default=30, metadata={"help": "Seconds between two polls of the metrics key"})
    wandb_project: Optional[str] = field(default=None, metadata={"help": "wandb project for the aggregated curves (offline when there is no network)"})
    save_checkpoint_step_interval: int = field(
        default=5, metadata={"help": "Pull a fresh copy of the training state from the peers every this many collaborative steps"})
    model_config_path: str = field(
        default="https://s3.amazonaws.com/models.huggingface.co/bert/albert-large-v2-config.json",
        metadata={"help": "config.json of the model to train (file, directory or URL)"})
    repo_path: Optional[str] = field(
        default=None,
        metadata={"help": "Local clone of the Hugging Face repository that receives model + optimizer_state.pt commits"})
    upload_interval: Optional[float] = field(
        default=None, metadata={"help": "Minimum seconds between two uploads"})
    max_iterations: int = field(default=0, metadata={"help": "Stop after this many refresh periods (0 = forever)"})
