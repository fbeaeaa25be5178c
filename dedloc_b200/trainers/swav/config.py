"""Hydra-style configuration of the SwAV trainer without hydra (not installable here).

    run_distributed_engines.py config=pretrain/swav/swav_1node_resnet_submit \
        config.DATA.TRAIN.BATCHSIZE_PER_REPLICA=64 +config.OPTIMIZER.lr=2.4 ...

`config=<name>` selects a YAML under trainers/swav/configs (or a path), `config.A.B=v` overrides an existing key,
`+config.A.B=v` adds a new one (both accepted for either case, like hydra with struct mode off), values are
parsed as YAML scalars / lists, `hydra.*` keys are ignored. The result is an `AttrDict` with the key names of
vissl/config/defaults.yaml; `infer_and_assert_config` derives the loss parameters the way
vissl/utils/hydra_config.py:301-324,352-440 does (num_prototypes / embedding_dim from the head, num_crops from the
multicrop transform, local queue length from the world size)."""
from __future__ import annotations

import copy
import os
from typing import Any, Dict, List, Sequence

import yaml

CONFIG_ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "configs")


class AttrDict(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    @classmethod
    def wrap(cls, obj):
        if isinstance(obj, dict):
            return cls({k: cls.wrap(v) for k, v in obj.items()})
        if isinstance(obj, list):
            return [cls.wrap(v) for v in obj]
        return obj


DEFAULTS: Dict[str, Any] = {
    "VERBOSE": False, "LOG_FREQUENCY": 10, "TEST_ONLY": False, "TEST_MODEL": False, "SEED_VALUE": 0,
    "MULTI_PROCESSING_METHOD": "fork",
    "HOOKS": {"LOG_GPU_STATS": False, "MEMORY_SUMMARY": {"PRINT_MEMORY_SUMMARY": False, "LOG_ITERATION_NUM": 0},
              "PERF_STATS": {"MONITOR_PERF_STATS": False, "ROLLING_BTIME_FREQ": -1},
              "MODEL_COMPLEXITY": {"COMPUTE_COMPLEXITY": False, "INPUT_SHAPE": [3, 224, 224]},
              "PROFILER": {"PROFILE_FORWARD": False},
              "TENSORBOARD_SETUP": {"USE_TENSORBOARD": False, "LOG_DIR": ".", "EXPERIMENT_LOG_DIR": "tensorboard",
                                    "FLUSH_EVERY_N_MIN": 5, "LOG_PARAMS": True, "LOG_PARAMS_GRADIENTS": True,
                                    "LOG_PARAMS_EVERY_N_ITERS": 310}},
    "DATA": {"NUM_DATALOADER_WORKERS": 4,
             "TRAIN": {"DATA_SOURCES": ["synthetic"], "DATASET_NAMES": ["synthetic"], "DATA_PATHS": [],
                       "BATCHSIZE_PER_REPLICA": 32, "LABEL_TYPE": "sample_index", "TRANSFORMS": [],
                       "COLLATE_FUNCTION": "default_collate", "DROP_LAST": True, "DATA_LIMIT": -1,
                       "MMAP_MODE": True, "COPY_TO_LOCAL_DISK": False, "COPY_DESTINATION_DIR": "",
                       # mixed into the shuffle seed of this peer; empty = something unique to the process
                       "PEER_SHUFFLE_SALT": ""}},
    "TRAINER": {"TRAIN_STEP_NAME": "standard_train_step", "TASK_NAME": "self_supervision_task"},
    "METERS": {"name": ""},
    "MODEL": {"TRUNK": {"NAME": "resnet", "TRUNK_PARAMS": {"RESNETS": {"DEPTH": 50, "WIDTH_MULTIPLIER": 1,
                                                                       "ZERO_INIT_RESIDUAL": False}}},
              "HEAD": {"PARAMS": []}, "TEMP_FROZEN_PARAMS_ITER_MAP": [],
              "SYNC_BN_CONFIG": {"CONVERT_BN_TO_SYNC_BN": False, "SYNC_BN_TYPE": "pytorch", "GROUP_SIZE": -1},
              "AMP_PARAMS": {"USE_AMP": False, "AMP_ARGS": {"opt_level": "O1"}},
              "ACTIVATION_CHECKPOINTING": {"USE_ACTIVATION_CHECKPOINTING": False,
                                           "NUM_ACTIVATION_CHECKPOINTING_SPLITS": 2},
              "SINGLE_PASS_EVERY_CROP": False},
    "LOSS": {"name": "swav_loss",
             "swav_loss": {"temperature": 0.1, "use_double_precision": False, "normalize_last_layer": True,
                           "num_iters": 3, "epsilon": 0.05, "crops_for_assign": [0, 1], "temp_hard_assignment_iters": 0,
                           "output_dir": ".", "queue": {"queue_length": 0, "start_iter": 0}}},
    "OPTIMIZER": {"name": "sgd", "use_larc": False, "larc_config": {"clip": False, "trust_coefficient": 0.001,
                                                                   "eps": 1e-8},
                  "weight_decay": 1e-4, "momentum": 0.9, "nesterov": False, "num_epochs": 1,
                  "regularize_bn": False, "regularize_bias": True, "head_optimizer_params": {},
                  "param_schedulers": {"lr": {"name": "constant", "value": 0.1, "update_interval": "step",
                                              "auto_lr_scaling": {"auto_scale": False, "base_value": 0.1,
                                                                  "base_lr_batch_size": 256}}}},
    "DISTRIBUTED": {"BACKEND": "nccl", "NUM_NODES": 1, "NUM_PROC_PER_NODE": 1, "RUN_ID": "auto",
                    "INIT_METHOD": "tcp", "NCCL_DEBUG": False},
    "MACHINE": {"DEVICE": "gpu"},
    "CHECKPOINT": {"DIR": ".", "AUTO_RESUME": True, "CHECKPOINT_FREQUENCY": 1, "CHECKPOINT_ITER_FREQUENCY": -1,
                   "OVERWRITE_EXISTING": True},
}


def _merge(dst: dict, src: dict) -> dict:
    for k, v in src.items():
        if isinstance(v, dict) and isinstance(dst.get(k), dict):
            _merge(dst[k], v)
        else:
            dst[k] = copy.deepcopy(v)
    return dst


def _set_path(cfg: dict, dotted: str, value: Any) -> None:
    keys = dotted.split(".")
    node = cfg
    for k in keys[:-1]:
        if k not in node or not isinstance(node[k], dict):
            node[k] = {}
        node = node[k]
    node[keys[-1]] = value


def resolve_config_file(name: str) -> str:
    for cand in (name, name + ".yaml", os.path.join(CONFIG_ROOT, name), os.path.join(CONFIG_ROOT, name + ".yaml")):
        if os.path.isfile(cand):
            return cand
    raise FileNotFoundError(f"config {name!r} not found (looked under {CONFIG_ROOT})")


def compose(overrides: Sequence[str]) -> AttrDict:
    """hydra.compose("defaults", overrides) equivalent."""
    cfg = copy.deepcopy(DEFAULTS)
    rest: List[str] = []
    for ov in overrides:
        if ov.startswith("config="):
            with open(resolve_config_file(ov.split("=", 1)[1])) as f:
                loaded = yaml.safe_load(f) or {}
            _merge(cfg, loaded.get("config", loaded))
        else:
            rest.append(ov)
    for ov in rest:
        key, _, raw = ov.partition("=")
        key = key.lstrip("+")
        if key.startswith("hydra."):
            continue
        if not key.startswith("config."):
            raise ValueError(f"unsupported override {ov!r}: expected config=<file> or [+]config.<KEY>=<value>")
        _set_path(cfg, key[len("config."):], yaml.safe_load(raw))
    return infer_and_assert_config(AttrDict.wrap(cfg))


def infer_and_assert_config(cfg: AttrDict) -> AttrDict:
    if cfg.LOSS.name == "swav_loss":
        loss = cfg.LOSS.swav_loss
        head = [p for p in cfg.MODEL.HEAD.PARAMS if p[0] == "swav_head"]
        assert head, "swav_loss needs a swav_head in MODEL.HEAD.PARAMS"
        loss["num_prototypes"] = list(head[0][1]["num_clusters"])
        loss["embedding_dim"] = head[0][1]["dims"][-1]
        crops = [t for t in cfg.DATA.TRAIN.TRANSFORMS if t.get("name") == "ImgPilToMultiCrop"]
        loss["num_crops"] = int(crops[0]["total_num_crops"]) if crops else int(loss.get("num_crops", 2))
        world = cfg.DISTRIBUTED.NUM_NODES * cfg.DISTRIBUTED.NUM_PROC_PER_NODE
        assert loss.queue.queue_length % world == 0, "queue length must be divisible by the world size"
        loss.queue["local_queue_length"] = loss.queue.queue_length // world
        bs = cfg.DATA.TRAIN.BATCHSIZE_PER_REPLICA
        if loss.queue.local_queue_length > 0:
            assert loss.queue.local_queue_length >= bs, "the local queue must hold at least one batch"
    return cfg
