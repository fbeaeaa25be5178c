"""ClassyVision-shaped optimizer layer of the SwAV trainer (SURVEY 2.3 `classy_vision/optim`, W2).

* `register_optimizer` / `build_optimizer`: the decorator registry (`optim/__init__.py:114,124`).
* `get_optimizer_param_groups`: regularised vs unregularised (BN / bias) parameter groups
  (vissl/optimizers/optimizer_helper.py:102-299, restricted to what the recipe sets).
* `SGDCollaborative` ("sgd_collaborative"): SGD(momentum) -> LARC -> CollaborativeOptimizer with
  target_batch_size 32768, FLOAT16 compression, averaging_expiration 5, metadata_expiration 30,
  averaging_timeout 30, LinearWarmupCosineAnnealingLR stepped per collaborative step
  (classy_vision/optim/sgd_collaborative.py:86-249). `get_classy_state` deliberately returns nothing: optimizer
  state is never part of VISSL checkpoints, peers resynchronise through `load_state_from_peers`.
  On a GPU the SGD+LARC pair is the fused LARS kernel (optim/fused.FusedLARS) inside the all-reduce.
"""
from __future__ import annotations

import logging
from typing import Any, Callable, Dict, List, Optional

import torch

from ...collab.optimizer import CollaborativeOptimizer
from ...dht import DHT
from ...optim import LinearWarmupCosineAnnealingLR, build_lars_sgd
from ...optim.fused import FusedLARS
from ...utils.compression import CompressionType

logger = logging.getLogger(__name__)
OPTIMIZER_REGISTRY: Dict[str, type] = {}


def register_optimizer(name: str):
    def deco(cls):
        if name in OPTIMIZER_REGISTRY:
            raise ValueError(f"optimizer {name} already registered")
        OPTIMIZER_REGISTRY[name] = cls
        return cls

    return deco


def build_optimizer(config: Dict[str, Any]):
    return OPTIMIZER_REGISTRY[config["name"]].from_config(dict(config))


def get_optimizer_param_groups(model: torch.nn.Module, optimizer_config: Dict[str, Any]) -> List[Dict[str, Any]]:
    wd = float(optimizer_config.get("weight_decay", 0.0))
    reg_bn = bool(optimizer_config.get("regularize_bn", False))
    reg_bias = bool(optimizer_config.get("regularize_bias", True))
    regularized, unregularized = [], []
    for module_name, module in model.named_modules():
        is_bn = isinstance(module, (torch.nn.modules.batchnorm._BatchNorm, torch.nn.LayerNorm, torch.nn.GroupNorm))
        for pname, p in module.named_parameters(recurse=False):
            if not p.requires_grad:
                continue
            p._dedloc_name = f"{module_name}.{pname}" if module_name else pname
            if (is_bn and not reg_bn) or (pname == "bias" and not is_bn and not reg_bias):
                unregularized.append(p)
            else:
                regularized.append(p)
    groups = [{"params": regularized, "weight_decay": wd}]
    if unregularized:
        groups.append({"params": unregularized, "weight_decay": 0.0})
    return groups


class ClassyOptimizer:
    """Minimal ClassyOptimizer contract used by the train step (classy_optimizer.py:47-279)."""

    def __init__(self):
        self.optimizer = None

    def set_param_groups(self, param_groups, **kwargs):
        self.prepare(param_groups)

    def prepare(self, param_groups):
        raise NotImplementedError

    def zero_grad(self):
        self.optimizer.zero_grad()

    def step(self, *args, closure: Optional[Callable] = None, where: float = None):
        self.optimizer.step()

    def on_epoch(self, where: float) -> None:
        pass

    def get_classy_state(self) -> Dict[str, Any]:
        return {"optim": self.optimizer.state_dict()}

    def set_classy_state(self, state: Dict[str, Any]) -> None:
        self.optimizer.load_state_dict(state["optim"])


@register_optimizer("sgd")
class SGD(ClassyOptimizer):
    """Plain SGD (+ optional LARC), the non-collaborative baseline optimizer (classy_vision/optim/sgd.py:41-47)."""

    def __init__(self, lr=0.1, momentum=0.0, weight_decay=0.0, nesterov=False, use_larc=False, larc_config=None):
        super().__init__()
        self._cfg = dict(lr=lr, momentum=momentum, weight_decay=weight_decay, nesterov=nesterov)
        self._use_larc, self._larc = use_larc, larc_config or {"clip": True, "eps": 1e-8, "trust_coefficient": 0.02}

    def prepare(self, param_groups):
        if self._use_larc:
            self.optimizer = build_lars_sgd(param_groups, **self._cfg, trust_coefficient=self._larc["trust_coefficient"],
                                            clip=self._larc["clip"], eps=self._larc["eps"])
        else:
            self.optimizer = torch.optim.SGD(param_groups, **self._cfg)

    @classmethod
    def from_config(cls, config):
        lr = config.get("lr", config.get("param_schedulers", {}).get("lr", {}).get("value", 0.1))
        return cls(lr=lr, momentum=config.get("momentum", 0.0), weight_decay=config.get("weight_decay", 0.0),
                   nesterov=config.get("nesterov", False), use_larc=config.get("use_larc", False),
                   larc_config=config.get("larc_config"))


# The config keys (and their defaults) are the contract with swav_1node_resnet_submit.yaml and the
# `+config.OPTIMIZER.<key>=...` overrides of swav/README.md:17-32; everything else here is this repo's.
#   key -> (default, type check or None, used)
_SGD_COLLAB_SCHEMA: Dict[str, tuple] = {
    # plain SGD + LARC (apex semantics)
    "lr": (0, (int, float), True),
    "momentum": (0.0, float, True),
    "weight_decay": (0.0, (int, float), True),
    "nesterov": (False, bool, True),
    "use_larc": (False, bool, True),
    "larc_config": ({"clip": True, "eps": 1e-08, "trust_coefficient": 0.02}, dict, True),
    # collaboration
    "dht_initial_peers": ([], (list, tuple), True),
    "batch_size_for_tracking": (200, int, True),
    "exp_prefix": ("test_exp", str, True),
    "dht_listen_on_port": ("1337", (str, int), True),
    "averager_listen_on_port": ("1338", (str, int), True),
    "target_group_size": (5, int, True),
    # accepted and stored, as in the reference, but nothing reads them
    "max_allowed_epoch_difference": (1, int, False),
    "total_steps_in_epoch": (6405, int, False),
    "averaging_steps_period": (1, int, False),
    "averaging_time_period": (0, (int, float), False),
    "report_progress_expiration": (60, (int, float), False),
    "timeout": (None, None, False),
    # learning-rate schedule, stepped once per COLLABORATIVE step
    "warmup_epochs": (0, (int, float), True),
    "max_epochs": (200, (int, float), True),
    "warmup_start_lr": (0, (int, float), True),
    "eta_min": (0.006 * 200 / 256, (int, float), True),
}
_SGD_COLLAB_EXTRA = ("target_batch_size", "compression", "dht", "collab_kwargs", "scheduler_cls")


def _listen_address(port) -> str:
    return "127.0.0.1:" + ("*" if str(port) in ("any", "*", "0") else str(port))


@register_optimizer("sgd_collaborative")
class SGDCollaborative(ClassyOptimizer):
    TARGET_BATCH_SIZE = 32768   # hard-coded in the reference (sgd_collaborative.py:153)

    def __init__(self, **options):
        super().__init__()
        unknown = set(options) - set(_SGD_COLLAB_SCHEMA) - set(_SGD_COLLAB_EXTRA)
        if unknown:
            raise TypeError(f"sgd_collaborative: unknown option(s) {sorted(unknown)}")
        opts = {k: (options[k] if k in options else spec[0]) for k, spec in _SGD_COLLAB_SCHEMA.items()}
        self.options = opts
        self._lr, self._momentum = opts["lr"], opts["momentum"]
        self._weight_decay, self._nesterov = opts["weight_decay"], opts["nesterov"]
        self._use_larc, self._larc_config, self._exp_prefix = opts["use_larc"], opts["larc_config"], opts["exp_prefix"]
        self._listen_on_dht = _listen_address(opts["dht_listen_on_port"])
        self._listen_on_averager = _listen_address(opts["averager_listen_on_port"])
        self._dht_initial_peers = list(opts["dht_initial_peers"])
        self._target_group_size, self._batch_size_for_tracking = opts["target_group_size"], opts["batch_size_for_tracking"]
        self._scheduler_cls = options.get("scheduler_cls")
        self._target_batch_size = options.get("target_batch_size") or self.TARGET_BATCH_SIZE
        self._compression = options.get("compression", "FLOAT16")
        self._dht, self._collab_kwargs = options.get("dht"), dict(options.get("collab_kwargs") or {})
        self.dht: Optional[DHT] = None

    def prepare(self, param_groups):
        for group in param_groups:
            group["lr"] = self._lr
        if not self._use_larc:
            raise ValueError("sgd_collaborative runs SGD behind LARC only: set OPTIMIZER.use_larc=True")
        params = [p for g in param_groups for p in g["params"]]
        on_gpu = bool(params) and params[0].device.type == "cuda"
        if on_gpu:
            inner = FusedLARS(param_groups, lr=self._lr, momentum=self._momentum, weight_decay=self._weight_decay,
                              trust_coefficient=self._larc_config["trust_coefficient"], eps=self._larc_config["eps"],
                              nesterov=self._nesterov)
            assert not self._larc_config.get("clip", False), "the fused LARS kernel implements clip=False (the recipe)"
        else:
            inner = build_lars_sgd(param_groups, lr=self._lr, momentum=self._momentum, weight_decay=self._weight_decay,
                                   nesterov=self._nesterov, trust_coefficient=self._larc_config["trust_coefficient"],
                                   clip=self._larc_config.get("clip", False), eps=self._larc_config["eps"])
        self.dht = self._dht or DHT(listen_on=self._listen_on_dht, initial_peers=self._dht_initial_peers, start=True,
                                    listen=True)
        # sgd_collaborative.py:145-157 fixes these; `collab_kwargs` (tests, benchmarks) may override any of them
        options = dict(verbose=True, compression=CompressionType.Value(self._compression), prefix=self._exp_prefix,
                       target_group_size=self._target_group_size, target_batch_size=self._target_batch_size,
                       batch_size_per_step=self._batch_size_for_tracking, listen_on=self._listen_on_averager,
                       averaging_expiration=5.0, metadata_expiration=30, averaging_timeout=30, start=True)
        options.update(self._collab_kwargs)
        self.optimizer = CollaborativeOptimizer(
            opt=inner, scheduler=self._scheduler_cls(inner) if self._scheduler_cls else None, dht=self.dht, **options)

    def get_classy_state(self) -> Dict[str, Any]:
        return None

    def set_classy_state(self, state: Dict[str, Any]) -> None:
        return

    def on_epoch(self, where: float) -> None:
        return

    def step(self, *args, closure: Optional[Callable] = None, where: float = None) -> None:
        self.optimizer.step()

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "SGDCollaborative":
        """Build from the OPTIMIZER section of the config: fill in the defaults of `_SGD_COLLAB_SCHEMA`, check types
        and ranges, and attach the warm-up + cosine schedule unless the caller supplied one."""
        for key, (default, kind, _) in _SGD_COLLAB_SCHEMA.items():
            if key not in config:
                config[key] = dict(default) if isinstance(default, dict) else (list(default) if isinstance(default, list) else default)
            elif kind is not None and not isinstance(config[key], kind):
                wanted = kind.__name__ if isinstance(kind, type) else " or ".join(k.__name__ for k in kind)
                raise TypeError(f"sgd_collaborative: option '{key}' must be {wanted}, got {type(config[key]).__name__}")
        if not 0.0 <= config["momentum"] < 1.0:
            raise ValueError(f"sgd_collaborative: momentum {config['momentum']} is outside [0, 1)")
        logger.info("sgd_collaborative options: %s", {k: config[k] for k in _SGD_COLLAB_SCHEMA})
        options = {k: config[k] for k in _SGD_COLLAB_SCHEMA}
        options.update({k: config[k] for k in _SGD_COLLAB_EXTRA if k in config})
        if "scheduler_cls" not in options:
            schedule = {k: config[k] for k in ("warmup_epochs", "max_epochs", "warmup_start_lr", "eta_min")}
            options["scheduler_cls"] = lambda optim: LinearWarmupCosineAnnealingLR(optim, **schedule)
        return cls(**options)
