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


@register_optimizer("sgd_collaborative")
class SGDCollaborative(ClassyOptimizer):
    TARGET_BATCH_SIZE = 32768   # hard-coded in the reference (sgd_collaborative.py:153)

    def __init__(self, larc_config: Dict[str, Any] = None, lr: float = 0, momentum: float = 0.0,
                 weight_decay: float = 0.0, nesterov: bool = False, use_larc: bool = False,
                 dht_initial_peers: List[str] = (), batch_size_for_tracking: int = 200,
                 dht_listen_on_port: str = "1337", averager_listen_on_port: str = "1338",
                 exp_prefix: str = "test_exp", target_group_size: int = 5, max_allowed_epoch_difference: int = 1,
                 total_steps_in_epoch: int = 6405, scheduler_cls=None, averaging_steps_period: int = 5,
                 averaging_time_period: float = 0, report_progress_expiration: int = 60, timeout: float = None,
                 target_batch_size: Optional[int] = None, compression: str = "FLOAT16", dht: Optional[DHT] = None,
                 collab_kwargs: Optional[Dict[str, Any]] = None):
        super().__init__()
        self._lr, self._momentum, self._weight_decay, self._nesterov = lr, momentum, weight_decay, nesterov
        self._use_larc, self._larc_config, self._exp_prefix = use_larc, larc_config, exp_prefix
        port = lambda p: "*" if str(p) in ("any", "*", "0") else str(p)
        self._listen_on_dht = f"127.0.0.1:{port(dht_listen_on_port)}"
        self._listen_on_averager = f"127.0.0.1:{port(averager_listen_on_port)}"
        self._dht_initial_peers = list(dht_initial_peers)
        self._target_group_size, self._batch_size_for_tracking = target_group_size, batch_size_for_tracking
        # accepted for CLI parity; unused by the reference as well
        self._max_allowed_epoch_difference, self._total_steps_in_epoch = max_allowed_epoch_difference, total_steps_in_epoch
        self._averaging_steps_period, self._averaging_time_period = averaging_steps_period, averaging_time_period
        self._report_progress_expiration, self._timeout = report_progress_expiration, timeout
        self._scheduler_cls = scheduler_cls
        self._target_batch_size = target_batch_size or self.TARGET_BATCH_SIZE
        self._compression, self._dht, self._collab_kwargs = compression, dht, dict(collab_kwargs or {})
        self.dht: Optional[DHT] = None

    def prepare(self, param_groups):
        for group in param_groups:
            group["lr"] = self._lr
        assert self._use_larc, "we can't use collab sgd without larc"
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
        self.optimizer = CollaborativeOptimizer(
            opt=inner, scheduler=self._scheduler_cls(inner) if self._scheduler_cls else None, dht=self.dht, verbose=True,
            compression=CompressionType.Value(self._compression), prefix=self._exp_prefix,
            target_group_size=self._target_group_size, target_batch_size=self._target_batch_size,
            batch_size_per_step=self._batch_size_for_tracking, listen_on=self._listen_on_averager,
            averaging_expiration=5.0, metadata_expiration=30, averaging_timeout=30, start=True, **self._collab_kwargs)

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
        config.setdefault("lr", 0)
        config.setdefault("momentum", 0.0)
        config.setdefault("weight_decay", 0.0)
        config.setdefault("nesterov", False)
        config.setdefault("use_larc", False)
        config.setdefault("larc_config", {"clip": True, "eps": 1e-08, "trust_coefficient": 0.02})
        assert isinstance(config["momentum"], float) and 0.0 <= config["momentum"] < 1.0, \
            "Config must contain a 'momentum' in [0, 1) for SGD optimizer"
        assert isinstance(config["nesterov"], bool), "Config must contain a boolean 'nesterov' param for SGD optimizer"
        assert isinstance(config["use_larc"], bool), "Config must contain a boolean 'use_larc' param for SGD optimizer"
        for key, default in (("dht_initial_peers", []), ("batch_size_for_tracking", 200), ("exp_prefix", "test_exp"),
                             ("dht_listen_on_port", "1337"), ("averager_listen_on_port", "1338"),
                             ("target_group_size", 5), ("max_allowed_epoch_difference", 1),
                             ("total_steps_in_epoch", 6405), ("averaging_steps_period", 1),
                             ("averaging_time_period", 0), ("report_progress_expiration", 60), ("timeout", None),
                             ("warmup_epochs", 0), ("max_epochs", 200), ("warmup_start_lr", 0),
                             ("eta_min", 0.006 * 200 / 256)):
            config.setdefault(key, default)
        logger.info(f"creating SGDCollaborative with config: {config}")

        def scheduler_cls(optim):
            return LinearWarmupCosineAnnealingLR(optim, warmup_epochs=config["warmup_epochs"],
                                                 max_epochs=config["max_epochs"],
                                                 warmup_start_lr=config["warmup_start_lr"], eta_min=config["eta_min"])

        config.setdefault("scheduler_cls", scheduler_cls)
        keys = ("larc_config", "lr", "momentum", "weight_decay", "nesterov", "use_larc", "dht_initial_peers",
                "dht_listen_on_port", "averager_listen_on_port", "exp_prefix", "target_group_size",
                "batch_size_for_tracking", "max_allowed_epoch_difference", "total_steps_in_epoch", "scheduler_cls",
                "averaging_steps_period", "averaging_time_period", "report_progress_expiration", "timeout")
        extra = {k: config[k] for k in ("target_batch_size", "compression", "dht", "collab_kwargs") if k in config}
        return cls(**{k: config[k] for k in keys}, **extra)
