"""Optimizers whose update is executed INSIDE the fused all-reduce kernel (K1 = LAMB, K2 = LARS+SGD-momentum).

`FusedLamb` / `FusedLARS` look like `torch.optim.Optimizer` objects (param_groups, state_dict with the per-
parameter `exp_avg` / `exp_avg_sq` / `momentum_buffer` layout of torch_optimizer.Lamb / torch.optim.SGD, so the
reference's `optimizer_state.pt` format is preserved -- run_first_peer.py:131-134), but on a GPU they own a
`PeerBuffers` set: all parameters live in ONE chunk-aligned flat vector with an fp32 master copy, a bf16 mirror
that the model's `nn.Parameter`s alias, and the optimizer moments -- all inside the peer's symmetric NVLink
buffer. `CollaborativeOptimizer` drives them; used standalone, `.step()` is a one-member group (no peers).
On CPU they degrade to the reference implementations (optim/lamb.py, optim/lars.py).
"""
from __future__ import annotations

import math
from typing import Any, Dict, List, Optional

import torch
from torch.optim import Optimizer

from .lamb import Lamb, lamb_bias_correction
from .lars import build_lars_sgd


class FusedOptimizerBase(Optimizer):
    MODE: int = 0

    def __init__(self, params, defaults: Dict[str, Any]):
        super().__init__(params, defaults)
        self.buffers = None          # parallel.fused_allreduce.PeerBuffers once materialized
        self.layout = None
        self.flat_params: List[torch.nn.Parameter] = []
        self.grad_flat: Optional[torch.Tensor] = None     # flat gradient buffer in the parameter dtype
        self.main_grad_flat: Optional[torch.Tensor] = None  # fp32 flat buffer for fused-wgrad accumulation
        self.opt_step = 0            # number of optimizer updates applied (LAMB bias correction)
        self._cpu_impl: Optional[Optimizer] = None
        self.tensor_flags: Dict[int, int] = {}            # id(param) -> DEDLOC_TF_* flags

    # ---------------------------------------------------------------- materialisation
    @property
    def is_materialized(self) -> bool:
        return self.buffers is not None

    def materialize(self, device: Optional[torch.device] = None, use_main_grad: bool = True) -> None:
        """Move every parameter into the flat symmetric buffer (GPU only; idempotent)."""
        if self.buffers is not None:
            return
        from .. import lib as native
        from ..parallel.flat import FlatLayout
        from ..parallel.fused_allreduce import PeerBuffers

        native.require_native_on_gpu()
        params = [p for g in self.param_groups for p in g["params"]]
        assert params, "optimizer has no parameters"
        device = device or params[0].device
        assert device.type == "cuda", "materialize() is the CUDA path; CPU uses the reference optimizers"
        names = [getattr(p, "_dedloc_name", f"p{i}") for i, p in enumerate(params)]
        self.layout = FlatLayout([p.shape for p in params], names)
        self.buffers = buf = PeerBuffers(self.layout, device.index if device.index is not None
                                         else torch.cuda.current_device(), with_state=True)
        self.flat_params = params
        dtypes = {p.dtype for p in params}
        assert len(dtypes) == 1 and next(iter(dtypes)) in (torch.bfloat16, torch.float32), \
            "all parameters must share one dtype (bf16 or fp32)"
        pdtype = next(iter(dtypes))
        p32_views, p16_views = self.layout.views(buf.p32), self.layout.views(buf.p16)
        self.grad_flat = torch.zeros(self.layout.total, dtype=pdtype, device=device)
        grad_views = self.layout.views(self.grad_flat)
        if use_main_grad:
            self.main_grad_flat = torch.zeros(self.layout.total, dtype=torch.float32, device=device)
            main_views = self.layout.views(self.main_grad_flat)
        with torch.no_grad():
            for i, p in enumerate(params):
                master = getattr(p, "_dedloc_master", None)       # optional fp32 initial value
                p32_views[i].copy_(master if master is not None else p.detach().float())
                p16_views[i].copy_(p32_views[i])
                p.data = p16_views[i] if pdtype == torch.bfloat16 else p32_views[i]
                p.grad = grad_views[i]
                if use_main_grad:
                    p.main_grad = main_views[i]
        self._fill_tensor_tables()

    def _fill_tensor_tables(self) -> None:
        buf = self.buffers
        wd, flags = [], []
        for g in self.param_groups:
            for p in g["params"]:
                wd.append(float(g.get("weight_decay", 0.0)))
                flags.append(int(self.tensor_flags.get(id(p), 0)))
        buf.t_wd.copy_(torch.tensor(wd, dtype=torch.float32))
        buf.t_flags.copy_(torch.tensor(flags, dtype=torch.int32))

    def refresh_lr_table(self) -> float:
        """Per-tensor lr multipliers relative to group 0 (the kernel takes one scalar lr). Returns base lr."""
        base = float(self.param_groups[0]["lr"])
        scales = []
        for g in self.param_groups:
            s = (float(g["lr"]) / base) if base != 0 else 0.0
            scales.extend([s] * len(g["params"]))
        cur = getattr(self, "_lr_scales", None)
        if cur != scales:
            self.buffers.t_lr_scale.copy_(torch.tensor(scales, dtype=torch.float32))
            self._lr_scales = scales
        return base

    def set_tensor_flag(self, param: torch.nn.Parameter, flag: int, on: bool = True) -> None:
        cur = self.tensor_flags.get(id(param), 0)
        new = (cur | flag) if on else (cur & ~flag)
        if new != cur:
            self.tensor_flags[id(param)] = new
            if self.buffers is not None:
                self._fill_tensor_tables()

    def hyper(self):
        raise NotImplementedError

    # ---------------------------------------------------------------- gradients
    @torch.no_grad()
    def fold_gradients(self, max_grad_norm: Optional[float]) -> None:
        """acc += clip(local gradient); local gradient buffers are zeroed (SURVEY C9 + H3 step 2)."""
        from ..ops import flat_ops

        buf = self.buffers
        if self.main_grad_flat is not None:
            flat_ops.accumulate_clip(self.main_grad_flat, self.grad_flat, None, 0.0, zero_grad=True)
            src = self.main_grad_flat
        else:
            src = self.grad_flat
        if max_grad_norm is not None and max_grad_norm > 0:
            flat_ops.sumsq(src, buf.norm_scratch, buf.normsq)
            flat_ops.accumulate_clip(buf.acc, src, buf.normsq, float(max_grad_norm), zero_grad=True)
        else:
            flat_ops.accumulate_clip(buf.acc, src, None, 0.0, zero_grad=True)

    def zero_grad(self, set_to_none: bool = False):
        if self.buffers is None:
            return super().zero_grad(set_to_none=set_to_none)
        self.grad_flat.zero_()
        if self.main_grad_flat is not None:
            self.main_grad_flat.zero_()

    # ---------------------------------------------------------------- stand-alone step (one-member group)
    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        params = [p for g in self.param_groups for p in g["params"]]
        if params and params[0].device.type != "cuda":
            if self._cpu_impl is None:
                self._cpu_impl = self._build_cpu_impl()
            self._cpu_impl.step()
            self.opt_step += 1
            return loss
        from ..parallel.fused_allreduce import GroupSpec, fused_allreduce
        from ..ops import flat_ops
        from ..utils.compression import CompressionType

        self.materialize()
        buf = self.buffers
        self.fold_gradients(None)
        flat_ops.encode(CompressionType.NONE, buf.acc, 1.0, buf.send, buf.aux)
        self.opt_step += 1
        self._solo_tag = getattr(self, "_solo_tag", 0) + 1
        group = GroupSpec(bases=[buf.symm.ptr], me=0, weights=[1.0], shards=[(0, self.layout.n_chunks)],
                          slots=[0], tag=self._solo_tag)
        fused_allreduce(buf, group, CompressionType.NONE, self.hyper(), xgpu_barriers=False)
        buf.acc.zero_()
        return loss

    def _build_cpu_impl(self) -> Optimizer:
        raise NotImplementedError

    # ---------------------------------------------------------------- checkpoint format
    def _state_names(self) -> List[str]:
        raise NotImplementedError

    def state_dict(self):
        if self.buffers is None:
            return (self._cpu_impl or self._build_cpu_impl_cached()).state_dict()
        buf = self.buffers
        m_views = self.layout.views(buf.m)
        v_views = self.layout.views(buf.v)
        state, idx = {}, 0
        groups = []
        for g in self.param_groups:
            ids = []
            for _ in g["params"]:
                state[idx] = self._export_param_state(m_views[idx], v_views[idx])
                ids.append(idx)
                idx += 1
            groups.append({**{k: v for k, v in g.items() if k != "params"}, "params": ids})
        return {"state": state, "param_groups": groups, "opt_step": self.opt_step}

    def _build_cpu_impl_cached(self):
        if self._cpu_impl is None:
            self._cpu_impl = self._build_cpu_impl()
        return self._cpu_impl

    def load_state_dict(self, sd):
        if self.buffers is None:
            self._build_cpu_impl_cached().load_state_dict({k: v for k, v in sd.items() if k != "opt_step"})
            self.opt_step = int(sd.get("opt_step", 0))
            return
        buf = self.buffers
        m_views, v_views = self.layout.views(buf.m), self.layout.views(buf.v)
        step = 0
        for idx, st in sd["state"].items():
            step = max(step, self._import_param_state(st, m_views[int(idx)], v_views[int(idx)]))
        for g, saved in zip(self.param_groups, sd["param_groups"]):
            for k, v in saved.items():
                if k != "params":
                    g[k] = v
        self.opt_step = int(sd.get("opt_step", step))
        self._fill_tensor_tables()

    def _export_param_state(self, m, v) -> Dict[str, Any]:
        raise NotImplementedError

    def _import_param_state(self, st, m, v) -> int:
        raise NotImplementedError


class FusedLamb(FusedOptimizerBase):
    """LAMB with torch_optimizer.Lamb semantics (albert/run_trainer.py:86-94) executed by kernel K1."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0, clamp_value=10.0,
                 adam=False, debias=False):
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, clamp_value=clamp_value,
                        adam=adam, debias=debias, initial_lr=lr)
        super().__init__(params, defaults)
        self.clamp_value, self.adam, self.debias = clamp_value, adam, debias
        assert not adam, "adam=True (no trust ratio) is not supported by the fused kernel"

    def hyper(self):
        from .. import lib as native
        from ..parallel.fused_allreduce import OptHyper

        g0 = self.param_groups[0]
        b1, b2 = g0["betas"]
        return OptHyper(mode=native.AR_LAMB, lr=self.refresh_lr_table(), beta1=b1, beta2=b2, eps=g0["eps"],
                        bias_correction=lamb_bias_correction(max(self.opt_step, 1), b1, b2, self.debias),
                        clamp_value=self.clamp_value)

    def _build_cpu_impl(self):
        g0 = self.param_groups[0]
        impl = Lamb(self.param_groups, lr=g0["lr"], betas=g0["betas"], eps=g0["eps"],
                    weight_decay=g0["weight_decay"], clamp_value=self.clamp_value, debias=self.debias)
        impl.param_groups = self.param_groups  # share: schedulers mutate these dicts
        return impl

    def _export_param_state(self, m, v):
        return {"step": self.opt_step, "exp_avg": m.clone(), "exp_avg_sq": v.clone()}

    def _import_param_state(self, st, m, v) -> int:
        m.copy_(st["exp_avg"])
        v.copy_(st["exp_avg_sq"])
        return int(st.get("step", 0))


class FusedLARS(FusedOptimizerBase):
    """apex-LARC(clip=False) + SGD-momentum (sgd_collaborative.py:135-144) executed by kernel K2."""

    def __init__(self, params, lr=0.1, momentum=0.9, weight_decay=1e-6, trust_coefficient=0.001, eps=1e-8,
                 nesterov=False):
        assert not nesterov, "nesterov momentum is not used by the reference recipe and not fused"
        defaults = dict(lr=lr, momentum=momentum, weight_decay=weight_decay, trust_coefficient=trust_coefficient,
                        eps=eps, nesterov=nesterov, initial_lr=lr)
        super().__init__(params, defaults)

    def hyper(self):
        from .. import lib as native
        from ..parallel.fused_allreduce import OptHyper

        g0 = self.param_groups[0]
        return OptHyper(mode=native.AR_LARS, lr=self.refresh_lr_table(), momentum=g0["momentum"],
                        trust=g0["trust_coefficient"], eps=g0["eps"])

    def _build_cpu_impl(self):
        g0 = self.param_groups[0]
        impl = build_lars_sgd(self.param_groups, lr=g0["lr"], momentum=g0["momentum"],
                              weight_decay=g0["weight_decay"], trust_coefficient=g0["trust_coefficient"],
                              clip=False, eps=g0["eps"])
        return impl

    def _export_param_state(self, m, v):
        return {"momentum_buffer": m.clone()}

    def _import_param_state(self, st, m, v) -> int:
        if st.get("momentum_buffer") is not None:
            m.copy_(st["momentum_buffer"])
        return 0
