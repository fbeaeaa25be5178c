"""LAMB with the semantics of `torch_optimizer.Lamb` as configured at /root/reference/albert/run_trainer.py:86-94
(SURVEY C11): m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; u = m/(sqrt(v)+eps) + wd p;
r = clamp(||p||, 0, clamp_value)/||u|| (1 if either norm is 0); p -= lr * bc * r * u with
bc = sqrt(1-b2^t)/(1-b1^t) when debias=True. This is the CPU path and the oracle of the fused K1 kernel.
"""
from __future__ import annotations

import math

import torch
from torch.optim import Optimizer


class Lamb(Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0, clamp_value=10.0,
                 adam=False, debias=False):
        if lr <= 0.0:
            raise ValueError(f"Invalid learning rate: {lr}")
        if eps < 0.0:
            raise ValueError(f"Invalid epsilon value: {eps}")
        if not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0:
            raise ValueError(f"Invalid betas: {betas}")
        if weight_decay < 0 or clamp_value < 0:
            raise ValueError("weight_decay and clamp_value must be non-negative")
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        self.clamp_value, self.adam, self.debias = clamp_value, adam, debias
        super().__init__(params, defaults)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                grad = p.grad
                if grad.is_sparse:
                    raise RuntimeError("Lamb does not support sparse gradients")
                state = self.state[p]
                if len(state) == 0:
                    state["step"] = 0
                    state["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    state["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                exp_avg, exp_avg_sq = state["exp_avg"], state["exp_avg_sq"]
                state["step"] += 1
                exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
                if self.debias:
                    bias_correction = math.sqrt(1 - beta2 ** state["step"]) / (1 - beta1 ** state["step"])
                else:
                    bias_correction = 1.0
                step_size = group["lr"] * bias_correction
                weight_norm = torch.norm(p).clamp(0, self.clamp_value)
                adam_step = exp_avg / exp_avg_sq.sqrt().add(group["eps"])
                if group["weight_decay"] != 0:
                    adam_step.add_(p, alpha=group["weight_decay"])
                adam_norm = torch.norm(adam_step)
                trust_ratio = 1.0 if (weight_norm == 0 or adam_norm == 0) else float(weight_norm / adam_norm)
                state["weight_norm"], state["adam_norm"], state["trust_ratio"] = weight_norm, adam_norm, trust_ratio
                if self.adam:
                    trust_ratio = 1.0
                p.add_(adam_step, alpha=-step_size * trust_ratio)
        return loss


def lamb_bias_correction(step: int, beta1: float, beta2: float, debias: bool = True) -> float:
    return math.sqrt(1 - beta2 ** step) / (1 - beta1 ** step) if debias else 1.0
