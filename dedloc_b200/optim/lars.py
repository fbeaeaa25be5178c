"""LARC (apex semantics, clip=False by default) wrapped around SGD with momentum -- the optimizer built at
/root/reference/swav/ClassyVision/classy_vision/optim/sgd_collaborative.py:135-144.
For every parameter with non-zero ||p|| and ||g||: lr_p = trust * ||p|| / (||g|| + wd ||p|| + eps);
g <- (g + wd p) * lr_p (clip=True: lr_p <- min(lr_p / lr, 1)); the group's own weight decay is zeroed for the
inner SGD step. CPU path and oracle of the fused K2 kernel.
"""
from __future__ import annotations

import torch
from torch.optim import SGD, Optimizer


class LARC(Optimizer):
    def __init__(self, optimizer: Optimizer, trust_coefficient: float = 0.02, clip: bool = True, eps: float = 1e-8):
        self.optim = optimizer
        self.trust_coefficient, self.clip, self.eps = trust_coefficient, clip, eps
        # share the inner optimizer's containers so that this object quacks like a torch Optimizer
        self.param_groups = optimizer.param_groups
        self.state = optimizer.state
        self.defaults = optimizer.defaults

    def __getstate__(self):
        return self.optim.__getstate__()

    def __setstate__(self, state):
        self.optim.__setstate__(state)

    def __repr__(self):
        return f"LARC({self.optim!r})"

    def state_dict(self):
        return self.optim.state_dict()

    def load_state_dict(self, state_dict):
        self.optim.load_state_dict(state_dict)
        self.param_groups = self.optim.param_groups
        self.state = self.optim.state

    def zero_grad(self, set_to_none: bool = True):
        self.optim.zero_grad(set_to_none=set_to_none)

    def add_param_group(self, param_group):
        self.optim.add_param_group(param_group)

    @torch.no_grad()
    def step(self, closure=None):
        weight_decays = []
        for group in self.optim.param_groups:
            wd = group.get("weight_decay", 0)
            weight_decays.append(wd)
            group["weight_decay"] = 0
            for p in group["params"]:
                if p.grad is None:
                    continue
                param_norm, grad_norm = torch.norm(p), torch.norm(p.grad)
                if param_norm != 0 and grad_norm != 0:
                    adaptive_lr = self.trust_coefficient * param_norm / (grad_norm + param_norm * wd + self.eps)
                    if self.clip:
                        adaptive_lr = min(adaptive_lr / group["lr"], 1)
                    p.grad.add_(p, alpha=wd)
                    p.grad.mul_(adaptive_lr)
        loss = self.optim.step(closure) if closure is not None else self.optim.step()
        for group, wd in zip(self.optim.param_groups, weight_decays):
            group["weight_decay"] = wd
        return loss


def build_lars_sgd(params, lr, momentum=0.9, weight_decay=1e-6, nesterov=False, trust_coefficient=0.001,
                   clip=False, eps=1e-8) -> LARC:
    sgd = SGD(params, lr=lr, momentum=momentum, weight_decay=weight_decay, nesterov=nesterov)
    return LARC(sgd, trust_coefficient=trust_coefficient, clip=clip, eps=eps)
