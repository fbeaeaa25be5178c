"""Small elementwise helpers used by the autograd wrappers."""
from __future__ import annotations

import math

import torch

_K = math.sqrt(2.0 / math.pi)


def gelu_new(x: torch.Tensor) -> torch.Tensor:
    return 0.5 * x * (1.0 + torch.tanh(_K * (x + 0.044715 * x * x * x)))


def gelu_new_grad(x: torch.Tensor) -> torch.Tensor:
    x2 = x * x
    t = torch.tanh(_K * (x + 0.044715 * x * x2))
    return 0.5 * (1.0 + t) + 0.5 * x * (1.0 - t * t) * _K * (1.0 + 3 * 0.044715 * x2)


def dgelu_mul(dy: torch.Tensor, pre: torch.Tensor) -> torch.Tensor:
    """dy * gelu_new'(pre), computed in fp32, returned in dy's dtype."""
    return (dy.float() * gelu_new_grad(pre.float())).to(dy.dtype)
