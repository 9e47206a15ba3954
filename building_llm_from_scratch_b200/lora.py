"""Drop-in for the reference's lora.py: same class names, constructor signatures, parameter
names (``linear``, ``lora.A``, ``lora.B``) and initialisation; forward runs the fused tcgen05
GEMM (base product and rank-r update accumulated in one TMEM tile) instead of four ATen ops."""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from . import ops


class B200Linear(nn.Linear):
    """nn.Linear whose forward is the b200llm GEMM. Being an nn.Linear subclass it is found and
    wrapped by replace_linear_with_lora exactly like the reference's plain nn.Linear."""

    def forward(self, x: torch.Tensor, residual: torch.Tensor | None = None) -> torch.Tensor:
        return ops.lora_linear(x, self.weight, self.bias, None, None, 0.0, residual)[0]


class LoRALayer(nn.Module):
    """lora.py:6-26. A: [in, r] kaiming-uniform(a=sqrt(5)); B: [r, out] zeros; scaling = alpha / rank."""

    def __init__(self, in_dim: int, out_dim: int, rank: int, alpha: float, dtype=torch.float32):
        super().__init__()
        self.rank = rank
        self.alpha = alpha
        self.scaling = self.alpha / self.rank
        self.A = nn.Parameter(torch.empty(in_dim, rank, dtype=dtype))
        nn.init.kaiming_uniform_(self.A, a=math.sqrt(5))
        self.B = nn.Parameter(torch.zeros(rank, out_dim, dtype=dtype))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        # stand-alone use (the wrapper below fuses this into the base GEMM): s * (x A) B
        zero_w = torch.zeros(self.B.shape[1], self.A.shape[0], device=x.device, dtype=x.dtype)
        return ops.lora_linear(x, zero_w, None, self.A, self.B, self.scaling, None)[0]


class LinearWithLoRA(nn.Module):
    """lora.py:29-46: output = linear(x) + lora(x)."""

    def __init__(self, linear: nn.Linear, rank: int, alpha: float, dtype=torch.float32):
        super().__init__()
        self.linear = linear
        self.lora = LoRALayer(in_dim=linear.in_features, out_dim=linear.out_features, rank=rank, alpha=alpha,
                              dtype=dtype)

    def forward(self, x: torch.Tensor, residual: torch.Tensor | None = None) -> torch.Tensor:
        return ops.lora_linear(x, self.linear.weight, self.linear.bias, self.lora.A, self.lora.B,
                               self.lora.scaling, residual)[0]


def replace_linear_with_lora(model: nn.Module, rank: int, alpha: float, dtype=torch.float32):
    """lora.py:49-65: recursively wrap every nn.Linear child (the vocab head included)."""
    for name, module in model.named_children():
        if isinstance(module, nn.Linear):
            setattr(model, name, LinearWithLoRA(module, rank, alpha, dtype))
        else:
            replace_linear_with_lora(module, rank, alpha, dtype)


def activate_lora(model: nn.Module, rank: int, alpha: float, dtype=torch.float32) -> nn.Module:
    """build_components.py:117-135 (`activate_lora(args, model, config)` there): freeze every parameter, then wrap
    every nn.Linear; the LoRA A / B tensors are created on the device of the linear they wrap."""
    for p in model.parameters():
        p.requires_grad = False
    dev = next(model.parameters()).device
    with torch.device(dev):
        replace_linear_with_lora(model, rank=rank, alpha=alpha, dtype=dtype)
    return model


def head_operands(head: nn.Module):
    """(weight, A, B, scaling) of a plain or LoRA-wrapped vocab head."""
    if isinstance(head, LinearWithLoRA):
        return head.linear.weight, head.lora.A, head.lora.B, head.lora.scaling
    return head.weight, None, None, 0.0
