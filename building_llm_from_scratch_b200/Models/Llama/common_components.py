"""Drop-in for Models/Llama/common_components.py (same names and signatures)."""
from __future__ import annotations

import torch
import torch.nn as nn

from ... import ops
from ... import kernels as K
from ...lora import B200Linear


def compute_rope(x, cos, sin):
    """common_components.py:6-35. x: [B, H, T, hd] -> rotated copy (stand-alone form; inside the
    attention module RoPE is applied by the fused attention op on the [B*T, H*hd] layout)."""
    B, H, T, hd = x.shape
    if x.dtype == torch.float32:      # fp32 parity mode
        from ... import f32
        flat = x.transpose(1, 2).reshape(B * T, H * hd).contiguous()
        return f32.rope(flat, cos, sin, T, H, hd).view(B, T, H, hd).transpose(1, 2)
    flat = x.transpose(1, 2).reshape(B * T, H * hd).contiguous()
    out = K.rope(flat, cos.to(torch.bfloat16), sin.to(torch.bfloat16), B, T, H, hd)
    return out.view(B, T, H, hd).transpose(1, 2).to(dtype=x.dtype)


def rescale_theta(theta_old, context_length_old, context_length_new):
    """common_components.py:38-51."""
    return theta_old * (context_length_new / context_length_old)


class RMSNorm(nn.Module):
    """common_components.py:54-70 (fp32 weight, eps 1e-5)."""

    def __init__(self, emb_dim, eps=1e-5):
        super().__init__()
        self.eps = eps
        self.emb_dim = emb_dim
        self.weight = nn.Parameter(torch.ones(emb_dim).float())

    def forward(self, x):
        return ops.rmsnorm(x, self.weight, self.eps)


class SiLU(nn.Module):
    """common_components.py:78-88 — kept for interface parity; FeedForward fuses it with the gate."""

    def forward(self, x):
        return ops.swiglu(x, torch.ones_like(x))


class FeedForward(nn.Module):
    """common_components.py:95-124: fc3(silu(fc1 x) * fc2 x), all bias-free."""

    def __init__(self, cfg):
        super().__init__()
        self.fc1 = B200Linear(cfg["emb_dim"], cfg["hidden_dim"], dtype=cfg["dtype"], bias=False)
        self.fc2 = B200Linear(cfg["emb_dim"], cfg["hidden_dim"], dtype=cfg["dtype"], bias=False)
        self.fc3 = B200Linear(cfg["hidden_dim"], cfg["emb_dim"], dtype=cfg["dtype"], bias=False)
        self.silu = SiLU()

    def forward(self, x, residual=None):
        fused = ops.fc12_swiglu(x, self.fc1, self.fc2)       # (a | b) and h from one GEMM epilogue
        if fused is not None:
            out = ops.swiglu_down_module(fused[0], self.fc3, residual, h=fused[1].detach())
            if out is not None:
                return out
        ab = ops.linears_packed(x, [self.fc1, self.fc2])
        if ab is not None:
            fused = ops.swiglu_down_module(ab, self.fc3, residual)
            if fused is not None:
                return fused
            h = ops.swiglu_packed(ab)
        else:
            a, b = ops.linears_of_same_input(x, [self.fc1, self.fc2])
            h = ops.swiglu(a, b)
        return self.fc3(h, residual)
