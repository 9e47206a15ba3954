"""Drop-in for Models/Llama/Llama3.py: same class names, constructor signatures, buffer and
parameter names (state_dict compatible: ``load_state_dict(reference.state_dict())`` works).
The tokenizer class is out of scope (needs a downloaded tokenizer.model)."""
from __future__ import annotations

import torch
import torch.nn as nn
from torch.utils.checkpoint import checkpoint_sequential

from ... import ops
from ...lora import B200Linear, head_operands
from .common_components import FeedForward, RMSNorm


def precompute_rope_params(head_dim, theta_base=10_000, context_length=4096, freq_config=None):
    """Llama3.py:79-104 — host-side, once per model; fp32 math, incl. Llama-3.1/3.2 smoothing."""
    assert head_dim % 2 == 0, "Embedding dimension must be even"
    inv_freq = 1.0 / (theta_base ** (torch.arange(0, head_dim, 2).float() / head_dim))
    if freq_config:
        ocl = freq_config["original_context_length"]
        low_wl, high_wl = ocl / freq_config["low_freq_factor"], ocl / freq_config["high_freq_factor"]
        wavelen = 2 * torch.pi / inv_freq
        slow = inv_freq / freq_config["factor"]
        base = torch.where(wavelen > low_wl, slow, inv_freq)
        mix = (ocl / wavelen - freq_config["low_freq_factor"]) / (
            freq_config["high_freq_factor"] - freq_config["low_freq_factor"])
        mid = (wavelen <= low_wl) & (wavelen >= high_wl)
        inv_freq = torch.where(mid, (1 - mix) * slow + mix * inv_freq, base)
    angles = torch.arange(context_length)[:, None] * inv_freq[None, :]
    angles = torch.cat([angles, angles], dim=1)
    return torch.cos(angles), torch.sin(angles)


class SharedBuffers:
    """Llama3.py:55-75: one (mask, cos, sin) triple per distinct configuration."""
    _buffers = {}

    @staticmethod
    def get_buffers(context_length, head_dim, rope_base, freq_config, dtype=torch.float32):
        key = (context_length, head_dim, rope_base, tuple(freq_config.values()) if freq_config else freq_config, dtype)
        if key not in SharedBuffers._buffers:
            # the fused kernel applies causality as a predicate; the mask buffer exists only so the
            # state_dict keeps the reference's keys. It is stored as a stride-0 expansion (no ctx^2 memory).
            cos, sin = precompute_rope_params(head_dim, rope_base, context_length, freq_config)
            if dtype:
                cos, sin = cos.to(dtype), sin.to(dtype)
            SharedBuffers._buffers[key] = (None, cos, sin)
        return SharedBuffers._buffers[key]


class GroupedQueryAttention(nn.Module):
    """Llama3.py:108-155."""

    def __init__(self, d_in, d_out, context_length, num_heads, num_kv_groups, rope_base=10_000, rope_config=None,
                 dtype=None):
        super().__init__()
        assert d_out % num_heads == 0
        assert num_heads % num_kv_groups == 0
        self.d_out = d_out
        self.num_heads = num_heads
        self.head_dim = d_out // num_heads
        self.num_kv_groups = num_kv_groups
        self.group_size = num_heads // num_kv_groups
        self.W_query = B200Linear(d_in, d_out, bias=False, dtype=dtype)
        self.W_key = B200Linear(d_in, num_kv_groups * self.head_dim, bias=False, dtype=dtype)
        self.W_value = B200Linear(d_in, num_kv_groups * self.head_dim, bias=False, dtype=dtype)
        self.out_proj = B200Linear(d_out, d_out, bias=False, dtype=dtype)
        _, cos, sin = SharedBuffers.get_buffers(context_length, self.head_dim, rope_base, rope_config, dtype)
        # `mask` is never read by the kernels (causality is a predicate) and is not kept resident
        # (16-64 MiB per layer, re-broadcast by DDP every step). state_dict() still
        # emits it and load_state_dict() accepts it, so checkpoints stay interchangeable.
        self.register_buffer("mask", torch.empty(0), persistent=False)
        self.register_buffer("cos", cos)
        self.register_buffer("sin", sin)

    def _save_to_state_dict(self, destination, prefix, keep_vars):
        super()._save_to_state_dict(destination, prefix, keep_vars)
        n = self.cos.shape[0]
        destination[prefix + "mask"] = torch.triu(torch.ones(n, n, device=self.cos.device), diagonal=1)

    def _load_from_state_dict(self, state_dict, prefix, *args):
        state_dict.pop(prefix + "mask", None)
        super()._load_from_state_dict(state_dict, prefix, *args)

    def forward(self, x, residual=None, cache=None, layer=0):
        b, num_tokens, _ = x.shape
        if x.dtype == torch.float32:      # fp32 parity mode: fp32 tables, plain-fp32 kernels
            cos, sin = self.cos.float(), self.sin.float()
        else:
            cos = self.cos if self.cos.dtype == torch.bfloat16 else self.cos.to(torch.bfloat16)
            sin = self.sin if self.sin.dtype == torch.bfloat16 else self.sin.to(torch.bfloat16)
        if cache is not None:       # generation with a KV cache (kvcache.py)
            from ...kvcache import attend_with_cache, packed_qkv
            ctx = attend_with_cache(packed_qkv(x, [self.W_query, self.W_key, self.W_value]), cos, sin, cache, layer,
                                    b, num_tokens, self.num_heads, self.num_kv_groups, self.head_dim)
            return self.out_proj(ctx.view(b, num_tokens, self.d_out), residual)
        ctx = ops.qkv_rope_attention(x, [self.W_query, self.W_key, self.W_value], cos, sin, b, num_tokens, self.num_heads,
                                     self.num_kv_groups, self.head_dim)
        if ctx is not None:     # projection + RoPE (GEMM epilogue) + attention as one op
            return self.out_proj(ctx.view(b, num_tokens, self.d_out), residual)
        qkv = ops.linears_packed(x, [self.W_query, self.W_key, self.W_value])
        if qkv is not None:     # one packed projection: RoPE and the backward dX GEMM run once over q|k|v
            ctx = ops.attention_qkv(qkv.view(b * num_tokens, -1), cos, sin, b, num_tokens, self.num_heads,
                                    self.num_kv_groups, self.head_dim)[0]
            return self.out_proj(ctx.view(b, num_tokens, self.d_out), residual)
        q, k, v = (t.view(b * num_tokens, -1)
                   for t in ops.linears_of_same_input(x, [self.W_query, self.W_key, self.W_value]))
        ctx = ops.attend(q, k, v, cos, sin, b, num_tokens, self.num_heads, self.num_kv_groups, self.head_dim)
        return self.out_proj(ctx.view(b, num_tokens, self.d_out), residual)


class TransformerBlock(nn.Module):
    """Llama3.py:159-181. The two residual adds are fused into the out_proj / fc3 GEMM epilogues."""

    def __init__(self, cfg):
        super().__init__()
        self.att = GroupedQueryAttention(d_in=cfg["emb_dim"], d_out=cfg["emb_dim"], context_length=cfg["context_length"],
                                         num_heads=cfg["n_heads"], num_kv_groups=cfg["n_kv_groups"],
                                         rope_base=cfg["rope_base"], rope_config=cfg["rope_freq"], dtype=cfg["dtype"])
        self.ff = FeedForward(cfg)
        self.norm1 = RMSNorm(cfg["emb_dim"], eps=1e-5)
        self.norm2 = RMSNorm(cfg["emb_dim"], eps=1e-5)

    def forward(self, x, cache=None, layer=0):
        h, res = ops.rmsnorm_with_residual(x, self.norm1.weight, self.norm1.eps)
        x = self.att(h.to(torch.bfloat16), residual=res, cache=cache, layer=layer)
        h, res = ops.rmsnorm_with_residual(x, self.norm2.weight, self.norm2.eps)
        return self.ff(h.to(torch.bfloat16), residual=res)


class Llama3Model(nn.Module):
    """Llama3.py:185-204."""

    def __init__(self, cfg, use_actv_ckpt=False):
        super().__init__()
        self.cfg = cfg
        self.use_actv_ckpt = use_actv_ckpt
        self.tok_emb = nn.Embedding(cfg["vocab_size"], cfg["emb_dim"], dtype=cfg["dtype"])
        self.trf_blocks = nn.Sequential(*[TransformerBlock(cfg) for _ in range(cfg["n_layers"])])
        self.final_norm = RMSNorm(cfg["emb_dim"], eps=1e-5)
        self.out_head = B200Linear(cfg["emb_dim"], cfg["vocab_size"], bias=False, dtype=cfg["dtype"])

    def hidden_states(self, in_idx):
        x = ops.embedding(in_idx, self.tok_emb.weight, None)
        if self.use_actv_ckpt:
            if not x.requires_grad:
                x.requires_grad_(True)  # frozen embeddings (LoRA): keep the checkpointed segments differentiable
            x = checkpoint_sequential(self.trf_blocks, self.cfg["n_layers"], x, use_reentrant=False)
        else:
            x = self.trf_blocks(x)
        return self.final_norm(x).to(torch.bfloat16)

    def forward(self, in_idx):
        return self.out_head(self.hidden_states(in_idx))

    def forward_last(self, in_idx):
        """Logits of the last position only, [B, V] (generate.py:42-45 needs nothing else)."""
        h = self.hidden_states(in_idx)
        return self.out_head(h[:, -1, :].contiguous())

    def new_kv_cache(self, batch, t_max=None):
        from ...kvcache import KVCache
        att = self.trf_blocks[0].att
        return KVCache(len(self.trf_blocks), batch, t_max or self.cfg["context_length"],
                       att.num_kv_groups * att.head_dim, self.tok_emb.weight.device)

    @torch.no_grad()
    def forward_cached(self, in_idx, cache):
        """Logits [B, V] of the last of the new tokens in_idx [B, T]; their keys / values join the cache
        (T > 1 only as the prefill of an empty cache)."""
        x = ops.embedding(in_idx, self.tok_emb.weight, None)
        for i, blk in enumerate(self.trf_blocks):
            x = blk(x, cache=cache, layer=i)
        cache.pos += in_idx.shape[1]
        h = self.final_norm(x).to(torch.bfloat16)
        return self.out_head(h[:, -1, :].contiguous())

    def forward_loss(self, in_idx, targets, capacity=0):
        """Fused vocab projection + cross-entropy (train.py:90-91) without a logits tensor
        handed back to Python; used by Trainer.calc_loss_batch in training mode."""
        w, A, B, s = head_operands(self.out_head)
        return ops.linear_cross_entropy(self.hidden_states(in_idx), w, A, B, s, targets, capacity, torch.is_grad_enabled())[0]
