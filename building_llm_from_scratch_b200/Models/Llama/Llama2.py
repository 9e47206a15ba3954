"""Drop-in for the model classes of Models/Llama/Llama2.py (:34-190): MHA + RoPE (theta 10 000, no
frequency smoothing, fp32 cos/sin buffers), RMSNorm, SwiGLU. Llama-2 is GQA with kv groups == heads,
so it runs on the same kernels as Llama-3. The sentencepiece tokenizer class is out of scope.

Note: the reference's driver path for Llama-2 is broken; these modules are exercised at
module level, like the reference's can be."""
from __future__ import annotations

import torch
import torch.nn as nn

from ... import ops
from ...lora import B200Linear, head_operands
from .common_components import FeedForward, RMSNorm


def precompute_rope_params(head_dim, theta_base=10_000, context_length=4096):
    """Llama2.py:34-55."""
    assert head_dim % 2 == 0, "Embedding dimension must be even"
    inv_freq = 1.0 / (theta_base ** (torch.arange(0, head_dim, 2).float() / head_dim))
    angles = torch.arange(context_length)[:, None] * inv_freq[None, :]
    angles = torch.cat([angles, angles], dim=1)
    return torch.cos(angles), torch.sin(angles)


class MultiHeadAttention(nn.Module):
    """Llama2.py:61-118."""

    def __init__(self, d_in, d_out, context_length, num_heads, dtype=None):
        super().__init__()
        assert d_out % num_heads == 0, "d_out must be divisible by num_heads"
        self.d_out = d_out
        self.num_heads = num_heads
        self.head_dim = d_out // num_heads
        self.context_length = context_length
        self.W_query = B200Linear(d_in, d_out, bias=False, dtype=dtype)
        self.W_key = B200Linear(d_in, d_out, bias=False, dtype=dtype)
        self.W_value = B200Linear(d_in, d_out, bias=False, dtype=dtype)
        self.out_proj = B200Linear(d_out, d_out, bias=False, dtype=dtype)
        self.register_buffer("mask", torch.empty(0), persistent=False)   # see Llama3.GroupedQueryAttention
        cos, sin = precompute_rope_params(head_dim=self.head_dim, context_length=context_length)
        self.register_buffer("cos", cos)      # fp32, as in the reference
        self.register_buffer("sin", sin)
        self._tables_bf16 = None

    def _save_to_state_dict(self, destination, prefix, keep_vars):
        super()._save_to_state_dict(destination, prefix, keep_vars)
        n = self.context_length
        destination[prefix + "mask"] = torch.triu(torch.ones(n, n, device=self.cos.device), diagonal=1)

    def _load_from_state_dict(self, state_dict, prefix, *args):
        state_dict.pop(prefix + "mask", None)
        super()._load_from_state_dict(state_dict, prefix, *args)

    def _tables(self):
        # the rotation kernel reads bf16 tables; the fp32 buffers stay the state_dict truth
        t = self._tables_bf16
        if t is None or t[0].device != self.cos.device:
            t = self._tables_bf16 = (self.cos.to(torch.bfloat16), self.sin.to(torch.bfloat16))
        return t

    def forward(self, x, residual=None, cache=None, layer=0):
        b, num_tokens, _ = x.shape
        cos, sin = (self.cos, self.sin) if x.dtype == torch.float32 else self._tables()
        if cache is not None:       # generation with a KV cache (kvcache.py)
            from ...kvcache import attend_with_cache, packed_qkv
            ctx = attend_with_cache(packed_qkv(x, [self.W_query, self.W_key, self.W_value]), cos, sin, cache, layer,
                                    b, num_tokens, self.num_heads, self.num_heads, self.head_dim)
            return self.out_proj(ctx.view(b, num_tokens, self.d_out), residual)
        ctx = ops.qkv_rope_attention(x, [self.W_query, self.W_key, self.W_value], cos, sin, b, num_tokens, self.num_heads,
                                     self.num_heads, self.head_dim)
        if ctx is not None:     # projection + RoPE (GEMM epilogue) + attention as one op
            return self.out_proj(ctx.view(b, num_tokens, self.d_out), residual)
        qkv = ops.linears_packed(x, [self.W_query, self.W_key, self.W_value])
        if qkv is not None:
            ctx = ops.attention_qkv(qkv.view(b * num_tokens, -1), cos, sin, b, num_tokens, self.num_heads,
                                    self.num_heads, self.head_dim)[0]
            return self.out_proj(ctx.view(b, num_tokens, self.d_out), residual)
        q, k, v = (t.view(b * num_tokens, -1)
                   for t in ops.linears_of_same_input(x, [self.W_query, self.W_key, self.W_value]))
        ctx = ops.attend(q, k, v, cos, sin, b, num_tokens, self.num_heads, self.num_heads, self.head_dim)
        return self.out_proj(ctx.view(b, num_tokens, self.d_out), residual)


class TransformerBlock(nn.Module):
    """Llama2.py:124-155."""

    def __init__(self, cfg):
        super().__init__()
        self.att = MultiHeadAttention(d_in=cfg["emb_dim"], d_out=cfg["emb_dim"], context_length=cfg["context_length"],
                                      num_heads=cfg["n_heads"], dtype=cfg["dtype"])
        self.ff = FeedForward(cfg)
        self.norm1 = RMSNorm(cfg["emb_dim"])
        self.norm2 = RMSNorm(cfg["emb_dim"])

    def forward(self, x, cache=None, layer=0):
        h, res = ops.rmsnorm_with_residual(x, self.norm1.weight, self.norm1.eps)
        x = self.att(h, residual=res, cache=cache, layer=layer)
        h, res = ops.rmsnorm_with_residual(x, self.norm2.weight, self.norm2.eps)
        return self.ff(h, residual=res)


class Llama2Model(nn.Module):
    """Llama2.py:161-190."""

    def __init__(self, cfg):
        super().__init__()
        self.cfg = cfg
        self.tok_emb = nn.Embedding(cfg["vocab_size"], cfg["emb_dim"], dtype=cfg["dtype"])
        self.trf_blocks = nn.Sequential(*[TransformerBlock(cfg) for _ in range(cfg["n_layers"])])
        self.final_norm = RMSNorm(cfg["emb_dim"])
        self.out_head = B200Linear(cfg["emb_dim"], cfg["vocab_size"], bias=False, dtype=cfg["dtype"])

    def hidden_states(self, in_idx):
        x = ops.embedding(in_idx, self.tok_emb.weight, None)
        return self.final_norm(self.trf_blocks(x))

    def forward(self, in_idx):
        return self.out_head(self.hidden_states(in_idx))

    def forward_last(self, in_idx):
        """Logits of the last position only, [B, V] (generate.py:42-45 needs nothing else)."""
        h = self.hidden_states(in_idx)
        return self.out_head(h[:, -1, :].contiguous())

    def new_kv_cache(self, batch, t_max=None):
        from ...kvcache import KVCache
        return KVCache(len(self.trf_blocks), batch, t_max or self.cfg["context_length"], self.cfg["emb_dim"],
                       self.tok_emb.weight.device)

    @torch.no_grad()
    def forward_cached(self, in_idx, cache):
        """Logits [B, V] of the last of the new tokens; their keys / values join the cache."""
        x = ops.embedding(in_idx, self.tok_emb.weight, None)
        for i, blk in enumerate(self.trf_blocks):
            x = blk(x, cache=cache, layer=i)
        cache.pos += in_idx.shape[1]
        return self.out_head(self.final_norm(x)[:, -1, :].contiguous())

    def forward_loss(self, in_idx, targets, capacity=0):
        w, A, B, s = head_operands(self.out_head)
        return ops.linear_cross_entropy(self.hidden_states(in_idx), w, A, B, s, targets, capacity, torch.is_grad_enabled())[0]
