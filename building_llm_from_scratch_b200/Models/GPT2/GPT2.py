"""Drop-in for Models/GPT2/GPT2.py: same class names, constructor signatures and state_dict keys.

Deviations (documented in DESIGN.md): the per-forward debug print (GPT2.py:123) is dropped; the
modules compute in bf16 (the GPU configs of BASELINE.json; the reference's GPT-2 ignores
cfg["dtype"]); dropout (embedding, residual and attention-weight sites) uses the
b200llm counter-based masks, so it is statistically but not bit-wise the reference's Philox stream
(per-op parity runs use drop_rate 0)."""
from __future__ import annotations

import torch
import torch.nn as nn
from torch.utils.checkpoint import checkpoint_sequential

from ... import ops
from ...lora import B200Linear, head_operands


class B200LayerNorm(nn.LayerNorm):
    def forward(self, x):
        return ops.layernorm(x, self.weight, self.bias, self.eps)[0]


class B200Dropout(nn.Dropout):
    """nn.Dropout whose training-mode forward is the counter-based b200llm dropout kernel."""

    def forward(self, x, residual=None):
        if self.training and self.p > 0:
            return ops.dropout(x, residual, self.p, ops.new_seed())
        return x if residual is None else x + residual


class GELU(nn.GELU):
    def forward(self, x):
        return ops.gelu(x)


class MultiHeadAttention(nn.Module):
    """GPT2.py:6-49."""

    def __init__(self, d_in, d_out, context_length, dropout, num_heads, qkv_bias=False):
        super().__init__()
        assert d_out % num_heads == 0, "Output dimension must be divisible by number of heads"
        self.d_out = d_out
        self.num_heads = num_heads
        self.head_dim = d_out // num_heads
        self.context_length = context_length
        self.W_query = B200Linear(d_in, d_out, bias=qkv_bias)
        self.W_key = B200Linear(d_in, d_out, bias=qkv_bias)
        self.W_value = B200Linear(d_in, d_out, bias=qkv_bias)
        self.out_proj = B200Linear(d_out, d_out)
        self.dropout = B200Dropout(dropout)   # applied to the attention weights inside the fused kernel
        self.register_buffer("mask", torch.empty(0), persistent=False)  # see Llama3.GroupedQueryAttention

    def _save_to_state_dict(self, destination, prefix, keep_vars):
        super()._save_to_state_dict(destination, prefix, keep_vars)
        n = self.context_length
        destination[prefix + "mask"] = torch.triu(torch.ones(n, n, device=self.out_proj.weight.device), diagonal=1)

    def _load_from_state_dict(self, state_dict, prefix, *args):
        state_dict.pop(prefix + "mask", None)
        super()._load_from_state_dict(state_dict, prefix, *args)

    def _project_out(self, ctx, residual, drop):
        """out_proj(ctx) [+ residual], or residual + dropout(out_proj(ctx)) in one op when drop = (p, seed) is given."""
        if drop is not None and residual is not None:
            w, bias, A, B, s = ops.linear_operands(self.out_proj)
            return ops.linear_dropout_residual(ctx, w, bias, A, B, s, residual, drop[0], drop[1])[0]
        return self.out_proj(ctx, residual)

    def forward(self, x, residual=None, cache=None, layer=0, drop=None):
        b, num_tokens, _ = x.shape
        if cache is not None:       # generation with a KV cache (kvcache.py); no RoPE, no dropout in eval
            from ...kvcache import attend_with_cache, packed_qkv
            ctx = attend_with_cache(packed_qkv(x, [self.W_query, self.W_key, self.W_value]), None, None, cache, layer,
                                    b, num_tokens, self.num_heads, self.num_heads, self.head_dim)
            return self.out_proj(ctx.view(b, num_tokens, self.d_out), residual)
        p_drop = self.dropout.p if self.training else 0.0
        qkv = ops.linears_packed(x, [self.W_query, self.W_key, self.W_value])     # None with qkv_bias
        if qkv is not None:
            ctx = ops.attention_qkv(qkv.view(b * num_tokens, -1), None, None, b, num_tokens, self.num_heads,
                                    self.num_heads, self.head_dim, p_drop, ops.new_seed() if p_drop > 0 else 0)[0]
            return self._project_out(ctx.view(b, num_tokens, self.d_out), residual, drop)
        q, k, v = (t.view(b * num_tokens, -1)
                   for t in ops.linears_of_same_input(x, [self.W_query, self.W_key, self.W_value]))
        ctx = ops.attend(q, k, v, None, None, b, num_tokens, self.num_heads, self.num_heads, self.head_dim,
                         p_drop, ops.new_seed() if p_drop > 0 else 0)
        return self._project_out(ctx.view(b, num_tokens, self.d_out), residual, drop)


class FeedForward(nn.Module):
    """GPT2.py:52-65."""

    def __init__(self, cfg):
        super().__init__()
        self.layers = nn.Sequential(B200Linear(cfg["emb_dim"], 4 * cfg["emb_dim"]), GELU(),
                                    B200Linear(4 * cfg["emb_dim"], cfg["emb_dim"]))

    def forward(self, x, residual=None, drop=None):
        """layers.2(gelu(layers.0 x)) [+ residual]; with drop = (p, seed): residual + dropout(...). One fused op: GELU
        in the first GEMM's epilogue, dropout + residual in the second's, GELU backward in the input-gradient GEMM's."""
        if x.dtype == torch.bfloat16 and x.is_cuda:
            w1, b1, A1, B1, s1 = ops.linear_operands(self.layers[0])
            w2, b2, A2, B2, s2 = ops.linear_operands(self.layers[2])
            if s1 == s2 or A1 is None or A2 is None:
                p, seed = drop if (drop is not None and residual is not None) else (0.0, 0)
                y = ops.gelu_mlp(x, w1, b1, A1, B1, w2, b2, A2, B2, s1 if A1 is not None else s2, residual, p, seed)[0]
                if drop is not None and residual is None:
                    y = ops.dropout(y, None, drop[0], drop[1])
                return y
        h = self.layers[1](self.layers[0](x))
        y = self.layers[2](h, None if drop is not None else residual)
        return ops.dropout(y, residual, drop[0], drop[1]) if drop is not None else y


class TransformerBlock(nn.Module):
    """GPT2.py:68-88."""

    def __init__(self, cfg):
        super().__init__()
        self.att = MultiHeadAttention(d_in=cfg["emb_dim"], d_out=cfg["emb_dim"], context_length=cfg["context_length"],
                                      num_heads=cfg["n_heads"], dropout=cfg["drop_rate"], qkv_bias=cfg["qkv_bias"])
        self.ff = FeedForward(cfg)
        self.norm1 = B200LayerNorm(cfg["emb_dim"])
        self.norm2 = B200LayerNorm(cfg["emb_dim"])
        self.dropout = B200Dropout(cfg["drop_rate"])

    def forward(self, x, cache=None, layer=0):
        if cache is not None:
            x = self.att(self.norm1(x), residual=x, cache=cache, layer=layer)
            return self.ff(self.norm2(x), residual=x)
        # the norm hands x on as the residual stream: its backward kernel adds the shortcut's gradient while it writes dx
        n1, n2 = self.norm1, self.norm2
        if self.training and self.dropout.p > 0:
            # x + drop_shortcut(att(...)) and x + drop_shortcut(ff(...)) (GPT2.py:85,87): dropout and the residual add
            # run in the epilogue of out_proj / the second MLP linear
            h, res = ops.layernorm_with_residual(x, n1.weight, n1.bias, n1.eps)
            x = self.att(h, residual=res, drop=(self.dropout.p, ops.new_seed()))
            h, res = ops.layernorm_with_residual(x, n2.weight, n2.bias, n2.eps)
            return self.ff(h, residual=res, drop=(self.dropout.p, ops.new_seed()))
        h, res = ops.layernorm_with_residual(x, n1.weight, n1.bias, n1.eps)
        x = self.att(h, residual=res)
        h, res = ops.layernorm_with_residual(x, n2.weight, n2.bias, n2.eps)
        return self.ff(h, residual=res)


class GPTModel(nn.Module):
    """GPT2.py:91-124."""

    def __init__(self, cfg, use_actv_ckpt=False):
        super().__init__()
        self.cfg = cfg
        self.use_actv_ckpt = use_actv_ckpt
        self.tok_emb = nn.Embedding(cfg["vocab_size"], cfg["emb_dim"])
        self.pos_emb = nn.Embedding(cfg["context_length"], cfg["emb_dim"])
        self.drop_emb = B200Dropout(cfg["drop_rate"])
        self.blocks = nn.Sequential(*[TransformerBlock(cfg) for _ in range(cfg["n_layers"])])
        self.norm = B200LayerNorm(cfg["emb_dim"])
        self.output_head = B200Linear(cfg["emb_dim"], cfg["vocab_size"], bias=False)

    def hidden_states(self, x):
        x = ops.embedding(x, self.tok_emb.weight, self.pos_emb.weight)
        x = self.drop_emb(x)
        if self.use_actv_ckpt:
            if not x.requires_grad:
                x.requires_grad_(True)
            x = checkpoint_sequential(self.blocks, segments=self.cfg["n_layers"], input=x, use_reentrant=False)
        else:
            x = self.blocks(x)
        return self.norm(x)

    def forward(self, x):
        return self.output_head(self.hidden_states(x))

    def forward_last(self, x):
        """Logits of the last position only, [B, V] (generate.py:42-45 needs nothing else)."""
        h = self.hidden_states(x)
        return self.output_head(h[:, -1, :].contiguous())

    def new_kv_cache(self, batch, t_max=None):
        from ...kvcache import KVCache
        return KVCache(len(self.blocks), batch, t_max or self.cfg["context_length"], self.cfg["emb_dim"],
                       self.tok_emb.weight.device, dtype=self.tok_emb.weight.dtype)

    @torch.no_grad()
    def forward_cached(self, x, cache):
        """Logits [B, V] of the last of the new tokens x [B, T]; positions continue at cache.pos."""
        T = x.shape[1]
        h = ops.embedding(x, self.tok_emb.weight, None) + self.pos_emb.weight[cache.pos:cache.pos + T]
        for i, blk in enumerate(self.blocks):
            h = blk(h, cache=cache, layer=i)
        cache.pos += T
        return self.output_head(self.norm(h)[:, -1, :].contiguous())

    def forward_loss(self, x, targets, capacity=0):
        w, A, B, s = head_operands(self.output_head)
        return ops.linear_cross_entropy(self.hidden_states(x), w, A, B, s, targets, capacity, torch.is_grad_enabled())[0]
