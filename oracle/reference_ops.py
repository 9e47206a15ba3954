"""ORACLE — test infrastructure only (never imported by the product path).

A CPU restatement, in plain PyTorch tensor arithmetic, of the reference's decoder training step
(chemphenoms/Building_LLM_from_scratch). Each function cites the reference file:line it follows
(paths relative to the reference root). It is written functionally over a state_dict that uses
the reference's own key names, so the same weights drive the reference modules, this oracle and
the CUDA path.

Pinning: the reference ships no golden vectors. The oracle is pinned against
outputs of the live reference modules, generated in the build container by oracle/make_golden.py
and committed under tests/golden/; tests/test_oracle_golden.py replays them on CPU.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.
"""
from __future__ import annotations

import math

import torch

# --------------------------------------------------------------------------------------
# Llama pieces
# --------------------------------------------------------------------------------------


def rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float = 1e-5) -> torch.Tensor:
    """Models/Llama/common_components.py:66-70. All ops run in x.dtype (so bf16 inputs round
    after every op); the fp32 weight multiply promotes, the result is cast back to x.dtype."""
    ms = x.pow(2).mean(dim=-1, keepdim=True)
    xn = x * torch.rsqrt(ms + eps)
    return (xn * weight).to(x.dtype)


def precompute_rope_params(head_dim, theta_base=10_000, context_length=4096, freq_config=None):
    """Models/Llama/Llama3.py:79-104 (incl. the Llama-3.1/3.2 frequency smoothing)."""
    inv_freq = 1.0 / (theta_base ** (torch.arange(0, head_dim, 2).float() / head_dim))
    if freq_config:
        ocl = freq_config["original_context_length"]
        low_wl = ocl / freq_config["low_freq_factor"]
        high_wl = ocl / freq_config["high_freq_factor"]
        wavelen = 2 * torch.pi / inv_freq
        scaled = inv_freq / freq_config["factor"]
        adjusted = torch.where(wavelen > low_wl, scaled, inv_freq)
        smooth = (ocl / wavelen - freq_config["low_freq_factor"]) / (
            freq_config["high_freq_factor"] - freq_config["low_freq_factor"])
        blended = (1 - smooth) * scaled + smooth * inv_freq
        medium = (wavelen <= low_wl) & (wavelen >= high_wl)
        inv_freq = torch.where(medium, blended, adjusted)
    pos = torch.arange(context_length)
    ang = pos[:, None] * inv_freq[None, :]
    ang = torch.cat([ang, ang], dim=1)
    return torch.cos(ang), torch.sin(ang)


def rescale_theta(theta_old, ctx_old, ctx_new):
    """Models/Llama/common_components.py:38-51."""
    return theta_old * (ctx_new / ctx_old)


def rope(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """Models/Llama/common_components.py:6-35. x: [B, H, T, hd]; rotate-half."""
    T, hd = x.shape[-2], x.shape[-1]
    half = hd // 2
    x1, x2 = x[..., :half], x[..., half:]
    c = cos[:T, :][None, None]
    s = sin[:T, :][None, None]
    rot = torch.cat((-x2, x1), dim=-1)
    return ((x * c) + (rot * s)).to(x.dtype)


def linear(x, w, b=None):
    y = x @ w.t()
    return y if b is None else y + b


def lora_linear(x, sd, prefix, scaling):
    """nn.Linear or LinearWithLoRA (lora.py:26,46): linear(x) + scaling * (x @ A @ B).
    `prefix` names the module in the state_dict; LoRA-wrapped linears have X.linear.weight."""
    if prefix + ".linear.weight" in sd:
        y = linear(x, sd[prefix + ".linear.weight"], sd.get(prefix + ".linear.bias"))
        return y + scaling * (x @ sd[prefix + ".lora.A"] @ sd[prefix + ".lora.B"])
    return linear(x, sd[prefix + ".weight"], sd.get(prefix + ".bias"))


def causal_attention(q, k, v, dropout_p=0.0):
    """Score / mask / softmax / value product shared by Models/Llama/Llama3.py:148-153 and
    Models/GPT2/GPT2.py:38-46. q,k,v: [B, H, T, hd] (k, v already expanded to H heads)."""
    T, hd = q.shape[-2], q.shape[-1]
    scores = (q @ k.transpose(2, 3)) / hd ** 0.5
    mask = torch.triu(torch.ones(T, T, device=q.device), diagonal=1).bool()
    scores = scores.masked_fill(mask, -torch.inf)
    w = torch.softmax(scores, dim=-1)
    assert dropout_p == 0.0, "oracle parity runs use drop_rate = 0 (SURVEY.md H5)"
    return w @ v


def gqa_attention(x, sd, prefix, n_heads, n_kv_groups, cos, sin, scaling=0.0):
    """GroupedQueryAttention.forward, Models/Llama/Llama3.py:131-155."""
    B, T, _ = x.shape
    q = lora_linear(x, sd, prefix + ".W_query", scaling)
    k = lora_linear(x, sd, prefix + ".W_key", scaling)
    v = lora_linear(x, sd, prefix + ".W_value", scaling)
    hd = q.shape[-1] // n_heads
    q = q.view(B, T, n_heads, hd).transpose(1, 2)
    k = k.view(B, T, n_kv_groups, hd).transpose(1, 2)
    v = v.view(B, T, n_kv_groups, hd).transpose(1, 2)
    q = rope(q, cos, sin)
    k = rope(k, cos, sin)
    g = n_heads // n_kv_groups
    k = k.repeat_interleave(g, dim=1)
    v = v.repeat_interleave(g, dim=1)
    ctx = causal_attention(q, k, v).transpose(1, 2).reshape(B, T, n_heads * hd)
    return lora_linear(ctx, sd, prefix + ".out_proj", scaling)


def swiglu_ffn(x, sd, prefix, scaling=0.0):
    """FeedForward.forward, Models/Llama/common_components.py:120-124 (SiLU at :87)."""
    a = lora_linear(x, sd, prefix + ".fc1", scaling)
    b = lora_linear(x, sd, prefix + ".fc2", scaling)
    h = (a * torch.sigmoid(a)) * b
    return lora_linear(h, sd, prefix + ".fc3", scaling)


def llama3_block(x, sd, prefix, cfg, scaling=0.0):
    """TransformerBlock.forward, Models/Llama/Llama3.py:176-181."""
    att_in = rmsnorm(x, sd[prefix + ".norm1.weight"]).to(torch.bfloat16)
    x = x + gqa_attention(att_in, sd, prefix + ".att", cfg["n_heads"], cfg["n_kv_groups"],
                          sd[prefix + ".att.cos"], sd[prefix + ".att.sin"], scaling)
    ff_in = rmsnorm(x, sd[prefix + ".norm2.weight"]).to(torch.bfloat16)
    return x + swiglu_ffn(ff_in, sd, prefix + ".ff", scaling)


def llama3_forward(idx, sd, cfg, scaling=0.0):
    """Llama3Model.forward, Models/Llama/Llama3.py:196-204 -> logits [B, T, V]."""
    x = sd["tok_emb.weight"][idx]
    for i in range(cfg["n_layers"]):
        x = llama3_block(x, sd, f"trf_blocks.{i}", cfg, scaling)
    x = rmsnorm(x, sd["final_norm.weight"])
    return lora_linear(x.to(torch.bfloat16), sd, "out_head", scaling)


# --------------------------------------------------------------------------------------
# GPT-2 pieces
# --------------------------------------------------------------------------------------


def layernorm(x, w, b, eps=1e-5):
    """nn.LayerNorm(emb_dim), Models/GPT2/GPT2.py:79-80,105 (biased variance, eps 1e-5)."""
    mu = x.mean(dim=-1, keepdim=True)
    var = ((x - mu) ** 2).mean(dim=-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * w + b


def gelu_erf(x):
    """nn.GELU() (approximate='none'), Models/GPT2/GPT2.py:60."""
    return 0.5 * x * (1.0 + torch.erf(x / math.sqrt(2.0)))


def mha_attention(x, sd, prefix, n_heads, scaling=0.0):
    """MultiHeadAttention.forward, Models/GPT2/GPT2.py:29-49 (dropout off)."""
    B, T, _ = x.shape
    q = lora_linear(x, sd, prefix + ".W_query", scaling)
    k = lora_linear(x, sd, prefix + ".W_key", scaling)
    v = lora_linear(x, sd, prefix + ".W_value", scaling)
    hd = q.shape[-1] // n_heads
    q, k, v = (t.view(B, T, n_heads, hd).transpose(1, 2) for t in (q, k, v))
    ctx = causal_attention(q, k, v).transpose(1, 2).reshape(B, T, n_heads * hd)
    return lora_linear(ctx, sd, prefix + ".out_proj", scaling)


def gpt2_block(x, sd, prefix, cfg, scaling=0.0):
    """TransformerBlock.forward, Models/GPT2/GPT2.py:83-88 (dropout off)."""
    h = layernorm(x, sd[prefix + ".norm1.weight"], sd[prefix + ".norm1.bias"])
    x = x + mha_attention(h, sd, prefix + ".att", cfg["n_heads"], scaling)
    h = layernorm(x, sd[prefix + ".norm2.weight"], sd[prefix + ".norm2.bias"])
    h = gelu_erf(lora_linear(h, sd, prefix + ".ff.layers.0", scaling))
    return x + lora_linear(h, sd, prefix + ".ff.layers.2", scaling)


def gpt2_forward(idx, sd, cfg, scaling=0.0):
    """GPTModel.forward, Models/GPT2/GPT2.py:108-124 (dropout off, debug print dropped)."""
    T = idx.shape[1]
    x = sd["tok_emb.weight"][idx] + sd["pos_emb.weight"][torch.arange(T, device=idx.device)][None]
    for i in range(cfg["n_layers"]):
        x = gpt2_block(x, sd, f"blocks.{i}", cfg, scaling)
    x = layernorm(x, sd["norm.weight"], sd["norm.bias"])
    return lora_linear(x, sd, "output_head", scaling)


# --------------------------------------------------------------------------------------
# loss / optimiser / schedule (train.py, build_components.py)
# --------------------------------------------------------------------------------------


def cross_entropy(logits: torch.Tensor, targets: torch.Tensor, ignore_index: int = -100):
    """train.py:91 — F.cross_entropy(logits.flatten(0,1), targets.flatten()): log-softmax in
    fp32, mean of -log p[target] over the rows whose target != ignore_index (NaN if none)."""
    lg = logits.reshape(-1, logits.shape[-1]).float()
    tg = targets.reshape(-1)
    lse = torch.logsumexp(lg, dim=-1)
    valid = tg != ignore_index
    picked = lg.gather(1, tg.clamp(min=0)[:, None])[:, 0]
    per_row = torch.where(valid, lse - picked, torch.zeros_like(lse))
    return (per_row.sum() / valid.sum()).to(logits.dtype)


def lr_at_step(step, peak_lr, total_steps, warmup_steps=10, initial_lr=1e-5, min_lr=1e-6):
    """train.py:100-104 (lr_increment from train.py:156)."""
    if step < warmup_steps:
        return initial_lr + step * (peak_lr - initial_lr) / warmup_steps
    progress = (step - warmup_steps) / (total_steps - warmup_steps)
    return min_lr + (peak_lr - min_lr) * 0.5 * (1 + math.cos(math.pi * progress))


def clip_coef(grads, max_norm=1.0):
    """train.py:115-120 -> torch.nn.utils.clip_grad_norm_: total L2 norm over all grads,
    coef = max_norm / (norm + 1e-6) clamped to 1."""
    total = torch.sqrt(sum((g.float() ** 2).sum() for g in grads))
    return total, torch.clamp(max_norm / (total + 1e-6), max=1.0)


def adamw_step(p, g, m, v, step, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.1):
    """torch.optim.AdamW (build_components.py:258, defaults + weight_decay=0.1), one tensor,
    fp32 math; returns (p, m, v). `step` is 1-based."""
    p = p * (1 - lr * weight_decay)
    m = beta1 * m + (1 - beta1) * g
    v = beta2 * v + (1 - beta2) * g * g
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    denom = v.sqrt() / math.sqrt(bc2) + eps
    return p - (lr / bc1) * (m / denom), m, v


# --------------------------------------------------------------------------------------
# data pipeline (integer, bit-exact)
# --------------------------------------------------------------------------------------


def shift_windows(tokens, max_length, stride):
    """DatasetPT.__init__, datautils/dataset.py:21-34: sliding windows, target = input + 1."""
    xs, ys = [], []
    for i in range(0, len(tokens) - max_length, stride):
        xs.append(list(tokens[i:i + max_length]))
        ys.append(list(tokens[i + 1:i + max_length + 1]))
    return xs, ys


def collate(batch, pad_token_id=50256, ignore_index=-100, allowed_max_length=None):
    """custom_collate_fn, datautils/dataloader_instruction_finetune.py:10-50.
    batch: list of (instruction_length, token_list). Returns int64 (inputs, targets)."""
    width = max(len(toks) for _, toks in batch) + 1
    ins, tgs = [], []
    for instr_len, toks in batch:
        padded = list(toks) + [pad_token_id]
        padded += [pad_token_id] * (width - len(padded))
        inp = padded[:-1]
        tgt = padded[1:]
        pad_pos = [i for i, t in enumerate(tgt) if t == pad_token_id]
        for i in pad_pos[1:]:          # keep the first pad as the end-of-text target
            tgt[i] = ignore_index
        # tgt[:instr_len-1] = ignore — Python slice semantics (instr_len 0 -> [: -1])
        stop = instr_len - 1
        n = len(tgt)
        hi = stop if stop >= 0 else max(n + stop, 0)
        for i in range(0, min(hi, n)):
            tgt[i] = ignore_index
        if allowed_max_length is not None:
            inp = inp[:allowed_max_length]
            tgt = tgt[:allowed_max_length]
        ins.append(inp)
        tgs.append(tgt)
    return torch.tensor(ins, dtype=torch.int64), torch.tensor(tgs, dtype=torch.int64)
