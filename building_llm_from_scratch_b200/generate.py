"""Drop-in for generate.py:4-75 — autoregressive decoding with top-k / temperature sampling.

Same signature and control flow as the reference. The model forward runs on the b200llm kernels; when
the model offers ``forward_last`` the vocab projection is evaluated for the last position only (the
reference projects all T positions and then slices ``[:, -1, :]``, generate.py:42-45). Models that offer
``new_kv_cache`` / ``forward_cached`` run the prompt once and then one token per step against a KV cache
(kvcache.py) as long as prompt + new tokens fit the context window; otherwise — and with
``use_cache=False`` — every new token costs a full forward of the window, as in the reference."""
from __future__ import annotations

import torch


def generate(model, idx: torch.Tensor, max_new_tokens: int, context_size: int, temperature: float = 0.0,
             top_k: int = None, eos_id: int = None, use_cache: bool = True) -> torch.Tensor:
    model.eval()
    device = next(model.parameters()).device
    idx = idx.to(device)
    last = getattr(model, "forward_last", None)
    cache = None
    if (use_cache and hasattr(model, "forward_cached") and idx.shape[1] + max_new_tokens <= context_size
            and idx.shape[1] > 1):
        cache = model.new_kv_cache(idx.shape[0], idx.shape[1] + max_new_tokens)
    with torch.no_grad():
        for step in range(max_new_tokens):
            if cache is not None:
                logits = model.forward_cached(idx if step == 0 else idx[:, -1:], cache)
            else:
                idx_cond = idx[:, -context_size:]
                logits = last(idx_cond) if last is not None else model(idx_cond)[:, -1, :]
            logits = logits.float()
            if top_k is not None:
                top_logits, _ = torch.topk(logits, top_k)
                min_val = top_logits[:, -1].unsqueeze(-1)
                logits = torch.where(logits < min_val, torch.full_like(logits, float("-inf")), logits)
            if temperature > 0.0:
                probs = torch.softmax(logits / temperature, dim=-1)
                idx_next = torch.multinomial(probs, num_samples=1)
            else:
                idx_next = torch.argmax(logits, dim=-1, keepdim=True)
            if eos_id is not None and bool((idx_next == eos_id).all()):
                break
            idx = torch.cat((idx, idx_next), dim=1)
    return idx
