"""Drop-in for generate.py:4-75 — autoregressive decoding with top-k / temperature sampling.

Same signature and control flow as the reference. The model forward runs on the b200llm kernels; when
the model offers ``forward_last`` the vocab projection is evaluated for the last position only (the
reference projects all T positions and then slices ``[:, -1, :]``, generate.py:42-45). There is no
KV cache yet (neither has the reference): every new token costs a full forward of the window."""
from __future__ import annotations

import torch


def generate(model, idx: torch.Tensor, max_new_tokens: int, context_size: int, temperature: float = 0.0,
             top_k: int = None, eos_id: int = None) -> torch.Tensor:
    model.eval()
    device = next(model.parameters()).device
    idx = idx.to(device)
    last = getattr(model, "forward_last", None)
    with torch.no_grad():
        for _ in range(max_new_tokens):
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
