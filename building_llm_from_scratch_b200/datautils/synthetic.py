"""Synthetic batches of the shapes SURVEY.md §8(d) defines (no dataset / tokenizer download)."""
from __future__ import annotations

import math

import torch

from .dataloader_instruction_finetune import custom_collate_fn
from .dataset import shift_windows


def alpaca_batch(gen: torch.Generator, batch_size: int, seq_len: int, vocab: int):
    """Alpaca-shaped: instruction 32..256 tokens, log-normal response, uniform token ids (so
    natural 50256s occur), one sample >= seq_len so the batch is [B, seq_len]; through the
    collate function -> int64 (inputs, targets) with prompt / pad positions = -100."""
    items = []
    for i in range(batch_size):
        instr = int(torch.randint(32, 257, (1,), generator=gen))
        resp = int(min(max(math.exp(math.log(200.0) + 0.8 * float(torch.randn(1, generator=gen))), 8), seq_len))
        n = seq_len + 8 if i == 0 else min(instr + resp, seq_len)
        toks = torch.randint(0, vocab, (n,), generator=gen).tolist()
        items.append((min(instr, n), toks))
    pad = 50256 if vocab > 50256 else vocab - 1
    return custom_collate_fn(items, pad_token_id=pad, allowed_max_length=seq_len)


def gutenberg_batch(gen: torch.Generator, batch_size: int, seq_len: int, vocab: int):
    """Gutenberg-shaped: a uniform token stream windowed like DatasetPT with stride = seq_len."""
    stream = torch.randint(0, vocab, (batch_size * seq_len + 1,), generator=gen)
    return shift_windows(stream, seq_len, seq_len)
