"""Drop-in for the collate step of datautils/dataloader_instruction_finetune.py (:10-50).

Same signature and bit-identical int64 outputs, but built with whole-batch tensor operations
instead of a per-item Python loop: pad, shift, keep only the first pad target per row, blank the
instruction prefix, truncate."""
from __future__ import annotations

import torch


def custom_collate_fn(batch, pad_token_id=50256, ignore_index=-100, allowed_max_length=None):
    n = len(batch)
    lens = torch.tensor([len(item) for _, item in batch], dtype=torch.int64)
    instr = torch.tensor([il for il, _ in batch], dtype=torch.int64)
    width = int(lens.max()) + 1                       # + one end-of-text token
    padded = torch.full((n, width), pad_token_id, dtype=torch.int64)
    for r, (_, item) in enumerate(batch):
        if len(item):
            padded[r, :len(item)] = torch.as_tensor(item, dtype=torch.int64)
    inputs = padded[:, :-1].clone()
    targets = padded[:, 1:].clone()
    # every pad-valued target except the first one in its row is ignored (this also hits pad ids
    # that occur naturally in the text, exactly like the reference)
    is_pad = targets == pad_token_id
    later_pad = is_pad & (is_pad.cumsum(dim=1) > 1)
    targets[later_pad] = ignore_index
    # targets[:instruction_length - 1] = ignore_index with Python slice semantics
    w = targets.shape[1]
    stop = instr - 1
    stop = torch.where(stop < 0, (w + stop).clamp(min=0), stop).clamp(max=w)
    cols = torch.arange(w)[None, :]
    targets[cols < stop[:, None]] = ignore_index
    if allowed_max_length is not None:
        inputs = inputs[:, :allowed_max_length]
        targets = targets[:, :allowed_max_length]
    return inputs.contiguous(), targets.contiguous()
