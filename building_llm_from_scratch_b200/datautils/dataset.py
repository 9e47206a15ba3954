"""Drop-in for datautils/dataset.py:21-34 (DatasetPT): sliding-window next-token pairs.

Windows start at range(0, len(tokens) - max_length, stride); input = tokens[i:i+L], target =
tokens[i+1:i+L+1]. Built with one strided view instead of a Python loop."""
from __future__ import annotations

import torch
from torch.utils.data import Dataset


def shift_windows(token_ids, max_length, stride):
    """-> (inputs [n, L], targets [n, L]) int64, same windows as the reference's loop."""
    tok = torch.as_tensor(token_ids, dtype=torch.int64)
    n = len(range(0, tok.numel() - max_length, stride))
    if n == 0:
        empty = torch.empty((0, max_length), dtype=torch.int64)
        return empty, empty.clone()
    inputs = tok.as_strided((n, max_length), (stride, 1)).clone()
    targets = tok.as_strided((n, max_length), (stride, 1), storage_offset=1).clone()
    return inputs, targets


class DatasetPT(Dataset):
    def __init__(self, txt, tokenizer, max_length, stride):
        token_ids = tokenizer.encode(txt, allowed_special={"<|endoftext|>"})
        self.inputs, self.targets = shift_windows(token_ids, max_length, stride)
        self.input_ids = list(self.inputs)
        self.target_ids = list(self.targets)

    def __len__(self):
        return self.inputs.shape[0]

    def __getitem__(self, idx):
        return self.inputs[idx], self.targets[idx]
