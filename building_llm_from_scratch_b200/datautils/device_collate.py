"""The data side of the step on the device.

`custom_collate_fn_device` produces the same int64 (inputs, targets) as custom_collate_fn
(datautils/dataloader_instruction_finetune.py:10-50) but on the GPU: the batch's token lists are packed
into one pinned host buffer, sent with a single asynchronous copy, and padded / shifted / masked by
`b200_collate_i64`; the step's H2D traffic is the real tokens only and nothing blocks on `.to(device)`
(train.py:89). `window_batch` gathers DatasetPT windows (datautils/dataset.py:21-34) from a token
stream that already lives on the device. Integer work: bit-exact against the host versions."""
from __future__ import annotations

import torch

from .. import _lib


def _stream():
    return torch.cuda.current_stream().cuda_stream


def custom_collate_fn_device(batch, pad_token_id=50256, ignore_index=-100, allowed_max_length=None, device="cuda"):
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("custom_collate_fn_device needs a CUDA device (use custom_collate_fn on the host)")
    n = len(batch)
    lens = [len(item) for _, item in batch]
    width = max(lens)
    if width == 0:
        raise ValueError("custom_collate_fn_device: every item of the batch is empty")
    out_cols = width if allowed_max_length is None else min(width, allowed_max_length)
    # one pinned staging buffer: [offsets (n + 1) | instruction lengths (n) | tokens]
    total = sum(lens)
    host = torch.empty(2 * n + 1 + total, dtype=torch.int64, pin_memory=True)
    off = 0
    for r, (il, item) in enumerate(batch):
        host[r] = off
        host[n + 1 + r] = il
        if lens[r]:
            host[2 * n + 1 + off:2 * n + 1 + off + lens[r]] = torch.as_tensor(item, dtype=torch.int64)
        off += lens[r]
    host[n] = off
    dev = host.to(device, non_blocking=True)
    inputs = torch.empty((n, out_cols), dtype=torch.int64, device=device)
    targets = torch.empty((n, out_cols), dtype=torch.int64, device=device)
    lib = _lib.load()
    base = dev.data_ptr()
    _lib.check(lib.b200_collate_i64(base + 8 * (2 * n + 1), base, base + 8 * (n + 1), n, width, out_cols,
                                    pad_token_id, ignore_index, inputs.data_ptr(), targets.data_ptr(), _stream()),
               "collate")
    inputs._b200_staging = (host, dev)      # keep the staging buffers alive until the kernel has run
    return inputs, targets


def window_batch(tokens_dev: torch.Tensor, starts, max_length: int):
    """inputs[b] = tokens[s_b : s_b + L], targets[b] = tokens[s_b + 1 : s_b + L + 1] for the window
    starts `starts` (list or tensor) of a device-resident int64 token stream."""
    if not tokens_dev.is_cuda or tokens_dev.dtype != torch.int64 or not tokens_dev.is_contiguous():
        raise RuntimeError("window_batch needs a contiguous int64 CUDA token stream")
    starts = torch.as_tensor(starts, dtype=torch.int64)
    if int(starts.min()) < 0 or int(starts.max()) + max_length + 1 > tokens_dev.numel():
        raise ValueError("window_batch: a window runs past the end of the token stream")
    s_dev = starts.to(tokens_dev.device, non_blocking=True)
    n = starts.numel()
    inputs = torch.empty((n, max_length), dtype=torch.int64, device=tokens_dev.device)
    targets = torch.empty_like(inputs)
    lib = _lib.load()
    _lib.check(lib.b200_window_batch_i64(tokens_dev.data_ptr(), tokens_dev.numel(), s_dev.data_ptr(), n, max_length,
                                         inputs.data_ptr(), targets.data_ptr(), _stream()), "window_batch")
    return inputs, targets
