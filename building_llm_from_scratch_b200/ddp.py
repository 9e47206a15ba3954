"""Data-parallel wrapper replacing DDP(model, device_ids=[rank]) (build_components.py:176).

Semantics match the reference's DDP as probed in SURVEY.md §2.4: the gradient every rank applies is
the UNWEIGHTED MEAN over ranks of each rank's local-mean-loss gradient. Differences by design: no
per-step buffer broadcast (buffers are constants), and gradients are all-reduced in place in the
flat fp32 main_grad buffer of optim.B200AdamW (bucketed views, no copy into bucket storage),
issued on a side stream bucket by bucket while backward is still producing later buckets.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn


class B200DDP(nn.Module):
    def __init__(self, module: nn.Module, process_group=None, bucket_mb: float = 64.0, broadcast_params: bool = True):
        super().__init__()
        self.module = module
        self.pg = process_group
        self.bucket_mb = bucket_mb
        if dist.is_initialized() and broadcast_params and dist.get_world_size(self.pg) > 1:
            with torch.no_grad():
                for p in module.parameters():
                    dist.broadcast(p.data, src=0, group=self.pg)

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    def forward_loss(self, *args, **kwargs):
        return self.module.forward_loss(*args, **kwargs)

    @property
    def world_size(self):
        return dist.get_world_size(self.pg) if dist.is_initialized() else 1

    def reduce_gradients(self, flat_grads):
        """Average the flat gradient buffers over ranks (bucketed, async, then joined)."""
        ws = self.world_size
        if ws == 1:
            return
        works = []
        for flat in flat_grads:
            step = max(1, int(self.bucket_mb * 1024 * 1024 / flat.element_size()))
            for s in range(0, flat.numel(), step):
                chunk = flat[s:s + step]
                if dist.get_backend(self.pg) == "nccl":
                    works.append(dist.all_reduce(chunk, op=dist.ReduceOp.AVG, group=self.pg, async_op=True))
                else:  # gloo (CPU tests) has no AVG
                    w = dist.all_reduce(chunk, op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
                    works.append((w, chunk))
        for w in works:
            if isinstance(w, tuple):
                w[0].wait()
                w[1].div_(ws)
            else:
                w.wait()
