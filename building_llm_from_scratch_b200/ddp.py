"""Data-parallel wrapper replacing DDP(model, device_ids=[rank]) (build_components.py:176).

Semantics match the reference's DDP as probed in SURVEY.md §2.4: the gradient every rank applies is
the UNWEIGHTED MEAN over ranks of each rank's local-mean-loss gradient. Differences by design: no
per-step buffer broadcast (buffers are constants), and gradients are all-reduced in place in the
flat fp32 main_grad buffer of optim.B200AdamW (bucket = a contiguous range of it, no copy into
bucket storage).

Two ways to run the exchange:
  reduce_gradients(flat_grads)   after backward: every bucket is issued at once, then joined.
  attach(optimizer) + begin_backward() ... finish_backward()
                                 overlapped: buckets are built over the parameters in reverse
                                 registration order (the order backward finishes them, as DDP does);
                                 ops.py reports the parameters whose weight-gradient GEMMs have been
                                 launched, autograd hooks report the ones whose gradient came back as
                                 .grad, and a bucket's all-reduce is issued (async, NCCL's own stream,
                                 ordered after the GEMMs by the stream dependency torch records) as
                                 soon as its last parameter is in, while backward keeps running.
Every rank runs the same program, so the buckets complete, and the collectives are issued, in the
same order on all ranks.
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

    # ---------------------------------------------------------------- overlapped exchange
    def attach(self, optimizer):
        """Build the buckets over optimizer's flat gradient buffers (optim.B200AdamW layout: `_groups`
        of dict(params, offsets, g)) and subscribe to gradient-ready notifications."""
        self.detach()
        self._opt = optimizer
        self._buckets = []
        self._bucket_of = {}
        self._active = False
        self._works = []
        for grp in optimizer._groups.values():
            flat = grp["g"]
            limit = max(1, int(self.bucket_mb * 1024 * 1024 / flat.element_size()))
            hi, members = flat.numel(), []
            pairs = list(zip(grp["params"], grp["offsets"]))
            for i in range(len(pairs) - 1, -1, -1):
                p, off = pairs[i]
                members.append(p)
                if hi - off >= limit or i == 0:
                    lo = 0 if i == 0 else off
                    b = dict(flat=flat, lo=lo, hi=hi, members=members, pending=0, issued=False)
                    for q in members:
                        self._bucket_of[id(q)] = b
                    self._buckets.append(b)
                    hi, members = lo, []
        self._hooks = [p.register_post_accumulate_grad_hook(self._autograd_hook)
                       for grp in optimizer._groups.values() for p in grp["params"]]
        self._subscribed = [p for grp in optimizer._groups.values() for p in grp["params"]]
        try:
            from . import ops
            ops.register_grad_ready(self._subscribed, self._on_ready)
        except Exception:        # CPU-only unit tests drive _on_ready themselves
            pass
        return self

    def detach(self):
        """Undo attach(): remove the autograd hooks and the gradient-ready subscriptions of this wrapper."""
        for h in getattr(self, "_hooks", []):
            h.remove()
        self._hooks = []
        subs = getattr(self, "_subscribed", [])
        if subs:
            try:
                from . import ops
                ops.unregister_grad_ready(subs)
            except Exception:
                pass
        self._subscribed = []

    def _autograd_hook(self, p):
        if not getattr(self, "_active", False) or p.grad is None:
            return               # the optimizer folds .grad into main_grad itself
        p.main_grad.add_(p.grad.to(torch.float32))
        p.grad = None
        self._on_ready([p])

    def begin_backward(self):
        for b in self._buckets:
            b["pending"] = len(b["members"])
            b["issued"] = False
        self._seen = set()
        self._works = []
        self._active = True

    def _on_ready(self, params):
        if not getattr(self, "_active", False):
            return
        for p in params:
            b = self._bucket_of.get(id(p))
            if b is None:
                continue
            if id(p) in self._seen:
                raise RuntimeError("B200DDP overlap: a parameter received a second gradient contribution after "
                                   "being reported complete (shared weights / gradient accumulation): use "
                                   "reduce_gradients() instead of begin_backward()/finish_backward()")
            self._seen.add(id(p))
            b["pending"] -= 1
            if b["pending"] == 0:
                self._issue(b)

    def _issue(self, b):
        b["issued"] = True
        if self.world_size == 1:
            return
        chunk = b["flat"][b["lo"]:b["hi"]]
        if dist.get_backend(self.pg) == "nccl":
            self._works.append(dist.all_reduce(chunk, op=dist.ReduceOp.AVG, group=self.pg, async_op=True))
        else:
            self._works.append((dist.all_reduce(chunk, op=dist.ReduceOp.SUM, group=self.pg, async_op=True), chunk))

    def finish_backward(self):
        """Flush what backward left queued, send the buckets that are still open (parameters that got
        no gradient this step keep their zeros) and join every all-reduce."""
        flush = getattr(self._opt, "flush_pending_grads", None)
        if flush is not None:
            flush()
        for b in self._buckets:
            if not b["issued"]:
                self._issue(b)
        self._active = False
        ws = self.world_size
        for w in self._works:
            if isinstance(w, tuple):
                w[0].wait()
                w[1].div_(ws)
            else:
                w.wait()
        self._works = []

    # ---------------------------------------------------------------- exchange after backward
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


def flat_checksum(tensors, chunk: int = 1 << 26) -> torch.Tensor:
    """Two position-sensitive 64-bit checksums over the raw bits of bf16 / fp32 tensors (device tensor [2], no host
    sync), computed in chunks so that a multi-GB flat buffer needs no multi-GB temporary."""
    acc = None
    for t in tensors:
        bits = t.reshape(-1).view(torch.int16 if t.element_size() == 2 else torch.int32)
        for s0 in range(0, bits.numel(), chunk):
            b = bits[s0:s0 + chunk].to(torch.int64)
            w = (torch.arange(s0, s0 + b.numel(), device=b.device, dtype=torch.int64) % 65521) + 1
            part = torch.stack([b.sum(), (b * w).sum()])
            acc = part if acc is None else acc + part
    return acc


def ranks_identical(tensors, group=None) -> bool:
    """True when every rank of the group holds bit-identical copies of `tensors` (checksums all-gathered)."""
    mine = flat_checksum(tensors)
    world = dist.get_world_size(group)
    every = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(every, mine, group=group)
    return all(bool(torch.equal(e, every[0])) for e in every)
