"""Static-shape training-step runner with CUDA-graph replay.

Runs exactly Trainer.train_batch's sequence (zero_grad, LR for the step, fused forward + CE,
backward, [data-parallel gradient mean], clip 1.0, AdamW) on static device buffers so the ~1.5 k
kernel launches of a step can be captured once and replayed without host launch cost
(SURVEY.md H8). Per-step scalars (learning rate, Adam bias corrections) are read by the kernels from
device memory that is refreshed before each replay.
"""
from __future__ import annotations

import os

import torch

from . import ops
from .ddp import B200DDP
from .optim import B200AdamW
from .train import Trainer


class TrainStep:
    def __init__(self, trainer: Trainer, batch_shape, use_graph: bool = True, capture_collective=None):
        assert isinstance(trainer.optimizer, B200AdamW), "TrainStep drives the fused optimizer"
        self.tr = trainer
        self.opt = trainer.optimizer
        self.model = trainer.model
        self.core = self.model.module if isinstance(self.model, B200DDP) else self.model
        self.ddp = self.model if isinstance(self.model, B200DDP) and self.model.world_size > 1 else None
        dev = next(self.core.parameters()).device
        self.x = torch.zeros(batch_shape, dtype=torch.int64, device=dev)
        self.y = torch.zeros(batch_shape, dtype=torch.int64, device=dev)
        self.loss = torch.zeros((), dtype=torch.float32, device=dev)
        self.use_graph = use_graph
        self.graphs = None            # graphs of the full-capacity step
        self.bucket_graphs = {}       # capacity -> graphs with a row-compacted vocab projection
        self.pool = None
        self.capacity = 0             # rows the vocab projection sees in the step being built (0 = all)
        self.compact_head = True
        # Data-parallel replay is ONE graph per step with the bucket all-reduces captured inside it, overlapping
        # backward (B200_GRAPH_NCCL=0 or capture_collective=False: two graphs with the exchange issued by the host
        # between them). NCCL refuses to destroy a communicator while graphs that captured its collectives
        # exist: call close() before destroy_process_group().
        if capture_collective is None:
            capture_collective = os.environ.get("B200_GRAPH_NCCL", "1") != "0"
        self.capture_collective = bool(capture_collective)
        # bucketed all-reduce overlapped with backward (eager steps and captured collectives); with the
        # default two-graph replay the exchange sits between the graphs and cannot overlap
        self.overlap = os.environ.get("B200_DDP_OVERLAP", "1") != "0"
        if self.ddp is not None:
            self.ddp.attach(self.opt)
        # vocab-projection capacity buckets (one captured graph each): 1/16 of the rows, so a batch pays for
        # at most ~6 % more rows than it has non-ignored targets
        self.bucket_rows = max(128, (batch_shape[0] * batch_shape[1]) // 16)
        self.launches_per_step = 0
        self.opt.graph_safe = True
        self.core.train()

    # -- the two halves of a step (split where the data-parallel exchange goes) -------------
    def _fwd_bwd(self, overlap: bool = False):
        # dropout masks are functions of (site seed, device step counter): bumping the counter inside the captured
        # step gives every replay its own masks (GPT-2's drop_rate 0.1 at 2L+1 sites + the attention weights)
        ops.advance_dropout_step(self.x.device)
        self.opt.zero_grad()
        if overlap:
            self.ddp.begin_backward()
        loss = self.core.forward_loss(self.x, self.y, capacity=self.capacity)
        loss.backward()
        # weight-gradient GEMMs still queued belong to THIS half of the step: with two graphs and the
        # all-reduce between them they would otherwise be launched by the optimizer, after the exchange
        self.opt.flush_pending_grads()
        self.loss.copy_(loss.detach())

    def _update(self):
        self.opt.step()

    def _exchange_step(self):
        """forward, backward with the gradient mean over ranks (overlapped when enabled), update."""
        if getattr(self.opt, "zero", False):     # ZeRO-1: the optimizer reduce-scatters the gradients itself
            self._fwd_bwd(False)
            self._update()
            return
        overlap = self.ddp is not None and self.overlap
        self._fwd_bwd(overlap)
        if overlap:
            self.ddp.finish_backward()
        elif self.ddp is not None:
            self.ddp.reduce_gradients(self.opt.flat_grads)
        self._update()

    def _eager_step(self):
        self._exchange_step()

    def close(self):
        """Drop the captured graphs (needed before destroy_process_group() when they hold collectives)."""
        import gc
        self.graphs = None
        self.bucket_graphs = {}
        gc.collect()
        torch.cuda.synchronize()

    def capture(self, warmup: int = 3):
        """Warm up eagerly on a side stream (lazy initialisation, allocator), then capture."""
        self._refresh_hyper(peek=True)
        # Warm-up steps must not move the weights or the optimizer state. They run the real kernels with neutral
        # hyper-parameters in the device buffer the AdamW kernel reads (lr 0, betas 1, weight decay 0: p, m and v
        # are rewritten with their own values) — no snapshot of the parameters and moments is needed, which for a
        # fully trained 8 B model would be 48 GB.
        saved_step = self.opt._step
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        from . import _lib
        with torch.cuda.stream(s):
            self.opt.freeze_updates(True)
            for _ in range(warmup):
                n0 = _lib.launch_count()
                self._eager_step()
                self.launches_per_step = _lib.launch_count() - n0   # b200llm kernels in one step
            self.opt.freeze_updates(False)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.opt._step = saved_step
        if not self.use_graph:
            return
        self.graphs = self._capture_graphs()

    def _capture_graphs(self):
        graphs = self._capture_graphs_impl()
        self.opt._step = self.tr.global_step + 1   # capturing ran B200AdamW.step() in Python once: undo its count
        return graphs

    def _capture_graphs_impl(self):
        kw = {} if self.pool is None else dict(pool=self.pool)
        if self.ddp is None:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, **kw):
                self._fwd_bwd()
                self._update()
            if self.pool is None:
                self.pool = g.pool()
            return (g,)
        if self.capture_collective or getattr(self.opt, "zero", False):
            # the NCCL all-reduce is captured too: the whole data-parallel step is ONE graph launch
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, **kw):
                self._exchange_step()
            if self.pool is None:
                self.pool = g.pool()
            return (g,)
        g1, g2 = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(g1, **kw):
            self._fwd_bwd()
        if self.pool is None:
            self.pool = g1.pool()
        with torch.cuda.graph(g2, pool=self.pool):
            self._update()
        return (g1, g2)

    def _graphs_for(self, n_valid):
        """Pick (capturing on first use) the graph whose vocab projection covers n_valid rows."""
        M = self.x.numel()
        if not self.compact_head or n_valid is None or n_valid <= 0:
            self.capacity = 0
            return self.graphs
        cap = (n_valid + self.bucket_rows - 1) // self.bucket_rows * self.bucket_rows
        if cap >= M:
            self.capacity = 0
            return self.graphs
        self.capacity = cap
        if self.use_graph and cap not in self.bucket_graphs:
            torch.cuda.synchronize()
            self.bucket_graphs[cap] = self._capture_graphs()
        return self.bucket_graphs.get(cap)

    def prepare(self, n_valid):
        """Capture (outside any timed region) the graph that a batch with n_valid non-ignored targets
        will replay."""
        self._graphs_for(n_valid)
        self.capacity = 0

    def _refresh_hyper(self, peek: bool = False):
        tr = self.tr
        step = tr.global_step + (1 if peek else 0)
        lr = tr.lr_for_step(max(step, 0))
        for pg in self.opt.param_groups:
            pg["lr"] = lr
        self.opt.set_hyper(lr, max(step, 0) + 1)
        return lr

    def run(self, x_src: torch.Tensor, y_src: torch.Tensor, n_valid=None):
        """One optimizer step on (x_src, y_src) — device tensors or pinned host tensors.
        n_valid: number of non-ignored targets in y_src if the caller knows it (counted here when
        y_src is a host tensor); it selects the row-compacted vocab projection.
        Returns the device scalar holding this step's loss (no synchronisation)."""
        tr = self.tr
        if n_valid is None and not y_src.is_cuda and self.compact_head:
            n_valid = int((y_src != -100).sum())
        graphs = self._graphs_for(n_valid)
        tr.global_step += 1
        lr = self._refresh_hyper()
        self.opt._step = tr.global_step          # steps taken before this one (an eager step() increments it)
        tr.track_lrs.append(lr)
        self.x.copy_(x_src, non_blocking=True)
        self.y.copy_(y_src, non_blocking=True)
        if graphs is None:
            self._eager_step()
        elif len(graphs) == 1:
            graphs[0].replay()
        else:
            graphs[0].replay()
            self.ddp.reduce_gradients(self.opt.flat_grads)
            graphs[1].replay()
        # replayed steps do not pass through B200AdamW.step(): keep its counter (saved in checkpoints, used for
        # the bias corrections of later eager steps) in line with the trainer's
        self.opt._step = tr.global_step + 1
        tr.tokens_seen += self.x.numel()
        return self.loss
