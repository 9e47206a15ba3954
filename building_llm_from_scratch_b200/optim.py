"""Fused AdamW + gradient clipping over flat buffers (replaces torch.optim.AdamW at
build_components.py:258 and clip_grad_norm_ at train.py:115-120 on the b200llm path).

All trainable parameters of one dtype live in ONE flat buffer (each nn.Parameter's .data is
re-pointed to a view of it); their gradients accumulate in ONE fp32 flat buffer whose views are
published as ``param.main_grad`` — the backward GEMMs of ops.py add into them directly, so there is
no per-parameter gradient tensor, no foreach launch train and the data-parallel all-reduce runs on
one contiguous buffer. A step is: grad-norm kernel -> clip coefficient (device scalar) -> one AdamW
kernel that also applies the coefficient. State (exp_avg, exp_avg_sq) has the parameter dtype,
like the reference's optimizer.

zero=True shards the optimizer state over the data-parallel group (the reference's --use_zero_opt,
build_components.py:228-231,250-256: ZeroRedundancyOptimizer around AdamW): each rank keeps exp_avg /
exp_avg_sq for 1/world of the flat buffer only; a step is reduce-scatter(AVG) of the flat gradients ->
global norm from the per-shard sums of squares (one scalar all-reduce) -> AdamW on the local shard ->
all-gather of the updated parameter shards. The result is the same update plain DDP + AdamW makes.
"""
from __future__ import annotations

import torch

from . import kernels as K


def _align(n, a=64):
    return (n + a - 1) // a * a


class B200AdamW(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, zero=False,
                 process_group=None, master_weights=False):
        params = [p for p in params]
        import torch.distributed as dist
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if (zero and dist.is_initialized()) else 1
        self.rank = dist.get_rank(process_group) if self.world > 1 else 0
        self.zero = bool(zero) and self.world > 1
        # master_weights: the bf16_hybrid policy of datautils/mixed_precision.py:24-28 — bf16 parameters keep an fp32
        # master copy and fp32 moments; every step updates the master and rewrites the bf16 copy the GEMMs read
        self.master_weights = bool(master_weights)
        if self.master_weights and self.zero:
            raise ValueError("B200AdamW: master_weights together with zero=True is not supported")
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self._check_groups()
        # the reference hands every parameter to AdamW; frozen ones never get a gradient
        self._trainable = [p for g in self.param_groups for p in g["params"] if p.requires_grad]
        if not self._trainable:
            raise ValueError("B200AdamW: no trainable parameters")
        dev = self._trainable[0].device
        if dev.type != "cuda":
            raise RuntimeError("B200AdamW needs CUDA parameters (no CPU fallback)")
        self._groups = {}
        for dtype in (torch.bfloat16, torch.float32):
            ps = [p for p in self._trainable if p.dtype == dtype]
            if not ps:
                continue
            offs, total = [], 0
            for p in ps:
                offs.append(total)
                total += _align(p.numel())
            if self.zero:                       # equal shards of whole 64-element groups
                total = _align(total, 64 * self.world)
            flat_p = torch.zeros(total, device=dev, dtype=dtype)
            flat_g = torch.zeros(total, device=dev, dtype=torch.float32)
            for p, o in zip(ps, offs):
                flat_p[o:o + p.numel()].copy_(p.data.reshape(-1))
                p.data = flat_p[o:o + p.numel()].view(p.shape)
                p.main_grad = flat_g[o:o + p.numel()].view(p.shape)
            if self.zero:
                shard = total // self.world
                lo = self.rank * shard
                self._groups[dtype] = dict(params=ps, offsets=offs, p=flat_p, g=flat_g,
                                           p_shard=flat_p[lo:lo + shard], g_shard=torch.zeros(shard, device=dev),
                                           m=torch.zeros(shard, device=dev, dtype=dtype),
                                           v=torch.zeros(shard, device=dev, dtype=dtype))
            elif self.master_weights and dtype == torch.bfloat16:
                self._groups[dtype] = dict(params=ps, offsets=offs, p=flat_p, g=flat_g, master=flat_p.float(),
                                           m=torch.zeros(total, device=dev), v=torch.zeros(total, device=dev))
            else:
                self._groups[dtype] = dict(params=ps, offsets=offs, p=flat_p, g=flat_g, m=torch.zeros_like(flat_p),
                                           v=torch.zeros_like(flat_p))
        other = [p for p in self._trainable if p.dtype not in self._groups]
        if other:
            raise RuntimeError(f"B200AdamW: unsupported parameter dtypes {set(p.dtype for p in other)}")
        self._step = 0
        self._norm_out = torch.zeros(2, device=dev, dtype=torch.float32)
        self._ws = torch.empty(K._lib.load().b200_grad_norm_workspace_bytes(1), device=dev, dtype=torch.uint8)
        self._hyper_dev = torch.zeros(7, device=dev, dtype=torch.float32)
        self._partial = torch.zeros(2 * len(self._groups), device=dev, dtype=torch.float32)
        self.max_norm = None  # set by Trainer (train.py:115-120 uses 1.0)
        self.graph_safe = False
        self._frozen = False

    def _check_groups(self):
        """One flat buffer per dtype is updated by one kernel launch with one set of hyper-parameters (the
        reference builds a single group, build_components.py:258): groups that differ are refused, not ignored."""
        g0 = self.param_groups[0]
        for g in self.param_groups[1:]:
            for k in ("lr", "betas", "eps", "weight_decay"):
                if g[k] != g0[k]:
                    raise ValueError(f"B200AdamW: param groups with different {k} ({g0[k]} vs {g[k]}) are not supported")

    # ------------------------------------------------------------------
    def flush_pending_grads(self):
        """Everything backward produced is in main_grad after this: queued weight-gradient GEMMs are
        launched and gradients autograd delivered as .grad (norm weights, biases) are folded in."""
        from . import ops
        ops.flush_weight_grads()
        self._absorb_autograd_grads()

    @property
    def flat_grads(self):
        self.flush_pending_grads()        # must land before anyone (all-reduce, norm) reads main_grad
        return [g["g"] for g in self._groups.values()]

    def zero_grad(self, set_to_none: bool = True):
        from . import ops
        ops.flush_weight_grads()
        for g in self._groups.values():
            g["g"].zero_()
        for p in self._trainable:
            p.grad = None
            p._b200_grad_fresh = True     # lets the first weight-gradient GEMM store instead of accumulate

    def _absorb_autograd_grads(self):
        # parameters whose gradient came back through autograd (ops outside this package) are folded in
        for p in self._trainable:
            if p.grad is not None:
                p.main_grad.add_(p.grad.to(torch.float32))
                p.grad = None

    def grad_norm(self, max_norm: float):
        """Global L2 norm over every gradient -> device tensor [norm, clip_coef]."""
        from . import ops
        ops.flush_weight_grads()
        groups = list(self._groups.values())
        if len(groups) == 1:
            return K.grad_norm(groups[0]["g"], max_norm, out=self._norm_out, ws=self._ws)
        parts = [K.grad_norm(g["g"], max_norm)[0] for g in groups]
        total = torch.sqrt(sum(p * p for p in parts))
        self._norm_out[0] = total
        self._norm_out[1] = torch.clamp(max_norm / (total + 1e-6), max=1.0)
        return self._norm_out

    @torch.no_grad()
    def step(self, closure=None):
        assert closure is None
        from . import ops
        ops.flush_weight_grads()
        self._absorb_autograd_grads()
        self._check_groups()
        self._step += 1
        if self.zero:
            return self._zero_step()
        scale = None
        if self.max_norm is not None:
            scale = self.grad_norm(self.max_norm)[1:2]
        pg = self.param_groups[0]
        lr, (b1, b2), eps, wd = pg["lr"], pg["betas"], pg["eps"], pg["weight_decay"]
        hyper = self._graph_hyper(lr)
        for g in self._groups.values():
            if "master" in g:
                K.adamw_master_(g["master"], g["g"], g["m"], g["v"], g["p"], self._step, lr, b1, b2, eps, wd,
                                grad_scale=scale, dev_hyper=hyper)
            else:
                K.adamw_(g["p"], g["g"], g["m"], g["v"], self._step, lr, b1, b2, eps, wd, grad_scale=scale,
                         dev_hyper=hyper)

    def _graph_hyper(self, lr):
        """graph_safe: the kernels read (lr, bias corrections) from device memory so that a captured step can be
        replayed with new values. Outside a capture the buffer is refreshed here from param_groups / the step
        counter, so an eager step() (Trainer.train_batch) is never run on stale values; during a capture it is
        left alone (TrainStep refreshes it before every replay)."""
        if not self.graph_safe:
            return None
        if not torch.cuda.is_current_stream_capturing() and not self._frozen:
            self.set_hyper(lr, max(self._step, 1))
        return self._hyper_dev

    def freeze_updates(self, on: bool):
        """While on, graph_safe steps run the real kernels with neutral hyper-parameters (lr 0, beta1 = beta2 = 1,
        weight decay 0, no bias correction): parameters and moments are rewritten with their own values. Used for
        the warm-up steps before a CUDA-graph capture."""
        if on and not self.graph_safe:
            raise RuntimeError("freeze_updates needs graph_safe (hyper-parameters read from device memory)")
        self._frozen = bool(on)
        if on:
            eps = self.param_groups[0]["eps"]
            host = torch.tensor([0.0, 1.0, 1.0, eps, 0.0, 0.0, 1.0], dtype=torch.float32)
            self._hyper_dev.copy_(host)

    def _zero_step(self):
        """ZeRO-1 update (see the module docstring); every collective is on the current stream's
        process group, so the whole thing can sit inside a captured graph."""
        import torch.distributed as dist
        groups = list(self._groups.values())
        for g in groups:
            dist.reduce_scatter_tensor(g["g_shard"], g["g"], op=dist.ReduceOp.AVG, group=self.pg)
        scale = None
        if self.max_norm is not None:
            sq = None
            for g in groups:
                part = K.grad_norm(g["g_shard"], self.max_norm)[0:1]
                sq = part * part if sq is None else sq + part * part
            dist.all_reduce(sq, group=self.pg)
            total = torch.sqrt(sq)
            self._norm_out[0:1] = total
            self._norm_out[1:2] = torch.clamp(self.max_norm / (total + 1e-6), max=1.0)
            scale = self._norm_out[1:2]
        pg = self.param_groups[0]
        lr, (b1, b2), eps, wd = pg["lr"], pg["betas"], pg["eps"], pg["weight_decay"]
        hyper = self._graph_hyper(lr)
        for g in groups:
            K.adamw_(g["p_shard"], g["g_shard"], g["m"], g["v"], self._step, lr, b1, b2, eps, wd, grad_scale=scale,
                     dev_hyper=hyper)
            dist.all_gather_into_tensor(g["p"], g["p_shard"], group=self.pg)

    # ---- checkpoint / resume (the reference saves the model only, train.py:231-257; a resume needs the
    # moments and the step count too) ------------------------------------------------------------------
    def state_dict(self):
        """Flat moments of the WHOLE parameter buffer + step. With ZeRO-1 every rank owns 1/world of the moments:
        they are all-gathered here, so this is a COLLECTIVE call (every rank of the group must make it) and the
        result is the same on every rank and independent of the sharding — a checkpoint written by rank 0 alone
        (Trainer.save_training_state) can be resumed with any world size, with or without ZeRO."""
        import torch.distributed as dist
        moments = {}
        for dt, g in self._groups.items():
            m, v = g["m"], g["v"]
            if self.zero:
                fm, fv = torch.empty_like(g["p"]), torch.empty_like(g["p"])
                dist.all_gather_into_tensor(fm, m, group=self.pg)
                dist.all_gather_into_tensor(fv, v, group=self.pg)
                m, v = fm, fv
            moments[str(dt)] = {"m": m.detach().cpu(), "v": v.detach().cpu(), "numel": g["p"].numel(),
                                "param_numel": sum(p.numel() for p in g["params"])}
            if "master" in g:
                moments[str(dt)]["master"] = g["master"].detach().cpu()
        return {"b200_adamw": 2, "step": self._step, "layout": "full", "zero": self.zero, "world": self.world,
                "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups],
                "moments": moments}

    def load_state_dict(self, state):
        if not isinstance(state, dict) or state.get("b200_adamw") not in (1, 2):
            raise ValueError("B200AdamW.load_state_dict: not a B200AdamW state (torch.optim.AdamW states are laid out per parameter)")
        if state["b200_adamw"] == 1 and state.get("zero"):
            raise ValueError("B200AdamW.load_state_dict: version-1 ZeRO checkpoints hold only the saving rank's shard of "
                             "the moments and cannot be resumed; re-save with this version")
        for dt, g in self._groups.items():
            src = state["moments"][str(dt)]
            n_par = sum(p.numel() for p in g["params"])
            if src.get("param_numel", n_par) != n_par:
                raise ValueError("B200AdamW.load_state_dict: flat buffer sizes differ (different set of trainable parameters)")
            # the flat buffers may be padded differently (ZeRO rounds the total up to a multiple of 64 * world);
            # the padding holds no parameter, so only the common prefix matters
            total = g["p"].numel()
            for key in ("m", "v"):
                full = torch.zeros(total, dtype=g[key].dtype)
                n = min(total, src[key].numel())
                if n < g["offsets"][-1] + g["params"][-1].numel():
                    raise ValueError("B200AdamW.load_state_dict: checkpoint moments are shorter than the parameters")
                full[:n].copy_(src[key][:n])
                if self.zero:
                    shard = total // self.world
                    full = full[self.rank * shard:(self.rank + 1) * shard]
                g[key].copy_(full)
            if "master" in g:
                if "master" in src:
                    n = min(total, src["master"].numel())
                    g["master"][:n].copy_(src["master"][:n])
                else:                    # checkpoint written without master weights: start them from the bf16 copy
                    g["master"].copy_(g["p"].float())
        self._step = int(state["step"])
        for pg, saved in zip(self.param_groups, state["param_groups"]):
            pg.update(saved)

    def set_hyper(self, lr: float, step: int):
        """Upload (lr, bias corrections for `step`) for the graph-captured step."""
        pg = self.param_groups[0]
        host = K.adam_hyper(lr, pg["betas"][0], pg["betas"][1], pg["eps"], pg["weight_decay"], step)
        self._hyper_dev.copy_(host, non_blocking=True)
        self._host_keepalive = host

    def last_grad_norm(self):
        return self._norm_out[0]
