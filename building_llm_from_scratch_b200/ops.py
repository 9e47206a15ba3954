"""torch.library custom ops (namespace ``b200llm``) — the middle face of the drop-in boundary.

One op per fused unit of the decoder training step; each has a fake (meta) implementation for
shape/dtype propagation and an autograd formula whose backward launches the matching sm_100a
kernels through the C-ABI (kernels.py). There is no CPU or library fallback: every op raises on
non-CUDA tensors.

Weight gradients: if a parameter carries a ``main_grad`` attribute (an fp32 view into the flat
gradient buffer owned by optim.B200AdamW), the backward GEMM accumulates straight into it with
split-K fp32 atomics and autograd receives ``None`` for that parameter; otherwise a bf16 gradient
tensor is returned and autograd accumulates it the usual way.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import os
import weakref

import torch
from torch import Tensor

from . import f32 as K32
from . import kernels as K

BF16 = torch.bfloat16
F32 = torch.float32      # fp32 parity mode: ops handed float32 activations run the plain-fp32 kernels (f32.py)


def _pitch_ok(t: Optional[Tensor]) -> Optional[Tensor]:
    """TMA needs 16-byte row pitches. LoRA's B [r, out] has an odd pitch when `out` is GPT-2's
    vocabulary (50257): hand the GEMM a copy whose rows are padded to a multiple of 8 elements."""
    if t is None or t.dim() != 2 or (t.stride(0) % 8 == 0 and t.stride(1) == 1 and t.data_ptr() % 16 == 0):
        return t
    ld = (t.shape[1] + 7) // 8 * 8
    buf = t.new_zeros((t.shape[0], ld))
    buf[:, :t.shape[1]].copy_(t)
    return buf[:, :t.shape[1]]


# Saved tensors are not always the Parameter objects: under torch.utils.checkpoint (use_reentrant=False) the backward of
# a recomputed segment gets DETACHED copies of whatever forward saved — same storage, no Python attributes. main_grad,
# _b200_grad_fresh and the gradient-ready subscriptions live on the Parameter, so every setup_context keeps the
# parameters it was called with (`ctx.p_*`) and backward looks attributes up on those; a LoRA / full-finetune block
# under activation checkpointing otherwise fell off the main_grad path (zero-filled fp32 temporaries, split-K atomics,
# cast, .grad, a second add: 7 weights x 4 extra launches and weight-gradient GEMMs at 2/3 of their speed per block).
def _main_grad(*params):
    """The fp32 main_grad accumulators of parameters (optim.B200AdamW publishes them) when ALL of them have one, else
    None: backward kernels that produce small gradients (norm weights, biases, embeddings) then add straight into the
    flat gradient buffer instead of returning a tensor that autograd stores in .grad and the optimizer folds in later
    (zero-fill + cast + copy + add per parameter: ~28 tiny launches per GPT-2 block)."""
    gs = [getattr(p, "main_grad", None) if p is not None else None for p in params]
    return gs if all(g is not None for g in gs) else None


def _flat2d(x: Tensor) -> Tensor:
    x2 = x.reshape(-1, x.shape[-1])
    return x2 if x2.stride(-1) == 1 else x2.contiguous()


# ======================================================================================
# RMSNorm  (Models/Llama/common_components.py:54-70)
# ======================================================================================
@torch.library.custom_op("b200llm::rmsnorm", mutates_args=())
def rmsnorm(x: Tensor, weight: Tensor, eps: float) -> Tensor:
    return K.rmsnorm_fwd(x.contiguous(), weight, eps)


@rmsnorm.register_fake
def _(x, weight, eps):
    return torch.empty_like(x)


def _rmsnorm_setup(ctx, inputs, output):
    x, w, eps = inputs
    ctx.save_for_backward(x, w)
    ctx.eps = eps
    ctx.p_w = w


def _rmsnorm_bwd(ctx, dy):
    x, w = ctx.saved_tensors
    w = ctx.p_w
    need_dw = ctx.needs_input_grad[1]
    mg = _main_grad(w) if need_dw else None
    dx, dw = K.rmsnorm_bwd(x.contiguous(), w, dy.contiguous(), ctx.eps, need_dw=need_dw, dw_out=mg[0] if mg else None)
    if mg:
        _notify_ready([w])
        dw = None
    return dx, dw, None


rmsnorm.register_autograd(_rmsnorm_bwd, setup_context=_rmsnorm_setup)


class _RMSNormWithResidual(torch.autograd.Function):
    """y = rmsnorm(x) and x itself handed on as the residual stream. In a pre-norm block x feeds both
    the norm and the residual add; autograd would sum the two gradient contributions with a separate
    add kernel, here the norm's backward kernel adds the residual-stream gradient while it writes dx
    (b200_rmsnorm_bwd_bf16's dres_in)."""

    @staticmethod
    def forward(ctx, x, weight, eps):
        y = K.rmsnorm_fwd(x.contiguous(), weight, eps)
        ctx.save_for_backward(x, weight)
        ctx.eps = eps
        ctx.p_w = weight
        ctx.set_materialize_grads(False)
        return y, x.view_as(x)

    @staticmethod
    def backward(ctx, dy, dres):
        x, w = ctx.saved_tensors
        w = ctx.p_w
        need_dw = ctx.needs_input_grad[1]
        if dy is None:
            return dres, None, None
        mg = _main_grad(w) if need_dw else None
        dx, dw = K.rmsnorm_bwd(x.contiguous(), w, dy.contiguous(), ctx.eps,
                               dres=None if dres is None else dres.contiguous(), need_dw=need_dw,
                               dw_out=mg[0] if mg else None)
        if mg:
            _notify_ready([w])
            dw = None
        return dx, dw, None


def rmsnorm_with_residual(x, weight, eps):
    """-> (rmsnorm(x), x as residual stream); see _RMSNormWithResidual."""
    if x.dtype != BF16:
        return rmsnorm(x, weight, eps), x
    return _RMSNormWithResidual.apply(x, weight, eps)


# ======================================================================================
# LayerNorm (nn.LayerNorm in Models/GPT2/GPT2.py:79-80,105)
# ======================================================================================
class _LayerNormWithResidual(torch.autograd.Function):
    """y = layernorm(x) and x itself handed on as the residual stream (GPT2.py:83-87: x feeds the norm and the shortcut
    add): the norm's backward kernel adds the residual-stream gradient while it writes dx, like _RMSNormWithResidual."""

    @staticmethod
    def forward(ctx, x, weight, bias, eps):
        y, mean, rstd = K.layernorm_fwd(x.contiguous(), weight, bias, eps)
        ctx.save_for_backward(x, mean, rstd)
        ctx.p_w, ctx.p_b = weight, bias
        ctx.set_materialize_grads(False)
        return y, x.view_as(x)

    @staticmethod
    def backward(ctx, dy, dres):
        x, mean, rstd = ctx.saved_tensors
        w, b = ctx.p_w, ctx.p_b
        if dy is None:
            return dres, None, None, None
        need_w, need_b = ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        mg = _main_grad(w, b) if (need_w and need_b) else None
        dx, dw, db = K.layernorm_bwd(x.contiguous(), w, dy.contiguous(), mean, rstd, need_dw=need_w or need_b,
                                     dw_out=mg[0] if mg else None, db_out=mg[1] if mg else None,
                                     dres=None if dres is None else dres.contiguous())
        if mg:
            _notify_ready([w, b])
            return dx, None, None, None
        if need_w or need_b:
            dw, db = dw.to(w.dtype), db.to(w.dtype)
        return dx, (dw if need_w else None), (db if need_b else None), None


def layernorm_with_residual(x, weight, bias, eps):
    """-> (layernorm(x), x as residual stream); see _LayerNormWithResidual."""
    if x.dtype != BF16:
        return layernorm(x, weight, bias, eps)[0], x
    return _LayerNormWithResidual.apply(x, weight, bias, eps)


@torch.library.custom_op("b200llm::layernorm", mutates_args=())
def layernorm(x: Tensor, weight: Tensor, bias: Tensor, eps: float) -> Tuple[Tensor, Tensor, Tensor]:
    return K.layernorm_fwd(x.contiguous(), weight, bias, eps)


@layernorm.register_fake
def _(x, weight, bias, eps):
    rows = x.numel() // x.shape[-1]
    return (torch.empty_like(x), x.new_empty((rows,), dtype=torch.float32),
            x.new_empty((rows,), dtype=torch.float32))


def _layernorm_setup(ctx, inputs, output):
    ctx.set_materialize_grads(False)
    x, w, b, eps = inputs
    _, mean, rstd = output
    ctx.save_for_backward(x, w, mean, rstd)
    ctx.bias_ref = b
    ctx.p_w = w


def _layernorm_bwd(ctx, dy, _dm, _dr):
    if dy is None:
        return None, None, None, None
    x, w, mean, rstd = ctx.saved_tensors
    w = ctx.p_w
    need_dw = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
    mg = _main_grad(w, ctx.bias_ref) if (ctx.needs_input_grad[1] and ctx.needs_input_grad[2]) else None
    dx, dw, db = K.layernorm_bwd(x.contiguous(), w, dy.contiguous(), mean, rstd, need_dw=need_dw,
                                 dw_out=mg[0] if mg else None, db_out=mg[1] if mg else None)
    if mg:
        _notify_ready([w, ctx.bias_ref])
        return dx, None, None, None
    if need_dw:
        dw, db = dw.to(w.dtype), db.to(w.dtype)
    return dx, dw, db, None


layernorm.register_autograd(_layernorm_bwd, setup_context=_layernorm_setup)


# ======================================================================================
# Linear / LinearWithLoRA  (lora.py:26,46; every nn.Linear of the two model families)
# ======================================================================================
@torch.library.custom_op("b200llm::lora_linear", mutates_args=())
def lora_linear(x: Tensor, weight: Tensor, bias: Optional[Tensor], A: Optional[Tensor], B: Optional[Tensor],
                scaling: float, residual: Optional[Tensor]) -> Tuple[Tensor, Tensor]:
    """y = x W^T (+ bias) (+ scaling (x A) B) (+ residual) -> (y, t) with t = scaling * x A."""
    x2 = _flat2d(x)
    out_f = weight.shape[0]
    res2 = None if residual is None else _flat2d(residual)
    if x.dtype == F32:
        y, t = K32.linear_fwd(x2, weight, bias, A, B, scaling, res2)
        return y.view(*x.shape[:-1], out_f), (t if t is not None else x2.new_empty((0,)))
    if A is not None:
        # A [in, r] and B [r, out] are read as MN-major operands: no transposes are materialised
        t = K.gemm_ex(x2, A, b1_mn=True, alpha=scaling)
        y = K.gemm_ex(x2, weight, a2=t, b2=_pitch_ok(B), b2_mn=True, bias=bias, residual=res2)
    else:
        t = x2.new_empty((0,))
        y = K.gemm_ex(x2, weight, bias=bias, residual=res2)
    return y.view(*x.shape[:-1], out_f), t


@lora_linear.register_fake
def _(x, weight, bias, A, B, scaling, residual):
    rows = x.numel() // x.shape[-1]
    y = x.new_empty((*x.shape[:-1], weight.shape[0]))
    t = x.new_empty((rows, A.shape[1])) if A is not None else x.new_empty((0,))
    return y, t


def _lora_linear_setup(ctx, inputs, output):
    ctx.set_materialize_grads(False)
    x, weight, bias, A, B, scaling, residual = inputs
    _, t = output
    ctx.save_for_backward(x, weight, A, B, t)
    ctx.scaling = scaling
    ctx.has_bias = bias is not None
    ctx.has_res = residual is not None
    ctx.bias_ref = bias
    ctx.p_w, ctx.p_A, ctx.p_B = weight, A, B


# Weight-gradient GEMMs into main_grad accumulators are off the critical path of backward: they are
# queued and issued as grouped launches (kernels.gemm_grouped) — one persistent kernel for up to 16
# of these skinny problems instead of one ~20 us launch each.
_wgrad_queue: list = []
WGRAD_FLUSH_AT = 16


# ddp.B200DDP wrappers subscribe here PER PARAMETER (several models / wrappers may live in one process): the
# subscriber of a parameter is called with the parameters whose weight-gradient GEMMs have just been launched
# into main_grad (their gradient is final for this backward once those kernels ran).
_grad_ready_cbs: dict = {}      # id(parameter) -> callback


def register_grad_ready(params, cb) -> None:
    for p in params:
        _grad_ready_cbs[id(p)] = cb


def unregister_grad_ready(params) -> None:
    for p in params:
        _grad_ready_cbs.pop(id(p), None)


def _notify_ready(params) -> None:
    if not _grad_ready_cbs:
        return
    by_cb: dict = {}
    for p in params:
        cb = _grad_ready_cbs.get(id(p))
        if cb is not None:
            by_cb.setdefault(id(cb), (cb, []))[1].append(p)
    for cb, ps in by_cb.values():
        cb(ps)


def flush_weight_grads():
    """Issue every queued weight-gradient GEMM (called before the optimizer / all-reduce read main_grad). The queue
    is shared by everything in the process that runs on the current stream: flushing it early is always correct."""
    global _wgrad_queue
    if _wgrad_queue:
        q, _wgrad_queue = _wgrad_queue, []
        K.gemm_grouped(q)
        _notify_ready([e["param"] for e in q])


def _weight_grad(a_mn_src: Tensor, b_mn_src: Tensor, param: Tensor, trans_out: bool = False):
    """param-shaped gradient = a_mn_src^T · b_mn_src (both stored [tokens, *])."""
    mg = getattr(param, "main_grad", None)
    if mg is not None:
        if b_mn_src.shape[1] <= 64:      # GEMM N = width of the second operand (the LoRA rank)
            _wgrad_queue.append(dict(a=a_mn_src, b=b_mn_src, out=mg, a_mn=True, b_mn=True, accumulate=True,
                                     trans_out=trans_out, param=param))
            if len(_wgrad_queue) >= WGRAD_FLUSH_AT:
                flush_weight_grads()
        elif getattr(param, "_b200_grad_fresh", False) and not trans_out:
            # first (usually only) contribution since zero_grad: a plain fp32 store through the CTA-pair
            # kernel instead of split-K atomics into a zeroed buffer (full-training dW = dY^T x)
            param._b200_grad_fresh = False
            K.gemm_ex(a_mn_src, b_mn_src, a1_mn=True, b1_mn=True, out=mg)
        else:
            K.gemm_ex(a_mn_src, b_mn_src, a1_mn=True, b1_mn=True, out=mg, accumulate=True, trans_out=trans_out)
        if b_mn_src.shape[1] > 64:
            _notify_ready([param])
        return None
    g32 = torch.zeros(param.shape, device=param.device, dtype=torch.float32)
    K.gemm_ex(a_mn_src, b_mn_src, a1_mn=True, b1_mn=True, out=g32, accumulate=True, trans_out=trans_out)
    return K.cast(g32, param.dtype) if param.dtype == BF16 else g32


def _lora_linear_bwd(ctx, dy, _dt):
    if dy is None:
        return (None,) * 7
    x, weight, A, B, t = ctx.saved_tensors
    wref, bref, aref, bbref = ctx.p_w, ctx.bias_ref, ctx.p_A, ctx.p_B     # the Parameters (see _main_grad)
    need = ctx.needs_input_grad
    x2 = _flat2d(x)
    dy2 = _flat2d(dy)
    if dy2.dtype == F32:
        dx, dW, dbias, dA, dB = K32.linear_bwd(dy2.contiguous(), x2, weight, ctx.bias_ref, A, B, t, ctx.scaling, need[0])
        return ((None if dx is None else dx.view(x.shape)), dW, dbias, dA, dB, None,
                (dy if (ctx.has_res and need[6]) else None))
    dx = dW = dbias = dA = dB = None
    u = None
    if A is not None and (need[0] or need[3]):
        u = K.gemm_ex(dy2, _pitch_ok(B), alpha=ctx.scaling)           # s * dY B^T   [M, r]
    if need[0]:
        if u is not None:
            dx2 = K.gemm_ex(dy2, weight, b1_mn=True, a2=u, b2=A)      # dY W + u A^T
        else:
            dx2 = K.gemm_ex(dy2, weight, b1_mn=True)
        dx = dx2.view(x.shape)
    if A is not None and need[3]:
        dA = _weight_grad(x2, u, aref)                                # x^T u          [in, r]
    if B is not None and need[4]:
        dB = _weight_grad(dy2, t, bbref, trans_out=True)              # (dY^T t)^T     [r, out]
    if need[1]:
        dW = _weight_grad(dy2, x2, wref)                              # dY^T x         [out, in]
    if ctx.has_bias and need[2]:
        mg = _main_grad(bref)
        if mg:
            K.colsum(dy2, out=mg[0])
            _notify_ready([bref])
        else:
            dbias = K.colsum(dy2, bref)
    dres = dy if (ctx.has_res and need[6]) else None
    return dx, dW, dbias, dA, dB, None, dres


lora_linear.register_autograd(_lora_linear_bwd, setup_context=_lora_linear_setup)


# --------------------------------------------------------------------------------------
# several bias-free (LoRA) linears of the SAME input: q/k/v of an attention block, fc1/fc2 of SwiGLU
# --------------------------------------------------------------------------------------
@torch.library.custom_op("b200llm::lora_linear_multi", mutates_args=())
def lora_linear_multi(x: Tensor, weights: Sequence[Tensor], As: Sequence[Tensor], Bs: Sequence[Tensor],
                      scaling: float) -> List[Tensor]:
    """[x W_i^T + scaling (x A_i) B_i for i] + [t_i = scaling x A_i for i] (As/Bs empty: plain linears).
    The rank-r projections of all members run as ONE grouped launch (x is streamed by one
    persistent kernel instead of len(weights) separate ~17 us launches)."""
    x2 = _flat2d(x)
    n = len(weights)
    ts: List[Tensor] = []
    if len(As) > 0:
        ts = [x2.new_empty((x2.shape[0], A.shape[1])) for A in As]
        K.gemm_grouped([dict(a=x2, b=A, out=t, b_mn=True, alpha=scaling) for A, t in zip(As, ts)])
        ys = [K.gemm_ex(x2, weights[i], a2=ts[i], b2=_pitch_ok(Bs[i]), b2_mn=True) for i in range(n)]
    else:
        ys = [K.gemm_ex(x2, weights[i]) for i in range(n)]
    return [y.view(*x.shape[:-1], y.shape[-1]) for y in ys] + ts


@lora_linear_multi.register_fake
def _(x, weights, As, Bs, scaling):
    rows = x.numel() // x.shape[-1]
    ys = [x.new_empty((*x.shape[:-1], w.shape[0])) for w in weights]
    return ys + [x.new_empty((rows, A.shape[1])) for A in As]


def _llm_setup(ctx, inputs, output):
    ctx.set_materialize_grads(False)
    x, weights, As, Bs, scaling = inputs
    n = len(weights)
    ctx.n, ctx.lora = n, len(As) > 0
    ctx.scaling = scaling
    ctx.p_weights, ctx.p_As, ctx.p_Bs = list(weights), list(As), list(Bs)
    ctx.save_for_backward(x, *weights, *As, *Bs, *output[n:])


def _llm_bwd(ctx, grads):
    n = ctx.n
    saved = ctx.saved_tensors
    x = saved[0]
    weights, As, Bs = ctx.p_weights, ctx.p_As, ctx.p_Bs          # the Parameters themselves (see _main_grad)
    ts = saved[1 + 3 * n:1 + 4 * n] if ctx.lora else ()
    dys = [None if g is None else _flat2d(g) for g in grads[:n]]
    live = [i for i in range(n) if dys[i] is not None]
    need_x = ctx.needs_input_grad[0]           # (needs_input_grad is flattened over the list arguments)
    need_w = need_b = True
    need_a = ctx.lora and any(A.requires_grad for A in As)
    x2 = _flat2d(x)
    us = {}
    if ctx.lora and live and (need_x or need_a):
        for i in live:
            us[i] = dys[i].new_empty((dys[i].shape[0], Bs[i].shape[0]))
        K.gemm_grouped([dict(a=dys[i], b=_pitch_ok(Bs[i]), out=us[i], alpha=ctx.scaling) for i in live])
    dx = None
    if need_x:
        for i in live:   # dX accumulates through the residual epilogue: no separate add kernels
            kw = dict(a2=us[i], b2=As[i]) if ctx.lora else {}
            dx = K.gemm_ex(dys[i], weights[i], b1_mn=True, residual=dx, out=dx, **kw)
        if dx is not None:
            dx = dx.view(x.shape)
    dWs = [None] * n
    dAs = [None] * n if ctx.lora else []
    dBs = [None] * n if ctx.lora else []
    for i in live:
        if ctx.lora and need_a and As[i].requires_grad:
            dAs[i] = _weight_grad(x2, us[i], As[i])
        if ctx.lora and need_b and Bs[i].requires_grad:
            dBs[i] = _weight_grad(dys[i], ts[i], Bs[i], trans_out=True)
        if need_w and weights[i].requires_grad:
            dWs[i] = _weight_grad(dys[i], x2, weights[i])
    return dx, dWs, dAs, dBs, None


lora_linear_multi.register_autograd(_llm_bwd, setup_context=_llm_setup)


def linears_of_same_input(x, modules):
    """Apply several B200Linear / LinearWithLoRA modules to the same x; fused when they are all
    bias-free and of the same kind, otherwise one op each."""
    from .lora import LinearWithLoRA
    lora = [isinstance(m, LinearWithLoRA) for m in modules]
    bases = [m.linear if l else m for m, l in zip(modules, lora)]
    if x.dtype != BF16 or any(b.bias is not None for b in bases) or (any(lora) and not all(lora)):
        return [m(x) for m in modules]
    n = len(modules)
    if all(lora):
        scal = modules[0].lora.scaling
        if any(m.lora.scaling != scal for m in modules):
            return [m(x) for m in modules]
        out = lora_linear_multi(x, [b.weight for b in bases], [m.lora.A for m in modules],
                                [m.lora.B for m in modules], scal)
    else:
        out = lora_linear_multi(x, [b.weight for b in bases], [], [], 0.0)
    return out[:n]


# --------------------------------------------------------------------------------------
# The same family with ONE packed output [.., sum(out_i)] (q|k|v, fc1|fc2): the consumers
# (attention_qkv, swiglu_packed) read column ranges of it in place and hand back ONE packed gradient,
# so backward is a single dX GEMM over K = sum(out_i) (+ n*r for LoRA) against the row-concatenated
# frozen weights instead of n GEMMs chained through residual-accumulate epilogues.
# --------------------------------------------------------------------------------------
_wcat_cache: dict = {}


def _derived_weight(cache: dict, weights, build):
    """A tensor derived from FROZEN weights (their concatenation, an interleaving), built once and rebuilt when
    any of them is written to (load_state_dict bumps the version counters). Entries are tied to the weight
    OBJECTS through weak references — a freed weight's address may be handed to another tensor with the same
    version count — and leave the cache when a weight dies, so dropped models do not pin their copies."""
    key = tuple(id(w) for w in weights)
    state = tuple((w.data_ptr(), w._version) for w in weights)
    hit = cache.get(key)
    if hit is not None and hit[0] == state and all(r() is w for r, w in zip(hit[2], weights)):
        return hit[1]
    with torch.no_grad():
        out = build([w.detach() for w in weights])
    refs = tuple(weakref.ref(w, lambda _r, k=key, c=cache: c.pop(k, None)) for w in weights)
    cache[key] = (state, out, refs)
    return out


def _adjacent_rows_view(weights):
    """The [sum(out_i), in] matrix that the weights ALREADY form in memory, or None. optim.B200AdamW re-points every
    trainable parameter into one flat buffer in registration order, so W_query / W_key / W_value (and fc1 / fc2) of a
    block sit back to back: their row-concatenation is a zero-copy view that is always current."""
    w0 = weights[0]
    if any(w.dim() != 2 or not w.is_contiguous() or w.shape[1] != w0.shape[1] or w.dtype != w0.dtype for w in weights):
        return None
    esz = w0.element_size()
    for a, b in zip(weights[:-1], weights[1:]):
        if b.data_ptr() != a.data_ptr() + a.numel() * esz or b.untyped_storage().data_ptr() != a.untyped_storage().data_ptr():
            return None
    rows = sum(w.shape[0] for w in weights)
    return torch.empty(0, dtype=w0.dtype, device=w0.device).set_(w0.untyped_storage(), w0.storage_offset(),
                                                               (rows, w0.shape[1]), (w0.shape[1], 1))


def _packed_weight(weights):
    """Row-concatenation of the weights: a cached copy for frozen weights; for trainable ones the in-place view of
    the optimizer's flat buffer when they are adjacent there (else None)."""
    if any(w.requires_grad for w in weights):
        return _adjacent_rows_view([w.detach() for w in weights])
    return _derived_weight(_wcat_cache, weights, lambda ws: torch.cat(ws, dim=0).contiguous())


@torch.library.custom_op("b200llm::lora_linear_packed", mutates_args=())
def lora_linear_packed(x: Tensor, weights: Sequence[Tensor], As: Sequence[Tensor], Bs: Sequence[Tensor],
                       scaling: float, w_cat: Optional[Tensor]) -> List[Tensor]:
    """-> [y_packed [.., sum(out_i)], t_cat [M, n*r]] with y_packed[.., off_i:off_i+out_i] =
    x W_i^T + scaling (x A_i) B_i and t_cat[:, i*r:(i+1)*r] = scaling x A_i."""
    x2 = _flat2d(x)
    M = x2.shape[0]
    n = len(weights)
    total = sum(w.shape[0] for w in weights)
    y = x2.new_empty((M, total))
    if len(As) > 0:
        r = As[0].shape[1]
        t_cat = x2.new_empty((M, n * r))
        K.gemm_grouped([dict(a=x2, b=As[i], out=t_cat[:, i * r:(i + 1) * r], b_mn=True, alpha=scaling) for i in range(n)])
    else:
        t_cat = x2.new_empty((0,))
    starts = [0]
    for w in weights[:-1]:
        starts.append(starts[-1] + w.shape[0])
    if (w_cat is not None and len(As) > 0 and n <= 4 and M >= 256 and all(st % 256 == 0 for st in starts)
            and all(B.shape[1] % 8 == 0 for B in Bs) and os.environ.get("B200_PACKED_FWD", "1") != "0"):
        # one GEMM for all members: the tile's segment picks its t block, B's sit side by side
        K.gemm_lora_packed(x2, w_cat, t_cat, torch.cat(list(Bs), dim=1), r, starts, out=y)
        return [y.view(*x.shape[:-1], total), t_cat]
    if w_cat is not None and len(As) == 0:
        K.gemm_ex(x2, w_cat, out=y)           # plain linears of one input against their concatenation: one GEMM
        return [y.view(*x.shape[:-1], total), t_cat]
    off = 0
    for i, w in enumerate(weights):
        o = y[:, off:off + w.shape[0]]
        if len(As) > 0:
            K.gemm_ex(x2, w, a2=t_cat[:, i * r:(i + 1) * r], b2=_pitch_ok(Bs[i]), b2_mn=True, out=o)
        else:
            K.gemm_ex(x2, w, out=o)
        off += w.shape[0]
    return [y.view(*x.shape[:-1], total), t_cat]


@lora_linear_packed.register_fake
def _(x, weights, As, Bs, scaling, w_cat):
    rows = x.numel() // x.shape[-1]
    total = sum(w.shape[0] for w in weights)
    t = x.new_empty((rows, len(As) * As[0].shape[1])) if len(As) > 0 else x.new_empty((0,))
    return [x.new_empty((*x.shape[:-1], total)), t]


def _llp_setup(ctx, inputs, output):
    ctx.set_materialize_grads(False)
    x, weights, As, Bs, scaling, w_cat = inputs
    ctx.n, ctx.lora, ctx.scaling = len(weights), len(As) > 0, scaling
    ctx.has_cat = w_cat is not None
    ctx.p_weights, ctx.p_As, ctx.p_Bs = list(weights), list(As), list(Bs)
    ctx.save_for_backward(x, *weights, *As, *Bs, output[1], *([w_cat] if w_cat is not None else []))


def _llp_bwd(ctx, grads):
    n = ctx.n
    saved = ctx.saved_tensors
    x = saved[0]
    weights, As, Bs = ctx.p_weights, ctx.p_As, ctx.p_Bs          # the Parameters themselves (see _main_grad)
    pos = 1 + (3 * n if ctx.lora else n)
    t_cat = saved[pos]
    w_cat = saved[pos + 1] if ctx.has_cat else None
    none = (None, [None] * n, [None] * len(As), [None] * len(Bs), None, None)
    if grads[0] is None:
        return none
    dy = _flat2d(grads[0])
    if dy.stride(1) != 1:
        dy = dy.contiguous()
    x2 = _flat2d(x)
    offs = [0]
    for w in weights:
        offs.append(offs[-1] + w.shape[0])
    dys = [dy[:, offs[i]:offs[i + 1]] for i in range(n)]
    need_x = ctx.needs_input_grad[0]
    need_a = ctx.lora and any(A.requires_grad for A in As)
    u_cat, r = None, 0
    if ctx.lora and (need_x or need_a):
        r = As[0].shape[1]
        u_cat = dy.new_empty((dy.shape[0], n * r))
        K.gemm_grouped([dict(a=dys[i], b=_pitch_ok(Bs[i]), out=u_cat[:, i * r:(i + 1) * r], alpha=ctx.scaling)
                        for i in range(n)])
    dx = None
    if need_x:
        if w_cat is not None:
            # dX = dY_packed W_cat + u_cat A_cat^T : one launch over K = sum(out_i) + n*r
            if ctx.lora:
                a_cat = torch.cat(list(As), dim=1)
                dx = K.gemm_ex(dy, w_cat, b1_mn=True, a2=u_cat, b2=a_cat)
            else:
                dx = K.gemm_ex(dy, w_cat, b1_mn=True)
        else:
            for i in range(n):
                kw = dict(a2=u_cat[:, i * r:(i + 1) * r], b2=As[i]) if ctx.lora else {}
                dx = K.gemm_ex(dys[i], weights[i], b1_mn=True, residual=dx, out=dx, **kw)
        dx = dx.view(x.shape)
    dWs = [None] * n
    dAs = [None] * n if ctx.lora else []
    dBs = [None] * n if ctx.lora else []
    if not ctx.lora and all(w.requires_grad and getattr(w, "_b200_grad_fresh", False) for w in weights):
        # full fine-tuning: the members' main_grad accumulators are adjacent in the flat gradient buffer exactly like
        # the weights, so dW of the whole group is one store-mode GEMM dY_packed^T x
        mgs = _main_grad(*weights)
        mg_cat = _adjacent_rows_view(mgs) if mgs else None
        if mg_cat is not None:
            K.gemm_ex(dy, x2, a1_mn=True, b1_mn=True, out=mg_cat)
            for w in weights:
                w._b200_grad_fresh = False
            _notify_ready(list(weights))
            return dx, dWs, dAs, dBs, None, None
    for i in range(n):
        if ctx.lora and need_a and As[i].requires_grad:
            dAs[i] = _weight_grad(x2, u_cat[:, i * r:(i + 1) * r], As[i])
        if ctx.lora and Bs[i].requires_grad:
            dBs[i] = _weight_grad(dys[i], t_cat[:, i * r:(i + 1) * r], Bs[i], trans_out=True)
        if weights[i].requires_grad:
            dWs[i] = _weight_grad(dys[i], x2, weights[i])
    return dx, dWs, dAs, dBs, None, None


lora_linear_packed.register_autograd(_llp_bwd, setup_context=_llp_setup)


# --------------------------------------------------------------------------------------
# q | k | v projection + RoPE + causal attention as ONE op (Models/Llama/Llama3.py:135-153): compute_rope runs in the
# epilogue of the packed projection GEMM and its transpose in the epilogues of the attention-backward kernels, so neither
# the rotation (one 4-bytes-per-element pass over q | k) nor its gradient (another) exists as a kernel.
# --------------------------------------------------------------------------------------
_rope_ok_cache: dict = {}


def rope_tables_fusable(cos: Tensor, sin: Tensor) -> bool:
    """The fused rotation reads one half of each table row for both halves of a head: allowed when the two column halves
    are equal — which precompute_rope_params guarantees (angles = cat([a, a])); checked once per table."""
    key = (cos.data_ptr(), sin.data_ptr(), cos._version, sin._version, tuple(cos.shape))
    hit = _rope_ok_cache.get(key)
    if hit is None:
        half = cos.shape[1] // 2
        hit = bool(cos.dtype == BF16 and sin.dtype == BF16 and cos.is_contiguous() and sin.is_contiguous()
                   and torch.equal(cos[:, :half], cos[:, half:]) and torch.equal(sin[:, :half], sin[:, half:]))
        if len(_rope_ok_cache) > 64:
            _rope_ok_cache.clear()
        _rope_ok_cache[key] = hit
    return hit


@torch.library.custom_op("b200llm::lora_qkv_rope_attention", mutates_args=())
def lora_qkv_rope_attention(x: Tensor, weights: Sequence[Tensor], As: Sequence[Tensor], Bs: Sequence[Tensor], scaling: float,
                            w_cat: Tensor, cos: Tensor, sin: Tensor, B: int, T: int, h: int, hkv: int,
                            hd: int) -> List[Tensor]:
    """-> [o [B*T, h*hd], qkv_rot [B*T, (h + 2 hkv) hd] (q, k rotated), lse, t_cat]."""
    x2 = _flat2d(x)
    M = x2.shape[0]
    n = len(weights)
    starts = [0]
    for w in weights[:-1]:
        starts.append(starts[-1] + w.shape[0])
    if len(As) > 0:
        r = As[0].shape[1]
        t_cat = x2.new_empty((M, n * r))
        K.gemm_grouped([dict(a=x2, b=As[i], out=t_cat[:, i * r:(i + 1) * r], b_mn=True, alpha=scaling) for i in range(n)])
        qkv = K.gemm_qkv_rope(x2, w_cat, t_cat, torch.cat(list(Bs), dim=1), r, starts, cos, sin, T, hd, (h + hkv) * hd)
    else:
        t_cat = x2.new_empty((0,))
        qkv = K.gemm_qkv_rope(x2, w_cat, None, None, 0, starts, cos, sin, T, hd, (h + hkv) * hd)
    nq, nk = h * hd, hkv * hd
    o, lse = K.attention_fwd(qkv[:, :nq], qkv[:, nq:nq + nk], qkv[:, nq + nk:], B, T, h, hkv, hd)
    return [o, qkv, lse, t_cat]


@lora_qkv_rope_attention.register_fake
def _(x, weights, As, Bs, scaling, w_cat, cos, sin, B, T, h, hkv, hd):
    rows = x.numel() // x.shape[-1]
    t = x.new_empty((rows, len(As) * As[0].shape[1])) if len(As) > 0 else x.new_empty((0,))
    return [x.new_empty((rows, h * hd)), x.new_empty((rows, (h + 2 * hkv) * hd)),
            x.new_empty((B, h, T), dtype=torch.float32), t]


def _lqra_setup(ctx, inputs, output):
    ctx.set_materialize_grads(False)
    x, weights, As, Bs, scaling, w_cat, cos, sin, B, T, h, hkv, hd = inputs
    o, qkv, lse, t_cat = output
    ctx.n, ctx.lora, ctx.scaling, ctx.has_cat = len(weights), len(As) > 0, scaling, True
    ctx.dims = (B, T, h, hkv, hd)
    ctx.p_weights, ctx.p_As, ctx.p_Bs = list(weights), list(As), list(Bs)
    # layout expected by _llp_bwd: x, weights, As, Bs, t_cat, w_cat — then this op's own tensors
    ctx.save_for_backward(x, *weights, *As, *Bs, t_cat, w_cat, qkv, o, lse, cos, sin)


def _lqra_bwd(ctx, grads):
    do = grads[0]
    n = ctx.n
    n_in = 13
    if do is None:
        return (None, [None] * n, [None] * (n if ctx.lora else 0), [None] * (n if ctx.lora else 0)) + (None,) * (n_in - 4)
    saved = ctx.saved_tensors
    qkv, o, lse, cos, sin = saved[-5:]
    B, T, h, hkv, hd = ctx.dims
    nq, nk = h * hd, hkv * hd
    dqkv = qkv.new_empty(qkv.shape)
    # dq / dk come out of the attention-backward epilogues already multiplied by the transposed rotation
    K.attention_bwd(qkv[:, :nq], qkv[:, nq:nq + nk], qkv[:, nq + nk:], o, _flat2d(do), lse, B, T, h, hkv, hd,
                    dq=dqkv[:, :nq], dk=dqkv[:, nq:nq + nk], dv=dqkv[:, nq + nk:], rope=(cos, sin))
    out = _llp_bwd(_BwdView(ctx, saved[:-5]), [dqkv.view(*saved[0].shape[:-1], dqkv.shape[-1]), None])
    return out + (None,) * (n_in - len(out))


class _BwdView:
    """ctx stand-in handing _llp_bwd the packed-projection part of another op's saved tensors."""

    def __init__(self, ctx, saved):
        self.n, self.lora, self.scaling, self.has_cat = ctx.n, ctx.lora, ctx.scaling, ctx.has_cat
        self.p_weights, self.p_As, self.p_Bs = ctx.p_weights, ctx.p_As, ctx.p_Bs
        self.saved_tensors = saved
        self.needs_input_grad = ctx.needs_input_grad


lora_qkv_rope_attention.register_autograd(_lqra_bwd, setup_context=_lqra_setup)


def qkv_rope_attention(x, modules, cos, sin, B, T, h, hkv, hd):
    """Context vectors [B*T, h*hd] of causal attention over RoPE'd q, k from the three projection modules of x, through
    lora_qkv_rope_attention; None when the fused form does not apply (caller falls back to separate ops)."""
    if os.environ.get("B200_PACKED", "1") == "0" or os.environ.get("B200_ROPE_EPILOGUE", "1") == "0":
        return None
    from .lora import LinearWithLoRA
    lora = [isinstance(m, LinearWithLoRA) for m in modules]
    bases = [m.linear if l else m for m, l in zip(modules, lora)]
    if any(b.bias is not None for b in bases) or (any(lora) and not all(lora)) or hd not in (64, 128):
        return None
    weights = [b.weight for b in bases]
    rows = x.numel() // x.shape[-1]
    if rows < 256 or any(w.shape[0] % 256 for w in weights) or x.dtype != BF16 or not rope_tables_fusable(cos, sin):
        return None
    w_cat = _packed_weight(weights)
    if w_cat is None:
        return None
    if all(lora):
        scal, r = modules[0].lora.scaling, modules[0].lora.A.shape[1]
        if any(m.lora.scaling != scal or m.lora.A.shape[1] != r for m in modules) or r % 8 or r > 64:
            return None
        if any(m.lora.B.shape[1] % 8 for m in modules):
            return None
        return lora_qkv_rope_attention(x, weights, [m.lora.A for m in modules], [m.lora.B for m in modules], scal, w_cat,
                                       cos, sin, B, T, h, hkv, hd)[0]
    return lora_qkv_rope_attention(x, weights, [], [], 0.0, w_cat, cos, sin, B, T, h, hkv, hd)[0]


# fc1 | fc2 with SwiGLU in the GEMM epilogue: same packed (a | b) output and same backward as
# lora_linear_packed, plus h = silu(a) * b straight from the accumulators (no separate 6-bytes-per-element
# SwiGLU kernel). Needs the frozen weights interleaved in 128-row blocks (cached like the concatenation).
_wil_cache: dict = {}


def _interleaved_weight(weights, blk=128):
    if len(weights) != 2 or any(w.requires_grad for w in weights) or weights[0].shape != weights[1].shape:
        return None
    F, Kd = weights[0].shape
    if F % blk:
        return None
    return _derived_weight(
        _wil_cache, weights,
        lambda ws: torch.stack([ws[0].view(F // blk, blk, Kd), ws[1].view(F // blk, blk, Kd)], dim=1)
        .reshape(2 * F, Kd).contiguous())


def _interleaved_lora_b(B1, B2, blk=128):
    """[2r, 2F]: rows 0..r-1 = B1 in the a columns of the 128-block interleaving (zeros in the b columns), rows
    r..2r-1 = B2 in the b columns."""
    r, F = B1.shape
    out = B1.new_zeros((2 * r, F // blk, 2, blk))
    out[:r, :, 0, :] = B1.view(r, F // blk, blk)
    out[r:, :, 1, :] = B2.view(r, F // blk, blk)
    return out.view(2 * r, 2 * F)


@torch.library.custom_op("b200llm::lora_fc12_swiglu", mutates_args=())
def lora_fc12_swiglu(x: Tensor, weights: Sequence[Tensor], As: Sequence[Tensor], Bs: Sequence[Tensor], scaling: float,
                     w_cat: Optional[Tensor], w_il: Tensor) -> List[Tensor]:
    """-> [(a | b) [.., 2F], t_cat [M, 2r], h [.., F]] for the two up projections of a SwiGLU feed-forward."""
    x2 = _flat2d(x)
    M = x2.shape[0]
    F = weights[0].shape[0]
    if len(As) > 0:
        r = As[0].shape[1]
        t_cat = x2.new_empty((M, 2 * r))
        K.gemm_grouped([dict(a=x2, b=As[i], out=t_cat[:, i * r:(i + 1) * r], b_mn=True, alpha=scaling) for i in range(2)])
        ab, h = K.gemm_swiglu_fwd(x2, w_il, t_cat, _interleaved_lora_b(Bs[0], Bs[1]))
    else:
        t_cat = x2.new_empty((0,))
        ab, h = K.gemm_swiglu_fwd(x2, w_il, None, None)
    return [ab.view(*x.shape[:-1], 2 * F), t_cat, h.view(*x.shape[:-1], F)]


@lora_fc12_swiglu.register_fake
def _(x, weights, As, Bs, scaling, w_cat, w_il):
    rows = x.numel() // x.shape[-1]
    F = weights[0].shape[0]
    t = x.new_empty((rows, 2 * As[0].shape[1])) if len(As) > 0 else x.new_empty((0,))
    return [x.new_empty((*x.shape[:-1], 2 * F)), t, x.new_empty((*x.shape[:-1], F))]


def _lfs_setup(ctx, inputs, output):
    _llp_setup(ctx, inputs[:6], output[:2])


def _lfs_bwd(ctx, grads):
    return _llp_bwd(ctx, grads[:2]) + (None,)


lora_fc12_swiglu.register_autograd(_lfs_bwd, setup_context=_lfs_setup)


def fc12_swiglu(x, fc1, fc2):
    """(ab, h) through lora_fc12_swiglu for two bias-free up projections (both plain or both LoRA-wrapped with the
    same rank and scaling, frozen base weights), or None when that does not apply."""
    if x.dtype != BF16 or os.environ.get("B200_PACKED", "1") == "0" or os.environ.get("B200_SWIGLU_FWD_EPILOGUE", "1") == "0":
        return None
    from .lora import LinearWithLoRA
    mods = [fc1, fc2]
    lora = [isinstance(m, LinearWithLoRA) for m in mods]
    bases = [m.linear if l else m for m, l in zip(mods, lora)]
    if any(b.bias is not None for b in bases) or lora[0] != lora[1]:
        return None
    rows = x.numel() // x.shape[-1]
    weights = [b.weight for b in bases]
    w_il = _interleaved_weight(weights)
    if w_il is None or rows < 256:
        return None
    if lora[0]:
        scal, r = fc1.lora.scaling, fc1.lora.A.shape[1]
        if fc2.lora.scaling != scal or fc2.lora.A.shape[1] != r or r % 8 or 2 * r > 128:
            return None
        out = lora_fc12_swiglu(x, weights, [fc1.lora.A, fc2.lora.A], [fc1.lora.B, fc2.lora.B], scal,
                               _packed_weight(weights), w_il)
    else:
        out = lora_fc12_swiglu(x, weights, [], [], 0.0, _packed_weight(weights), w_il)
    return out[0], out[2]


def linears_packed(x, modules):
    """Packed output of several bias-free linears (all plain or all LoRA-wrapped with one scaling and one
    rank) of the same x, or None when they cannot be fused that way."""
    if x.dtype != BF16 or os.environ.get("B200_PACKED", "1") == "0":
        return None
    from .lora import LinearWithLoRA
    lora = [isinstance(m, LinearWithLoRA) for m in modules]
    bases = [m.linear if l else m for m, l in zip(modules, lora)]
    if any(b.bias is not None for b in bases) or (any(lora) and not all(lora)):
        return None
    if any(b.weight.shape[0] % 8 for b in bases):
        return None
    weights = [b.weight for b in bases]
    if all(lora):
        scal, r = modules[0].lora.scaling, modules[0].lora.A.shape[1]
        if any(m.lora.scaling != scal or m.lora.A.shape[1] != r for m in modules) or r % 8:
            return None
        return lora_linear_packed(x, weights, [m.lora.A for m in modules], [m.lora.B for m in modules], scal,
                                  _packed_weight(weights))[0]
    return lora_linear_packed(x, weights, [], [], 0.0, _packed_weight(weights))[0]


def linear(x, weight, bias=None, residual=None):
    return lora_linear(x, weight, bias, None, None, 0.0, residual)[0]


# --------------------------------------------------------------------------------------
# GPT-2 block fusions (Models/GPT2/GPT2.py:58-62, 83-88)
#   linear_dropout_residual:  residual + dropout(linear(x))          -- `x + drop_shortcut(att(...))`, :85
#   gelu_mlp:                 residual + dropout(W2 gelu(W1 x + b1) + b2)   -- `x + drop_shortcut(ff(...))`, :87
# Forward: GELU and dropout + residual run in GEMM epilogues. Backward: the GELU's derivative runs in the epilogue of the
# second linear's input-gradient GEMM; the dropout mask is re-drawn on the incoming gradient by the stand-alone kernel
# (that masked gradient is an operand of three GEMMs, so it has to exist).
# --------------------------------------------------------------------------------------
def _linear_fwd(x2, weight, bias, A, B, scaling, residual=None, drop=None):
    """y = [residual +] [dropout](x W^T + bias + s (x A) B) -> (y, t); drop = (p, seed) or None."""
    t = K.gemm_ex(x2, A, b1_mn=True, alpha=scaling) if A is not None else None
    Bp = _pitch_ok(B) if A is not None else None
    if drop is not None and drop[0] > 0:
        if K.fused_epilogue_ok(x2.shape[0], weight.shape[0]) and residual is not None:
            y = K.gemm_dropout_residual(x2, weight, bias, residual, drop[0], drop[1], _drop_step(x2, drop[0]), t, Bp)
        else:
            y = K.gemm_ex(x2, weight, a2=t, b2=Bp, b2_mn=True, bias=bias)
            y = K.dropout(y, drop[0], drop[1], residual, seed_dev=_drop_step(x2, drop[0]))
    else:
        y = K.gemm_ex(x2, weight, a2=t, b2=Bp, b2_mn=True, bias=bias, residual=residual)
    return y, t


def _linear_bwd(dy2, x2, weight, bias, A, B, t, scaling, need_x, dx_fn=None):
    """Gradients of y = x W^T + bias + s (x A) B given dy: -> (dx or None, dW, dbias, dA, dB). dx_fn(dy2, weight, u, A)
    replaces the plain input-gradient GEMM (an epilogue-fused variant)."""
    dx = dW = dbias = dA = dB = None
    u = None
    lora_live = A is not None and (need_x or A.requires_grad)
    if lora_live:
        u = K.gemm_ex(dy2, _pitch_ok(B), alpha=scaling)                   # s * dY B^T   [M, r]
    if need_x:
        if dx_fn is not None:
            dx = dx_fn(dy2, weight, u, A if u is not None else None)
        elif u is not None:
            dx = K.gemm_ex(dy2, weight, b1_mn=True, a2=u, b2=A)
        else:
            dx = K.gemm_ex(dy2, weight, b1_mn=True)
    if A is not None and A.requires_grad:
        dA = _weight_grad(x2, u, A)
    if B is not None and B.requires_grad:
        dB = _weight_grad(dy2, t, B, trans_out=True)
    if weight.requires_grad:
        dW = _weight_grad(dy2, x2, weight)
    if bias is not None and bias.requires_grad:
        mg = _main_grad(bias)
        if mg:
            K.colsum(dy2, out=mg[0])
            _notify_ready([bias])
        else:
            dbias = K.colsum(dy2, bias)
    return dx, dW, dbias, dA, dB


@torch.library.custom_op("b200llm::linear_dropout_residual", mutates_args=())
def linear_dropout_residual(x: Tensor, weight: Tensor, bias: Optional[Tensor], A: Optional[Tensor], B: Optional[Tensor],
                            scaling: float, residual: Tensor, p: float, seed: int) -> Tuple[Tensor, Tensor]:
    """-> (residual + dropout(x W^T + bias + s (x A) B), t = s x A)."""
    x2 = _flat2d(x)
    y, t = _linear_fwd(x2, weight, bias, A, B, scaling, _flat2d(residual), (p, seed))
    return y.view(*x.shape[:-1], weight.shape[0]), (t if t is not None else x2.new_empty((0,)))


@linear_dropout_residual.register_fake
def _(x, weight, bias, A, B, scaling, residual, p, seed):
    rows = x.numel() // x.shape[-1]
    t = x.new_empty((rows, A.shape[1])) if A is not None else x.new_empty((0,))
    return x.new_empty((*x.shape[:-1], weight.shape[0])), t


def _ldr_setup(ctx, inputs, output):
    ctx.set_materialize_grads(False)
    x, weight, bias, A, B, scaling, residual, p, seed = inputs
    ctx.save_for_backward(x, weight, A, B, output[1])
    ctx.bias_ref = bias
    ctx.p_w, ctx.p_A, ctx.p_B = weight, A, B
    ctx.scaling, ctx.p, ctx.seed = scaling, p, seed


def _ldr_bwd(ctx, dy, _dt):
    if dy is None:
        return (None,) * 9
    x, weight, A, B, t = ctx.saved_tensors
    weight, A, B = ctx.p_w, ctx.p_A, ctx.p_B                     # the Parameters themselves (see _main_grad)
    need = ctx.needs_input_grad
    dy2 = _flat2d(dy)
    dym = K.dropout(dy2, ctx.p, ctx.seed, seed_dev=_drop_step(dy2, ctx.p)) if ctx.p > 0 else dy2
    dx, dW, dbias, dA, dB = _linear_bwd(dym, _flat2d(x), weight, ctx.bias_ref, A, B, t, ctx.scaling, need[0])
    return (None if dx is None else dx.view(x.shape)), dW, dbias, dA, dB, None, (dy if need[6] else None), None, None


linear_dropout_residual.register_autograd(_ldr_bwd, setup_context=_ldr_setup)


@torch.library.custom_op("b200llm::gelu_mlp", mutates_args=())
def gelu_mlp(x: Tensor, w1: Tensor, b1: Optional[Tensor], A1: Optional[Tensor], B1: Optional[Tensor], w2: Tensor,
             b2: Optional[Tensor], A2: Optional[Tensor], B2: Optional[Tensor], scaling: float, residual: Optional[Tensor],
             p: float, seed: int) -> List[Tensor]:
    """[residual + dropout_p](W2 gelu(W1 x + b1) + b2), LoRA pairs optional -> [y, z, h, t1, t2] (z: pre-activation,
    h = gelu(z), t_i = s (input of linear i) A_i; all but y only saved for backward)."""
    x2 = _flat2d(x)
    M = x2.shape[0]
    t1 = K.gemm_ex(x2, A1, b1_mn=True, alpha=scaling) if A1 is not None else None
    if M >= 256 and w1.shape[0] >= 256:
        h, z = K.gemm_gelu_fwd(x2, w1, b1, t1, _pitch_ok(B1) if A1 is not None else None)
    else:
        z = K.gemm_ex(x2, w1, a2=t1, b2=_pitch_ok(B1) if A1 is not None else None, b2_mn=True, bias=b1)
        h = K.gelu_fwd(z)
    res2 = _flat2d(residual) if residual is not None else None
    y, t2 = _linear_fwd(h, w2, b2, A2, B2, scaling, res2, (p, seed) if p > 0 else None)
    # (outputs of a custom op may not alias each other: two separate empties)
    return [y.view(*x.shape[:-1], w2.shape[0]), z, h, t1 if t1 is not None else x2.new_empty((0,)),
            t2 if t2 is not None else x2.new_empty((0,))]


@gelu_mlp.register_fake
def _(x, w1, b1, A1, B1, w2, b2, A2, B2, scaling, residual, p, seed):
    rows = x.numel() // x.shape[-1]
    return [x.new_empty((*x.shape[:-1], w2.shape[0])), x.new_empty((rows, w1.shape[0])), x.new_empty((rows, w1.shape[0])),
            x.new_empty((rows, A1.shape[1])) if A1 is not None else x.new_empty((0,)),
            x.new_empty((rows, A2.shape[1])) if A2 is not None else x.new_empty((0,))]


def _gelu_mlp_setup(ctx, inputs, output):
    ctx.set_materialize_grads(False)
    x, w1, b1, A1, B1, w2, b2, A2, B2, scaling, residual, p, seed = inputs
    ctx.save_for_backward(x, w1, A1, B1, w2, A2, B2, *output[1:])
    ctx.biases = (b1, b2)
    ctx.p_lin = (w1, A1, B1, w2, A2, B2)
    ctx.scaling, ctx.p, ctx.seed = scaling, p, seed
    ctx.has_res = residual is not None


def _gelu_mlp_bwd(ctx, grads):
    dy = grads[0]
    if dy is None:
        return (None,) * 13
    x, w1, A1, B1, w2, A2, B2, z, h, t1, t2 = ctx.saved_tensors
    w1, A1, B1, w2, A2, B2 = ctx.p_lin                           # the Parameters themselves (see _main_grad)
    b1, b2 = ctx.biases
    need = ctx.needs_input_grad
    x2, dy2 = _flat2d(x), _flat2d(dy)
    dym = K.dropout(dy2, ctx.p, ctx.seed, seed_dev=_drop_step(dy2, ctx.p)) if ctx.p > 0 else dy2
    M, F = z.shape

    def dz_fn(g, w, u, A):
        if K.fused_epilogue_ok(M, F):
            return K.gemm_gelu_bwd(g, w, z, u, A)          # dz = gelu'(z) * (g W2 + u A2^T), one launch
        dh = K.gemm_ex(g, w, b1_mn=True, a2=u, b2=A) if u is not None else K.gemm_ex(g, w, b1_mn=True)
        return K.gelu_bwd(z, dh)
    dz, dW2, db2, dA2, dB2 = _linear_bwd(dym, h, w2, b2, A2, B2, t2, ctx.scaling, True, dx_fn=dz_fn)
    dx, dW1, db1, dA1, dB1 = _linear_bwd(dz, x2, w1, b1, A1, B1, t1, ctx.scaling, need[0])
    return ((None if dx is None else dx.view(x.shape)), dW1, db1, dA1, dB1, dW2, db2, dA2, dB2, None,
            (dy if (ctx.has_res and need[10]) else None), None, None)


gelu_mlp.register_autograd(_gelu_mlp_bwd, setup_context=_gelu_mlp_setup)


def linear_operands(mod):
    """(weight, bias, A, B, scaling) of a B200Linear or a LinearWithLoRA."""
    from .lora import LinearWithLoRA
    if isinstance(mod, LinearWithLoRA):
        return mod.linear.weight, mod.linear.bias, mod.lora.A, mod.lora.B, mod.lora.scaling
    return mod.weight, mod.bias, None, None, 0.0


# ======================================================================================
# attention (+ RoPE)  (Models/Llama/Llama3.py:131-153, Models/GPT2/GPT2.py:29-47)
# ======================================================================================
@torch.library.custom_op("b200llm::attention", mutates_args=())
def attention(q: Tensor, k: Tensor, v: Tensor, cos: Optional[Tensor], sin: Optional[Tensor], B: int, T: int,
              h: int, hkv: int, hd: int, dropout_p: float = 0.0, seed: int = 0) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """q [B*T, h*hd], k/v [B*T, hkv*hd] (row-major 2-D views) -> (o, lse, q_rot, k_rot).
    dropout_p > 0 drops attention weights after the softmax (GPT2.py:43) with the counter-based mask
    of `seed`; backward regenerates the same mask."""
    if cos is not None:
        qr = K.rope(q, cos, sin, B, T, h, hd)
        kr = K.rope(k, cos, sin, B, T, hkv, hd)
        o, lse = K.attention_fwd(qr, kr, v, B, T, h, hkv, hd, dropout_p=dropout_p, seed=seed, seed_dev=_drop_step(q, dropout_p))
        return o, lse, qr, kr
    o, lse = K.attention_fwd(q, k, v, B, T, h, hkv, hd, dropout_p=dropout_p, seed=seed, seed_dev=_drop_step(q, dropout_p))
    return o, lse, q.new_empty((0,)), q.new_empty((0,))   # no rotated copies to hand back


@attention.register_fake
def _(q, k, v, cos, sin, B, T, h, hkv, hd, dropout_p=0.0, seed=0):
    o, lse = q.new_empty((B * T, h * hd)), q.new_empty((B, h, T), dtype=torch.float32)
    if cos is None:
        return o, lse, q.new_empty((0,)), q.new_empty((0,))
    return o, lse, q.new_empty((B * T, h * hd)), q.new_empty((B * T, hkv * hd))


def _attention_setup(ctx, inputs, output):
    ctx.set_materialize_grads(False)
    q, k, v, cos, sin, B, T, h, hkv, hd, dropout_p, seed = inputs
    o, lse, qr, kr = output
    if cos is None:
        qr, kr = q, k
    ctx.save_for_backward(qr, kr, v, o, lse, cos, sin)
    ctx.dims = (B, T, h, hkv, hd)
    ctx.drop = (dropout_p, seed)


def _attention_bwd(ctx, do, _dl, _dq, _dk):
    if do is None:
        return (None,) * 12
    qr, kr, v, o, lse, cos, sin = ctx.saved_tensors
    B, T, h, hkv, hd = ctx.dims
    v2 = v if v.stride(-1) == 1 else v.contiguous()
    dq, dk, dv = K.attention_bwd(qr, kr, v2, o, _flat2d(do), lse, B, T, h, hkv, hd, dropout_p=ctx.drop[0],
                                 seed=ctx.drop[1], seed_dev=_drop_step(qr, ctx.drop[0]))
    if cos is not None:
        K.rope_(dq, cos, sin, B, T, h, hd, inverse=True)
        K.rope_(dk, cos, sin, B, T, hkv, hd, inverse=True)
    return (dq, dk, dv) + (None,) * 9


attention.register_autograd(_attention_bwd, setup_context=_attention_setup)


# fp32 parity mode: the same attention with materialised scores in plain fp32 (f32.py); P is saved for backward
@torch.library.custom_op("b200llm::attention_f32", mutates_args=())
def attention_f32(q: Tensor, k: Tensor, v: Tensor, cos: Optional[Tensor], sin: Optional[Tensor], B: int, T: int, h: int,
                  hkv: int, hd: int) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """float32 q [B*T, h*hd], k / v [B*T, hkv*hd] -> (o, P [B, h, T, T], q_rot, k_rot)."""
    if cos is not None:
        qr, kr = K32.rope(q, cos, sin, T, h, hd), K32.rope(k, cos, sin, T, hkv, hd)
    else:
        qr, kr = q.contiguous(), k.contiguous()
    o, P = K32.attention_fwd(qr, kr, v, B, T, h, hkv, hd)
    if cos is None:
        return o, P, q.new_empty((0,)), q.new_empty((0,))
    return o, P, qr, kr


@attention_f32.register_fake
def _(q, k, v, cos, sin, B, T, h, hkv, hd):
    o, P = q.new_empty((B * T, h * hd)), q.new_empty((B, h, T, T))
    if cos is None:
        return o, P, q.new_empty((0,)), q.new_empty((0,))
    return o, P, q.new_empty((B * T, h * hd)), q.new_empty((B * T, hkv * hd))


def _attention_f32_setup(ctx, inputs, output):
    ctx.set_materialize_grads(False)
    q, k, v, cos, sin, B, T, h, hkv, hd = inputs
    o, P, qr, kr = output
    if cos is None:
        qr, kr = q, k
    ctx.save_for_backward(qr, kr, v, P, cos, sin)
    ctx.dims = (B, T, h, hkv, hd)


def _attention_f32_bwd(ctx, do, _dp, _dq, _dk):
    if do is None:
        return (None,) * 10
    qr, kr, v, P, cos, sin = ctx.saved_tensors
    B, T, h, hkv, hd = ctx.dims
    dq, dk, dv = K32.attention_bwd(qr, kr, v, P, _flat2d(do), B, T, h, hkv, hd)
    if cos is not None:
        dq, dk = K32.rope(dq, cos, sin, T, h, hd, inverse=True), K32.rope(dk, cos, sin, T, hkv, hd, inverse=True)
    return (dq, dk, dv) + (None,) * 7


attention_f32.register_autograd(_attention_f32_bwd, setup_context=_attention_f32_setup)


def attend(q, k, v, cos, sin, B, T, h, hkv, hd, dropout_p=0.0, seed=0):
    """Context vectors [B*T, h*hd] of causal attention over q / k (rotated when cos / sin are given) and v: the
    tcgen05 kernels for bf16, the plain-fp32 path for float32 (parity mode; no dropout there)."""
    if q.dtype == F32:
        if dropout_p > 0:
            raise RuntimeError("the fp32 parity mode runs with drop_rate = 0 (SURVEY.md H5)")
        return attention_f32(q, k, v, cos, sin, B, T, h, hkv, hd)[0]
    return attention(q, k, v, cos, sin, B, T, h, hkv, hd, dropout_p, seed)[0]


@torch.library.custom_op("b200llm::attention_qkv", mutates_args=())
def attention_qkv(qkv: Tensor, cos: Optional[Tensor], sin: Optional[Tensor], B: int, T: int, h: int, hkv: int,
                  hd: int, dropout_p: float = 0.0, seed: int = 0) -> Tuple[Tensor, Tensor, Tensor]:
    """attention() on a packed [B*T, (h + 2 hkv) * hd] projection (q heads | k heads | v heads):
    RoPE of q and k is one launch, the kernels read the column ranges in place. -> (o, lse, qk_rot)."""
    qkv = _flat2d(qkv)
    nq, nk = h * hd, hkv * hd
    if cos is not None:
        qk = K.rope(qkv, cos, sin, B, T, h + hkv, hd)          # [M, (h + hkv) * hd]
    else:
        qk = qkv
    o, lse = K.attention_fwd(qk[:, :nq], qk[:, nq:nq + nk], qkv[:, nq + nk:], B, T, h, hkv, hd,
                             dropout_p=dropout_p, seed=seed, seed_dev=_drop_step(qkv, dropout_p))
    return o, lse, (qk if cos is not None else qkv.new_empty((0,)))


@attention_qkv.register_fake
def _(qkv, cos, sin, B, T, h, hkv, hd, dropout_p=0.0, seed=0):
    o, lse = qkv.new_empty((B * T, h * hd)), qkv.new_empty((B, h, T), dtype=torch.float32)
    return o, lse, (qkv.new_empty((B * T, (h + hkv) * hd)) if cos is not None else qkv.new_empty((0,)))


def _attention_qkv_setup(ctx, inputs, output):
    ctx.set_materialize_grads(False)
    qkv, cos, sin, B, T, h, hkv, hd, dropout_p, seed = inputs
    o, lse, qk = output
    ctx.save_for_backward(qkv, qk, o, lse, cos, sin)
    ctx.dims = (B, T, h, hkv, hd)
    ctx.drop = (dropout_p, seed)


def _attention_qkv_bwd(ctx, do, _dl, _dqk):
    if do is None:
        return (None,) * 10
    qkv, qk, o, lse, cos, sin = ctx.saved_tensors
    B, T, h, hkv, hd = ctx.dims
    nq, nk = h * hd, hkv * hd
    qkv2 = _flat2d(qkv)
    src = qk if cos is not None else qkv2
    dqkv = qkv2.new_empty(qkv2.shape)
    K.attention_bwd(src[:, :nq], src[:, nq:nq + nk], qkv2[:, nq + nk:], o, _flat2d(do), lse, B, T, h, hkv, hd,
                    dq=dqkv[:, :nq], dk=dqkv[:, nq:nq + nk], dv=dqkv[:, nq + nk:], dropout_p=ctx.drop[0],
                    seed=ctx.drop[1], seed_dev=_drop_step(qkv2, ctx.drop[0]))
    if cos is not None:
        K.rope_(dqkv, cos, sin, B, T, h + hkv, hd, inverse=True)      # q and k heads in one launch
    return (dqkv.view(qkv.shape),) + (None,) * 9


attention_qkv.register_autograd(_attention_qkv_bwd, setup_context=_attention_qkv_setup)


# ======================================================================================
# dropout (nn.Dropout sites of Models/GPT2/GPT2.py:24,81,102)
# ======================================================================================
# Every dropout mask is a function of (site seed + per-device step counter, element index). The site seed is a
# host integer drawn when the op is called (new_seed()); under CUDA-graph replay it is frozen into the captured
# kernel arguments, so the kernels also add a counter they read from DEVICE memory when they run.
# engine.TrainStep bumps it inside the captured step (advance_dropout_step), which makes every replay draw
# fresh masks; forward, backward and an activation-checkpoint recompute of one step all see the same value.
_drop_steps: dict = {}


def dropout_step_tensor(device) -> Tensor:
    """The int64 device scalar added to every dropout seed on `device`."""
    device = torch.device(device)
    idx = device.index if device.index is not None else torch.cuda.current_device()
    t = _drop_steps.get(idx)
    if t is None:
        t = torch.zeros(1, dtype=torch.int64, device=torch.device("cuda", idx))
        _drop_steps[idx] = t
    return t


def advance_dropout_step(device) -> None:
    """counter += 1 as a stream-ordered device operation (no host sync; capturable)."""
    dropout_step_tensor(device).add_(1)


def _drop_step(like: Tensor, p: float):
    return dropout_step_tensor(like.device) if p > 0 else None


@torch.library.custom_op("b200llm::dropout", mutates_args=())
def dropout(x: Tensor, residual: Optional[Tensor], p: float, seed: int) -> Tensor:
    """(residual +) keep * x / (1 - p); keep is a counter-based function of (seed + step counter, element index)."""
    if x.dtype != BF16:
        raise RuntimeError("b200llm dropout is a bf16 kernel: the fp32 parity mode runs with drop_rate = 0 (SURVEY.md H5)")
    return K.dropout(x, p, seed, residual, seed_dev=_drop_step(x, p)).view(x.shape)


@dropout.register_fake
def _(x, residual, p, seed):
    return torch.empty_like(x)


def _dropout_setup(ctx, inputs, output):
    _, residual, p, seed = inputs
    ctx.p, ctx.seed, ctx.has_res = p, seed, residual is not None


def _dropout_bwd(ctx, dy):
    return (K.dropout(dy.contiguous(), ctx.p, ctx.seed, seed_dev=_drop_step(dy, ctx.p)).view(dy.shape),
            (dy if ctx.has_res else None), None, None)


dropout.register_autograd(_dropout_bwd, setup_context=_dropout_setup)


def new_seed() -> int:
    """A fresh 62-bit site seed from torch's CPU generator (follows torch.manual_seed and the RNG state that
    activation checkpointing restores for its recompute; no device sync)."""
    return int(torch.randint(0, 2 ** 62, (1,)).item())


# ======================================================================================
# SwiGLU / GELU
# ======================================================================================
@torch.library.custom_op("b200llm::swiglu", mutates_args=())
def swiglu(a: Tensor, b: Tensor) -> Tensor:
    if a.dtype == F32:
        return K32.swiglu_fwd(a, b)
    return K.swiglu_fwd(_flat2d(a), _flat2d(b)).view(a.shape)


@swiglu.register_fake
def _(a, b):
    return torch.empty(a.shape, device=a.device, dtype=a.dtype)


def _swiglu_setup(ctx, inputs, output):
    ctx.save_for_backward(*inputs)


def _swiglu_bwd(ctx, dh):
    a, b = ctx.saved_tensors
    if a.dtype == F32:
        da, db = K32.swiglu_bwd(a, b, dh)
        return da.view(a.shape), db.view(b.shape)
    da, db = K.swiglu_bwd(_flat2d(a), _flat2d(b), dh.contiguous())
    return da.view(a.shape), db.view(b.shape)


swiglu.register_autograd(_swiglu_bwd, setup_context=_swiglu_setup)


@torch.library.custom_op("b200llm::swiglu_packed", mutates_args=())
def swiglu_packed(ab: Tensor) -> Tensor:
    """swiglu() on a packed [.., 2F] = (fc1 x | fc2 x); backward writes one packed [.., 2F] gradient."""
    ab2 = _flat2d(ab)
    F = ab2.shape[1] // 2
    return K.swiglu_fwd(ab2[:, :F], ab2[:, F:]).view(*ab.shape[:-1], F)


@swiglu_packed.register_fake
def _(ab):
    return ab.new_empty((*ab.shape[:-1], ab.shape[-1] // 2))


def _swiglu_packed_setup(ctx, inputs, output):
    ctx.save_for_backward(inputs[0])


def _swiglu_packed_bwd(ctx, dh):
    (ab,) = ctx.saved_tensors
    ab2 = _flat2d(ab)
    F = ab2.shape[1] // 2
    dab = ab2.new_empty(ab2.shape)
    K.swiglu_bwd(ab2[:, :F], ab2[:, F:], dh.contiguous(), da=dab[:, :F], db=dab[:, F:])
    return dab.view(ab.shape)


swiglu_packed.register_autograd(_swiglu_packed_bwd, setup_context=_swiglu_packed_setup)


# --------------------------------------------------------------------------------------
# SwiGLU + down projection as one op (common_components.py:121-124: fc3(silu(fc1 x) * fc2 x)), so that
# backward can form (da | db) in the epilogue of the fc3 input-gradient GEMM: dh never reaches memory and
# the 10-bytes-per-element SwiGLU backward kernel disappears.
# --------------------------------------------------------------------------------------
@torch.library.custom_op("b200llm::swiglu_down", mutates_args=())
def swiglu_down(ab: Tensor, weight: Tensor, A: Optional[Tensor], B: Optional[Tensor], scaling: float,
                residual: Optional[Tensor], h_in: Optional[Tensor]) -> List[Tensor]:
    """ab [.., 2F] = (fc1 x | fc2 x) -> [y = fc3(h) (+ residual), t = scaling h A, h = silu(a) * b].
    h_in: h already formed by the up projections' epilogue (lora_fc12_swiglu), else it is computed here."""
    ab2 = _flat2d(ab)
    F = ab2.shape[1] // 2
    h = _flat2d(h_in) if h_in is not None else K.swiglu_fwd(ab2[:, :F], ab2[:, F:])
    res2 = _flat2d(residual) if residual is not None else None
    if A is not None:
        t = K.gemm_ex(h, A, b1_mn=True, alpha=scaling)
        y = K.gemm_ex(h, weight, a2=t, b2=_pitch_ok(B), b2_mn=True, residual=res2)
    else:
        t = h.new_empty((0,))
        y = K.gemm_ex(h, weight, residual=res2)
    # h is handed back only to be saved for backward: a fresh tensor unless it came in (outputs must not alias inputs)
    return [y.view(*ab.shape[:-1], weight.shape[0]), t, h if h_in is None else h.new_empty((0,))]


@swiglu_down.register_fake
def _(ab, weight, A, B, scaling, residual, h_in):
    rows = ab.numel() // ab.shape[-1]
    F = ab.shape[-1] // 2
    t = ab.new_empty((rows, A.shape[1])) if A is not None else ab.new_empty((0,))
    return [ab.new_empty((*ab.shape[:-1], weight.shape[0])), t, ab.new_empty((rows, F) if h_in is None else (0,))]


def _swiglu_down_setup(ctx, inputs, output):
    ctx.set_materialize_grads(False)
    ab, weight, A, B, scaling, residual, h_in = inputs
    ctx.save_for_backward(ab, weight, A, B, output[1], output[2] if h_in is None else h_in)
    ctx.scaling = scaling
    ctx.has_res = residual is not None
    ctx.p_w, ctx.p_A, ctx.p_B = weight, A, B


def _swiglu_down_bwd(ctx, grads):
    dy = grads[0]
    if dy is None:
        return (None,) * 7
    ab, weight, A, B, t, h = ctx.saved_tensors
    weight, A, B = ctx.p_w, ctx.p_A, ctx.p_B                     # the Parameters themselves (see _main_grad)
    h = _flat2d(h)
    need = ctx.needs_input_grad
    ab2, dy2 = _flat2d(ab), _flat2d(dy)
    M, F = h.shape
    u = None
    if A is not None and (need[0] or A.requires_grad):
        u = K.gemm_ex(dy2, _pitch_ok(B), alpha=ctx.scaling)            # s * dY B^T   [M, r]
    dab = None
    if need[0]:
        fused = (M >= 256 and F % 256 == 0 and dy2.shape[1] % 8 == 0
                 and os.environ.get("B200_SWIGLU_EPILOGUE", "1") != "0")
        if fused:
            dab = K.gemm_swiglu_bwd(dy2, weight, u, A if u is not None else None, ab2)
        else:
            dh = K.gemm_ex(dy2, weight, b1_mn=True, a2=u, b2=A) if u is not None else K.gemm_ex(dy2, weight, b1_mn=True)
            dab = ab2.new_empty(ab2.shape)
            K.swiglu_bwd(ab2[:, :F], ab2[:, F:], dh, da=dab[:, :F], db=dab[:, F:])
        dab = dab.view(ab.shape)
    dW = dA = dB = None
    if A is not None and A.requires_grad:
        dA = _weight_grad(h, u, A)
    if B is not None and B.requires_grad:
        dB = _weight_grad(dy2, t, B, trans_out=True)
    if weight.requires_grad:
        dW = _weight_grad(dy2, h, weight)
    return dab, dW, dA, dB, None, (dy if (ctx.has_res and need[5]) else None), None


swiglu_down.register_autograd(_swiglu_down_bwd, setup_context=_swiglu_down_setup)


def swiglu_down_module(ab, fc3, residual, h=None):
    """swiglu_down for a bias-free fc3 module (plain or LoRA-wrapped); None when it does not apply."""
    if ab.dtype != BF16 or os.environ.get("B200_PACKED", "1") == "0":
        return None
    from .lora import LinearWithLoRA
    if isinstance(fc3, LinearWithLoRA):
        if fc3.linear.bias is not None:
            return None
        return swiglu_down(ab, fc3.linear.weight, fc3.lora.A, fc3.lora.B, fc3.lora.scaling, residual, h)[0]
    if getattr(fc3, "bias", None) is not None:
        return None
    return swiglu_down(ab, fc3.weight, None, None, 0.0, residual, h)[0]


@torch.library.custom_op("b200llm::gelu", mutates_args=())
def gelu(x: Tensor) -> Tensor:
    return K.gelu_fwd(x.contiguous())


@gelu.register_fake
def _(x):
    return torch.empty_like(x)


def _gelu_setup(ctx, inputs, output):
    ctx.save_for_backward(inputs[0])


def _gelu_bwd(ctx, dy):
    (x,) = ctx.saved_tensors
    return K.gelu_bwd(x.contiguous(), dy.contiguous())


gelu.register_autograd(_gelu_bwd, setup_context=_gelu_setup)


# ======================================================================================
# embedding  (Llama3.py:197; GPT2.py:110-112)
# ======================================================================================
@torch.library.custom_op("b200llm::embedding", mutates_args=())
def embedding(idx: Tensor, tok_w: Tensor, pos_w: Optional[Tensor]) -> Tensor:
    return K.embedding_fwd(idx, tok_w, pos_w)


@embedding.register_fake
def _(idx, tok_w, pos_w):
    return tok_w.new_empty((*idx.shape, tok_w.shape[1]))


def _embedding_setup(ctx, inputs, output):
    idx, tok_w, pos_w = inputs
    ctx.save_for_backward(idx)
    ctx.weights = (tok_w, pos_w)
    ctx.vocab = tok_w.shape[0]
    ctx.ctx_len = None if pos_w is None else pos_w.shape[0]
    ctx.dtype = tok_w.dtype


def _embedding_bwd(ctx, dout):
    (idx,) = ctx.saved_tensors
    if not (ctx.needs_input_grad[1] or ctx.needs_input_grad[2]):
        return None, None, None
    tok_w, pos_w = ctx.weights
    both = ctx.needs_input_grad[1] and (pos_w is None or ctx.needs_input_grad[2])
    mg = _main_grad(tok_w, *([pos_w] if pos_w is not None else [])) if both else None
    if mg:      # scatter-add straight into the flat fp32 gradient buffer (no [V, d] zero-fill, cast, copy, add)
        K.embedding_bwd(idx, dout.contiguous(), ctx.vocab, ctx.ctx_len, dtok_out=mg[0],
                        dpos_out=mg[1] if pos_w is not None else None)
        _notify_ready([tok_w] + ([pos_w] if pos_w is not None else []))
        return None, None, None
    dtok, dpos = K.embedding_bwd(idx, dout.contiguous(), ctx.vocab, ctx.ctx_len)
    dtok = K.cast(dtok, ctx.dtype) if ctx.dtype == BF16 else dtok
    if dpos is not None:
        dpos = K.cast(dpos, ctx.dtype) if ctx.dtype == BF16 else dpos
    return None, dtok, dpos


embedding.register_autograd(_embedding_bwd, setup_context=_embedding_setup)


# ======================================================================================
# cross-entropy and the vocab projection fused with it  (train.py:90-91)
# ======================================================================================
@torch.library.custom_op("b200llm::cross_entropy", mutates_args=())
def cross_entropy(logits: Tensor, targets: Tensor) -> Tuple[Tensor, Tensor]:
    """F.cross_entropy(logits.flatten(0,1), targets.flatten()) -> (loss fp32 [], dlogits)."""
    V = logits.shape[-1]
    M = logits.numel() // V
    ld = (V + 7) // 8 * 8
    buf = torch.empty((M, ld), device=logits.device, dtype=logits.dtype)
    buf[:, :V].copy_(logits.reshape(M, V))
    loss = (K32 if logits.dtype == F32 else K).cross_entropy_(buf[:, :V], targets)
    return loss.reshape(()), buf


@cross_entropy.register_fake
def _(logits, targets):
    V = logits.shape[-1]
    return logits.new_empty((), dtype=torch.float32), logits.new_empty((logits.numel() // V, (V + 7) // 8 * 8))


def _ce_setup(ctx, inputs, output):
    ctx.set_materialize_grads(False)
    ctx.save_for_backward(output[1])
    ctx.shape = inputs[0].shape


def _ce_bwd(ctx, dloss, _db):
    if dloss is None:
        return None, None
    (buf,) = ctx.saved_tensors
    V = ctx.shape[-1]
    g = buf[:, :V].reshape(ctx.shape)
    return g * dloss.to(g.dtype), None


cross_entropy.register_autograd(_ce_bwd, setup_context=_ce_setup)


@torch.library.custom_op("b200llm::linear_cross_entropy", mutates_args=())
def linear_cross_entropy(x: Tensor, weight: Tensor, A: Optional[Tensor], B: Optional[Tensor], scaling: float,
                         targets: Tensor, capacity: int, write_grad: bool = True) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """loss = CE(x W^T + scaling (x A) B, targets) -> (loss fp32 [], dlogits, t, x_rows, row_index).

    The logits live only in the returned work buffer, which the cross-entropy kernel overwrites in
    place with dloss/dlogits for the backward GEMMs. capacity > 0: only the rows whose target is not
    ignore_index are projected (they are the only ones that reach the loss or any gradient);
    capacity must be >= their count (otherwise the loss comes back NaN). write_grad=False (evaluation under
    no_grad) skips the dlogits pass of the cross-entropy kernel."""
    x2 = _flat2d(x)
    V = weight.shape[0]
    if x.dtype == F32:          # parity mode: every row, fp32 logits in the work buffer
        M, ld = x2.shape[0], (V + 7) // 8 * 8
        buf = torch.empty((M, ld), device=x.device, dtype=F32)
        K32.gemm(x2, weight, out=buf[:, :V])
        t = x2.new_empty((0,))
        if A is not None:
            t = K32.gemm(x2, A.t(), alpha=scaling)
            K32.gemm(t, B.t(), out=buf[:, :V], beta=1.0)
        loss = K32.cross_entropy_(buf[:, :V], targets, write_grad=write_grad)
        return loss.reshape(()), buf, t, x2.new_empty((0,)), x2.new_empty((0,), dtype=torch.int32)
    if capacity > 0 and capacity < x2.shape[0]:
        rows, tg, idx = K.compact_gather(x2, targets, capacity)
    else:
        rows, tg, idx = x2, targets, x2.new_empty((0,), dtype=torch.int32)
    M = rows.shape[0]
    ld = (V + 7) // 8 * 8
    buf = torch.empty((M, ld), device=x.device, dtype=BF16)
    if A is not None:
        t = K.gemm_ex(rows, A, b1_mn=True, alpha=scaling)
        K.gemm_ex(rows, weight, a2=t, b2=_pitch_ok(B), b2_mn=True, out=buf[:, :V])
    else:
        t = x2.new_empty((0,))
        K.gemm_ex(rows, weight, out=buf[:, :V])
    loss = K.cross_entropy_(buf[:, :V], tg, write_grad=write_grad)
    return loss.reshape(()), buf, t, (rows if idx.numel() else x2.new_empty((0,))), idx


@linear_cross_entropy.register_fake
def _(x, weight, A, B, scaling, targets, capacity, write_grad=True):
    M, V = x.numel() // x.shape[-1], weight.shape[0]
    compact = 0 < capacity < M
    R = capacity if compact else M
    t = x.new_empty((R, A.shape[1])) if A is not None else x.new_empty((0,))
    rows = x.new_empty((R, x.shape[-1])) if compact else x.new_empty((0,))
    idx = x.new_empty((R if compact else 0,), dtype=torch.int32)
    return x.new_empty((), dtype=torch.float32), x.new_empty((R, (V + 7) // 8 * 8)), t, rows, idx


def _lce_setup(ctx, inputs, output):
    ctx.set_materialize_grads(False)
    x, weight, A, B, scaling, targets, capacity, write_grad = inputs
    _, buf, t, rows, idx = output
    ctx.save_for_backward(x, weight, A, B, t, buf, rows, idx)
    ctx.scaling = scaling
    ctx.write_grad = write_grad
    ctx.p_w, ctx.p_A, ctx.p_B = weight, A, B


def _lce_bwd(ctx, dloss, _dbuf, _dt, _dr, _di):
    """The forward wrote dlogits for d(loss) = 1; the incoming gradient (a device scalar: 1 for loss.backward() as
    train.py:112 does, 1/k under gradient accumulation, a loss scale) is applied to that buffer in place by a kernel
    that returns immediately when it reads exactly 1 — so one backward per forward."""
    if dloss is None:
        return (None,) * 8
    if not ctx.write_grad:
        raise RuntimeError("linear_cross_entropy was run with write_grad=False (evaluation): it has no backward")
    x, weight, A, B, t, buf, rows, idx = ctx.saved_tensors
    if buf.dtype == F32:
        V = weight.shape[0]
        dl = K32.scale_by_device_scalar(buf, dloss.detach().to(F32))[:, :V]
        dx, dW, _, dA, dB = K32.linear_bwd(dl, _flat2d(x), weight, None, A, B, t if A is not None else None, ctx.scaling,
                                           ctx.needs_input_grad[0])
        return (None if dx is None else dx.view(x.shape)), dW, dA, dB, None, None, None, None
    K.scale_dev_(buf, dloss.detach().reshape(1).to(torch.float32))
    wref, aref, bref = ctx.p_w, ctx.p_A, ctx.p_B                 # the Parameters themselves (see _main_grad)
    need = ctx.needs_input_grad
    V = weight.shape[0]
    dl = buf[:, :V]
    compact = idx.numel() > 0
    x2 = rows if compact else _flat2d(x)
    dx = dW = dA = dB = None
    u = None
    if A is not None and (need[0] or need[2]):
        u = K.gemm_ex(dl, _pitch_ok(B), alpha=ctx.scaling)
    if need[0]:
        if u is not None:
            dx2 = K.gemm_ex(dl, weight, b1_mn=True, a2=u, b2=A)
        else:
            dx2 = K.gemm_ex(dl, weight, b1_mn=True)
        if compact:
            dx2 = K.scatter_rows(dx2, idx, x.numel() // x.shape[-1])
        dx = dx2.view(x.shape)
    if A is not None and need[2]:
        dA = _weight_grad(x2, u, aref)
    if B is not None and need[3]:
        dB = _weight_grad(dl, t, bref, trans_out=True)
    if need[1]:
        dW = _weight_grad(dl, x2, wref)
    return dx, dW, dA, dB, None, None, None, None


linear_cross_entropy.register_autograd(_lce_bwd, setup_context=_lce_setup)
