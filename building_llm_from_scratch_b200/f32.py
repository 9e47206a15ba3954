"""fp32 parity mode: the operators of ops.py evaluated by the plain-fp32 kernels of csrc/fp32.cu.

ops.py routes here whenever an op is handed float32 activations (the reference's GPT-2 is fp32 as written,
Models/GPT2/GPT2.py:20-23, and its Llama sub-modules accept dtype=float32), so the fp32 fixtures recorded from the
reference replay on the GPU at the north star's 1e-4. Everything is FFMA arithmetic through the C-ABI — no tensor
cores, no torch math — and nothing here is tuned: it is a correctness mode.
"""
from __future__ import annotations

import math

import torch

from . import _lib

F32 = torch.float32


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t) -> int:
    return 0 if t is None else t.data_ptr()


def _need(*ts):
    for t in ts:
        if t is not None and (not t.is_cuda or t.dtype != F32):
            raise RuntimeError("b200llm fp32 ops need float32 CUDA tensors")


def gemm(a, b, bias=None, out=None, alpha=1.0, beta=0.0):
    """out[M, N] = alpha * a[M, K] @ b[N, K]^T (+ bias) (+ beta * out); a, b, out: 2-D views with any strides."""
    _need(a, b, bias, out)
    M, Kd = a.shape
    N = b.shape[0]
    assert b.shape[1] == Kd
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=F32)
        beta = 0.0
    _lib.check(_lib.load().b200_gemm_f32(_p(a), 0, a.stride(0), a.stride(1), _p(b), 0, b.stride(0), b.stride(1), _p(bias),
                                         _p(out), 0, out.stride(0), out.stride(1), 1, M, N, Kd, float(alpha), float(beta),
                                         _stream()), "gemm_f32")
    return out


def bgemm(a, b, out, alpha=1.0, beta=0.0):
    """Batched: a [Bt, M, K], b [Bt, N, K], out [Bt, M, N] — strided views (a batch stride of 0 broadcasts)."""
    _need(a, b, out)
    Bt, M, Kd = a.shape
    N = b.shape[1]
    _lib.check(_lib.load().b200_gemm_f32(_p(a), a.stride(0), a.stride(1), a.stride(2), _p(b), b.stride(0), b.stride(1),
                                         b.stride(2), 0, _p(out), out.stride(0), out.stride(1), out.stride(2), Bt, M, N, Kd,
                                         float(alpha), float(beta), _stream()), "gemm_f32")
    return out


def _ew(a, b, op):
    a = a.contiguous()
    b = None if b is None else b.contiguous()
    _need(a, b)
    y = torch.empty_like(a)
    _lib.check(_lib.load().b200_elementwise_f32(_p(a), _p(b), _p(y), a.numel(), op, _stream()), "elementwise_f32")
    return y


def gelu_fwd(x):
    return _ew(x, None, 0)


def gelu_bwd(x, dy):
    return _ew(x, dy, 1)


def add(a, b):
    return _ew(a, b, 2)


def swiglu_fwd(a, b):
    return _ew(a, b, 3)


def scale_by_device_scalar(a, scalar):
    """a * scalar[0] (scalar: a float32 device tensor with one element)."""
    return _ew(a, scalar.reshape(1), 4)


def swiglu_bwd(a, b, dh):
    a, b, dh = a.contiguous(), b.contiguous(), dh.contiguous()
    _need(a, b, dh)
    da, db = torch.empty_like(a), torch.empty_like(a)
    _lib.check(_lib.load().b200_swiglu_bwd_f32(_p(a), _p(b), _p(dh), _p(da), _p(db), a.numel(), _stream()), "swiglu_bwd_f32")
    return da, db


def layernorm_fwd(x, w, b, eps):
    x2 = x.reshape(-1, x.shape[-1]).contiguous()
    _need(x2, w, b)
    M, d = x2.shape
    y = torch.empty_like(x2)
    mean, rstd = torch.empty(M, device=x.device, dtype=F32), torch.empty(M, device=x.device, dtype=F32)
    _lib.check(_lib.load().b200_layernorm_fwd_f32(_p(x2), _p(w), _p(b), _p(y), _p(mean), _p(rstd), M, d, eps, _stream()),
               "layernorm_fwd_f32")
    return y.view_as(x), mean, rstd


def layernorm_bwd(x, w, dy, mean, rstd, need_dw=True, dw_out=None, db_out=None):
    x2, dy2 = x.reshape(-1, x.shape[-1]).contiguous(), dy.reshape(-1, x.shape[-1]).contiguous()
    _need(x2, w, dy2, dw_out, db_out)
    M, d = x2.shape
    dx = torch.empty_like(x2)
    if dw_out is not None:
        dw, db = dw_out, db_out
    else:
        dw = torch.zeros(d, device=x.device, dtype=F32) if need_dw else None
        db = torch.zeros(d, device=x.device, dtype=F32) if need_dw else None
    _lib.check(_lib.load().b200_layernorm_bwd_f32(_p(x2), _p(w), _p(dy2), _p(mean), _p(rstd), _p(dx), _p(dw), _p(db), M, d,
                                                  _stream()), "layernorm_bwd_f32")
    return dx.view_as(x), dw, db


def rope(x2d, cos, sin, T, nheads, hd, inverse=False):
    """Rotate-half RoPE of the first nheads * hd columns of the contiguous [B*T, nheads*hd] tensor x2d."""
    x2d = x2d.contiguous()
    cos, sin = cos.to(F32).contiguous(), sin.to(F32).contiguous()
    _need(x2d, cos, sin)
    assert x2d.shape[1] == nheads * hd and cos.shape[1] == hd and cos.shape[0] >= T
    y = torch.empty_like(x2d)
    _lib.check(_lib.load().b200_rope_f32(_p(x2d), _p(y), _p(cos), _p(sin), x2d.shape[0], T, nheads, hd, int(inverse), _stream()),
               "rope_f32")
    return y


def embedding_fwd(idx, tok_w, pos_w=None):
    _need(tok_w, pos_w)
    B, T = idx.shape
    d = tok_w.shape[1]
    out = torch.empty((B, T, d), device=tok_w.device, dtype=F32)
    _lib.check(_lib.load().b200_embedding_fwd_f32(_p(idx.contiguous()), _p(tok_w), _p(pos_w), _p(out), B * T, T, d, _stream()),
               "embedding_fwd_f32")
    return out


def embedding_bwd(idx, dout, vocab, ctx=None, dtok_out=None, dpos_out=None):
    dout = dout.contiguous()
    _need(dout, dtok_out, dpos_out)
    B, T = idx.shape
    d = dout.shape[-1]
    dtok = dtok_out if dtok_out is not None else torch.zeros((vocab, d), device=dout.device, dtype=F32)
    dpos = dpos_out if dpos_out is not None else (torch.zeros((ctx, d), device=dout.device, dtype=F32) if ctx else None)
    _lib.check(_lib.load().b200_embedding_bwd_f32(_p(idx.contiguous()), _p(dout), _p(dtok), _p(dpos), B * T, T, d, _stream()),
               "embedding_bwd_f32")
    return dtok, dpos


def cross_entropy_(logits2d, targets, write_grad=True):
    """Loss (fp32 scalar tensor [1]); the contiguous-row fp32 logits2d [M, V] is overwritten with dloss/dlogits."""
    _need(logits2d)
    assert logits2d.stride(1) == 1
    lib = _lib.load()
    M, V = logits2d.shape
    tg = targets.reshape(-1).contiguous()
    nbytes = lib.b200_cross_entropy_f32_workspace_bytes(M)
    ws = torch.empty(nbytes, device=logits2d.device, dtype=torch.uint8)
    loss = torch.empty(1, device=logits2d.device, dtype=F32)
    _lib.check(lib.b200_cross_entropy_f32(_p(logits2d), logits2d.stride(0), _p(tg), _p(loss), M, V, int(write_grad), _p(ws),
                                          nbytes, _stream()), "cross_entropy_f32")
    return loss


# ---------------------------------------------------------------------------------------------------------------
# causal attention with materialised scores (parity-mode sizes): q [B*T, h*hd], k / v [B*T, hkv*hd] contiguous
# ---------------------------------------------------------------------------------------------------------------
def _heads(t2d, b, T, g0, G, hd, nheads):
    """[G, T, hd] strided view of heads g0 .. g0+G-1 of batch element b (G = 1 with stride 0 broadcasts a kv head)."""
    base = t2d[b * T:(b + 1) * T]
    return base.as_strided((G, T, hd), (hd if G > 1 else 0, nheads * hd, 1), base.storage_offset() + g0 * hd)


def attention_fwd(q, k, v, B, T, h, hkv, hd):
    """-> (o [B*T, h*hd], P [B, h, T, T]); P (the softmax weights) is what backward needs."""
    _need(q, k, v)
    q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
    G = h // hkv
    P = torch.empty((B, h, T, T), device=q.device, dtype=F32)
    o = torch.empty((B * T, h * hd), device=q.device, dtype=F32)
    for b in range(B):
        for g in range(hkv):
            Pg = P[b, g * G:(g + 1) * G]
            kg = _heads(k, b, T, g, 1, hd, hkv).expand(G, T, hd)
            bgemm(_heads(q, b, T, g * G, G, hd, h), kg, Pg)                       # S = Q K^T
    _lib.check(_lib.load().b200_softmax_causal_f32(_p(P), B * h * T, T, 1.0 / math.sqrt(hd), _stream()), "softmax_causal_f32")
    for b in range(B):
        for g in range(hkv):
            vg = _heads(v, b, T, g, 1, hd, hkv).expand(G, T, hd).transpose(1, 2)  # [G, hd, T]
            bgemm(P[b, g * G:(g + 1) * G], vg, _heads(o, b, T, g * G, G, hd, h))   # O = P V
    return o, P


def attention_bwd(q, k, v, P, do, B, T, h, hkv, hd):
    _need(q, k, v, P, do)
    q, k, v, do = q.contiguous(), k.contiguous(), v.contiguous(), do.contiguous()
    G = h // hkv
    dq, dk, dv = torch.empty_like(q), torch.zeros_like(k), torch.zeros_like(v)
    dS = torch.empty_like(P)
    for b in range(B):
        for g in range(hkv):
            vg = _heads(v, b, T, g, 1, hd, hkv).expand(G, T, hd)
            bgemm(_heads(do, b, T, g * G, G, hd, h), vg, dS[b, g * G:(g + 1) * G])          # dP = dO V^T
    _lib.check(_lib.load().b200_softmax_bwd_f32(_p(P), _p(dS), B * h * T, T, 1.0 / math.sqrt(hd), _stream()), "softmax_bwd_f32")
    for b in range(B):
        for g in range(hkv):
            dSg, Pg = dS[b, g * G:(g + 1) * G], P[b, g * G:(g + 1) * G]
            kg = _heads(k, b, T, g, 1, hd, hkv).expand(G, T, hd).transpose(1, 2)            # [G, hd, T]
            bgemm(dSg, kg, _heads(dq, b, T, g * G, G, hd, h))                               # dQ = dS K
            for j in range(G):      # the G query heads of a group accumulate into one kv head
                head = g * G + j
                qh = _heads(q, b, T, head, 1, hd, h)[0]
                doh = _heads(do, b, T, head, 1, hd, h)[0]
                gemm(dSg[j].t(), qh.t(), out=_heads(dk, b, T, g, 1, hd, hkv)[0], beta=1.0)  # dK += dS^T Q
                gemm(Pg[j].t(), doh.t(), out=_heads(dv, b, T, g, 1, hd, hkv)[0], beta=1.0)  # dV += P^T dO
    return dq, dk, dv


# ---------------------------------------------------------------------------------------------------------------
# linear (+ LoRA): forward and the gradient pieces, shared by ops.py's fp32 branches
# ---------------------------------------------------------------------------------------------------------------
def linear_fwd(x2, weight, bias, A, B, scaling, residual=None):
    """-> (y, t): y = x W^T + bias + s (x A) B (+ residual), t = s x A (or None)."""
    y = gemm(x2, weight, bias)
    t = None
    if A is not None:
        t = gemm(x2, A.t(), alpha=scaling)
        gemm(t, B.t(), out=y, beta=1.0)
    if residual is not None:
        y = add(y, residual)
    return y, t


def weight_grad(lhs_t, rhs_t, param):
    """param-shaped gradient = lhs_t @ rhs_t^T with lhs_t [rows(param), tokens], rhs_t [cols(param), tokens] (views);
    accumulated into param.main_grad when it exists (returns None then), else returned."""
    mg = getattr(param, "main_grad", None)
    if mg is not None:
        gemm(lhs_t, rhs_t, out=mg, beta=1.0)
        return None
    return gemm(lhs_t, rhs_t)


def bias_grad(dy2, bias):
    ones = torch.ones((1, dy2.shape[0]), device=dy2.device, dtype=F32)
    mg = getattr(bias, "main_grad", None)
    if mg is not None:
        gemm(dy2.t(), ones, out=mg.view(-1, 1), beta=1.0)
        return None
    return gemm(dy2.t(), ones).view(-1)


def linear_bwd(dy2, x2, weight, bias, A, B, t, scaling, need_x=True):
    """-> (dx, dW, dbias, dA, dB) for y = x W^T + bias + s (x A) B."""
    dx = dW = dbias = dA = dB = None
    u = gemm(dy2, B, alpha=scaling) if A is not None else None         # s dY B^T  [M, r]
    if need_x:
        dx = gemm(dy2, weight.t())
        if u is not None:
            gemm(u, A, out=dx, beta=1.0)
    if A is not None and A.requires_grad:
        dA = weight_grad(x2.t(), u.t(), A)
    if B is not None and B.requires_grad:
        dB = weight_grad(t.t(), dy2.t(), B)
    if weight.requires_grad:
        dW = weight_grad(dy2.t(), x2.t(), weight)
    if bias is not None and bias.requires_grad:
        dbias = bias_grad(dy2, bias)
    return dx, dW, dbias, dA, dB
