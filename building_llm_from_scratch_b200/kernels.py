"""Tensor-level wrappers over the C-ABI (one Python function per extern "C" entry point).

These take CUDA torch tensors, pass raw pointers / sizes / the current stream to libb200llm.so
and return tensors allocated by the PyTorch caching allocator (the library itself never
allocates). CPU tensors are rejected: there is no fallback path.
"""
from __future__ import annotations

import math
import os

import torch

from . import _lib

BF16 = torch.bfloat16
GEMM_BIAS, GEMM_GELU, GEMM_RESIDUAL = 1, 2, 4


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t) -> int:
    return 0 if t is None else t.data_ptr()


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("b200llm ops need CUDA tensors on an sm_100a device (no CPU fallback)")


def _rowmajor2d(t):
    """(tensor, ld) for a 2-D view whose last dim is contiguous."""
    assert t.dim() == 2 and t.stride(1) == 1, f"need a row-major 2-D view, got {t.shape} {t.stride()}"
    return t, t.stride(0)


class GemmProfile:
    """Opt-in per-launch timing of the GEMM (CUDA events on the launching stream) for bench.py's
    roofline figure: algorithmic FLOPs = 2*M*N*(K1+K2) per launch."""

    def __init__(self):
        self.records = []

    def reset(self):
        self.records = []

    def begin(self):
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def end(self, start, flops, shape):
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        self.records.append((start, e, flops, shape))

    def totals(self):
        torch.cuda.synchronize()
        flops = sum(r[2] for r in self.records)
        ms = sum(r[0].elapsed_time(r[1]) for r in self.records)
        return flops, ms, len(self.records)

    def by_shape(self):
        torch.cuda.synchronize()
        out = {}
        for s, e, f, shape in self.records:
            d = out.setdefault(shape, [0, 0.0, 0.0])
            d[0] += 1
            d[1] += s.elapsed_time(e)
            d[2] += f
        return out


_gemm_profile = None


def set_gemm_profile(p):
    global _gemm_profile
    _gemm_profile = p


_gemm_ws: dict = {}


def _ensure_gemm_workspace(device):
    """Register (once per device) the zero-filled scratch that lets the large-GEMM kernel split the
    tiles of a partly filled last wave along K (b200_gemm_set_workspace)."""
    if device.index in _gemm_ws:
        return
    if torch.cuda.is_current_stream_capturing():
        return                      # first use inside a capture: run without until an eager call registers it
    lib = _lib.load()
    with torch.cuda.device(device):
        ws = torch.zeros(int(lib.b200_gemm_workspace_bytes()), dtype=torch.uint8, device=device)
        _lib.check(lib.b200_gemm_set_workspace(_p(ws), ws.numel()), "gemm_set_workspace")
    _gemm_ws[device.index] = ws


def gemm_ex(a1, b1, *, a1_mn=False, b1_mn=False, a2=None, b2=None, a2_mn=False, b2_mn=False, bias=None,
            residual=None, gelu=False, out=None, out_dtype=BF16, accumulate=False, trans_out=False,
            alpha=1.0, splitk=0, M=None, N=None):
    """C[M,N] = alpha*(op(a1)·op(b1)^T + op(a2)·op(b2)^T) (+bias)(gelu)(+residual).

    K-major operands are [rows, K]; MN-major (x_mn=True) are stored [K, rows]."""
    _need_cuda(a1, b1, a2, b2, bias, residual, out)
    lib = _lib.load()
    _ensure_gemm_workspace(a1.device)
    a1, lda1 = _rowmajor2d(a1)
    b1, ldb1 = _rowmajor2d(b1)
    m1, k1 = (a1.shape[1], a1.shape[0]) if a1_mn else a1.shape
    n1, kb = (b1.shape[1], b1.shape[0]) if b1_mn else b1.shape
    assert k1 == kb, (a1.shape, b1.shape)
    M = m1 if M is None else M
    N = n1 if N is None else N
    K2 = 0
    lda2 = ldb2 = 0
    if a2 is not None:
        a2, lda2 = _rowmajor2d(a2)
        b2, ldb2 = _rowmajor2d(b2)
        K2 = a2.shape[0] if a2_mn else a2.shape[1]
        assert K2 == (b2.shape[0] if b2_mn else b2.shape[1])
    if out is None:
        shape = (N, M) if trans_out else (M, N)
        out = (torch.zeros if accumulate else torch.empty)(shape, device=a1.device, dtype=out_dtype)
    if (STREAMK and N <= 64 and a2 is None and bias is None and residual is None and not gelu and splitk == 0
            and M == m1 and N == n1 and out.dim() == 2):
        # skinny product (LoRA rank-r projection or weight gradient): HBM-bound stream, stream-K scheduled
        gemm_grouped([dict(a=a1, b=b1, out=out, a_mn=a1_mn, b_mn=b1_mn, accumulate=accumulate,
                           trans_out=trans_out, alpha=alpha)])
        return out
    out2, ldc = _rowmajor2d(out)
    flags = (GEMM_BIAS if bias is not None else 0) | (GEMM_GELU if gelu else 0) | (
        GEMM_RESIDUAL if residual is not None else 0)
    ldr = 0
    if residual is not None:
        residual, ldr = _rowmajor2d(residual)
    prof = _gemm_profile
    start = prof.begin() if prof is not None else None
    rc = lib.b200_gemm_ex(_p(a1), lda1, int(a1_mn), _p(b1), ldb1, int(b1_mn), k1, _p(a2), lda2, int(a2_mn),
                          _p(b2), ldb2, int(b2_mn), K2, _p(bias), _p(residual), ldr, _p(out2), ldc, M, N, flags,
                          int(out.dtype == torch.float32), int(accumulate), int(trans_out), float(alpha),
                          int(splitk), _stream())
    _lib.check(rc, "gemm_ex")
    if prof is not None:
        prof.end(start, 2.0 * M * N * (k1 + K2), (M, N, k1, K2, int(a1_mn), int(b1_mn), int(accumulate)))
    return out


def gemm_lora_packed(x, w_cat, t_cat, b_cat, r, seg_starts, out=None):
    """y[M, N] = x w_cat^T + t_cat[:, seg*r:(seg+1)*r] b_cat[:, n]  (one launch for several LoRA linears of x)."""
    _need_cuda(x, w_cat, t_cat, b_cat, out)
    lib = _lib.load()
    _ensure_gemm_workspace(x.device)
    x, ldx = _rowmajor2d(x)
    w_cat, ldw = _rowmajor2d(w_cat)
    t_cat, ldt = _rowmajor2d(t_cat)
    b_cat, ldb = _rowmajor2d(b_cat)
    M, Kd = x.shape
    N = w_cat.shape[0]
    assert w_cat.shape[1] == Kd and b_cat.shape == (r, N) and t_cat.shape == (M, r * len(seg_starts))
    if out is None:
        out = torch.empty((M, N), device=x.device, dtype=BF16)
    out2, ldc = _rowmajor2d(out)
    segs = (_lib.C.c_int64 * len(seg_starts))(*seg_starts)
    prof = _gemm_profile
    start = prof.begin() if prof is not None else None
    _lib.check(lib.b200_gemm_lora_packed_bf16(_p(x), ldx, _p(w_cat), ldw, Kd, _p(t_cat), ldt, _p(b_cat), ldb, r,
                                              len(seg_starts), segs, _p(out2), ldc, M, N, _stream()), "gemm_lora_packed")
    if prof is not None:
        prof.end(start, 2.0 * M * N * (Kd + r), (M, N, Kd, r, 0, 0, 0))
    return out


def gemm_qkv_rope(x, w_cat, t_cat, b_cat, r, seg_starts, cos, sin, T, hd, rope_cols, out=None):
    """The packed q|k|v projection (gemm_lora_packed; t_cat / b_cat None for plain linears) with RoPE applied to the
    first rope_cols output columns in the epilogue; row m is position m % T."""
    _need_cuda(x, w_cat, t_cat, b_cat, cos, sin, out)
    lib = _lib.load()
    _ensure_gemm_workspace(x.device)
    x, ldx = _rowmajor2d(x)
    w_cat, ldw = _rowmajor2d(w_cat)
    M, Kd = x.shape
    N = w_cat.shape[0]
    ldt = ldb = 0
    if t_cat is not None:
        t_cat, ldt = _rowmajor2d(t_cat)
        b_cat, ldb = _rowmajor2d(b_cat)
        assert b_cat.shape == (r, N) and t_cat.shape == (M, r * len(seg_starts))
    else:
        r = 0
    assert cos.dtype == BF16 and sin.dtype == BF16 and cos.shape[1] == hd and cos.shape[0] >= T and cos.is_contiguous() and sin.is_contiguous()
    if out is None:
        out = torch.empty((M, N), device=x.device, dtype=BF16)
    out2, ldc = _rowmajor2d(out)
    segs = (_lib.C.c_int64 * len(seg_starts))(*seg_starts)
    prof = _gemm_profile
    start = prof.begin() if prof is not None else None
    _lib.check(lib.b200_gemm_qkv_rope_bf16(_p(x), ldx, _p(w_cat), ldw, Kd, _p(t_cat), ldt, _p(b_cat), ldb, r, len(seg_starts), segs,
                                           _p(cos), _p(sin), T, hd, rope_cols, _p(out2), ldc, M, N, _stream()), "gemm_qkv_rope")
    if prof is not None:
        prof.end(start, 2.0 * M * N * (Kd + r), (M, N, Kd, r, 0, 0, 0))
    return out


def gemm_swiglu_bwd(dy, w, u, A, ab, out=None):
    """(da | db) [M, 2F] = SwiGLU'(a, b) * (dy w + u A^T) with (a | b) = ab [M, 2F]; w [K, F] (fc3.weight), u [M, r]
    and A [F, r] or both None."""
    _need_cuda(dy, w, u, A, ab, out)
    lib = _lib.load()
    dy, lddy = _rowmajor2d(dy)
    w, ldw = _rowmajor2d(w)
    ab, ldab = _rowmajor2d(ab)
    M, Kd = dy.shape
    F = w.shape[1]
    assert w.shape[0] == Kd and ab.shape == (M, 2 * F)
    r = ldu = lda = 0
    if u is not None:
        u, ldu = _rowmajor2d(u)
        A, lda = _rowmajor2d(A)
        r = u.shape[1]
        assert A.shape == (F, r)
    if out is None:
        out = torch.empty((M, 2 * F), device=dy.device, dtype=BF16)
    out2, ldo = _rowmajor2d(out)
    prof = _gemm_profile
    start = prof.begin() if prof is not None else None
    _lib.check(lib.b200_gemm_swiglu_bwd_bf16(_p(dy), lddy, _p(w), ldw, Kd, _p(u), ldu, _p(A), lda, r, _p(ab), ldab,
                                             _p(out2), ldo, M, F, _stream()), "gemm_swiglu_bwd")
    if prof is not None:
        prof.end(start, 2.0 * M * F * (Kd + r), ("swiglu_bwd_epilogue", M, F, Kd, r))
    return out


def gemm_swiglu_fwd(x, w_il, t2, b_il, ab=None, h=None):
    """(a | b) [M, 2F] = x [W1; W2]^T (+ LoRA), h [M, F] = silu(a) * b in one launch. w_il [2F, K]: rows of W1 / W2
    interleaved in blocks of 128; t2 [M, 2r], b_il [2r, 2F] (see b200_gemm_swiglu_fwd_bf16) or both None."""
    _need_cuda(x, w_il, t2, b_il, ab, h)
    lib = _lib.load()
    _ensure_gemm_workspace(x.device)
    x, ldx = _rowmajor2d(x)
    w_il, ldw = _rowmajor2d(w_il)
    M, Kd = x.shape
    F = w_il.shape[0] // 2
    r2 = ldt = ldb = 0
    if t2 is not None:
        t2, ldt = _rowmajor2d(t2)
        b_il, ldb = _rowmajor2d(b_il)
        r2 = t2.shape[1]
        assert b_il.shape == (r2, 2 * F)
    ab = torch.empty((M, 2 * F), device=x.device, dtype=BF16) if ab is None else ab
    h = torch.empty((M, F), device=x.device, dtype=BF16) if h is None else h
    prof = _gemm_profile
    start = prof.begin() if prof is not None else None
    _lib.check(lib.b200_gemm_swiglu_fwd_bf16(_p(x), ldx, _p(w_il), ldw, Kd, _p(t2), ldt, _p(b_il), ldb, r2, _p(ab),
                                             ab.stride(0), _p(h), h.stride(0), M, F, _stream()), "gemm_swiglu_fwd")
    if prof is not None:
        prof.end(start, 2.0 * M * 2 * F * (Kd + r2 // 2), (M, 2 * F, Kd, r2, 0, 0, 0))
    return ab, h


def fused_epilogue_ok(M, N):
    """Shapes the GELU-backward / dropout-residual epilogues (staged TMA epilogue of the CTA-pair kernel) accept."""
    return M >= 256 and N >= 256 and N % 64 == 0


def gemm_gelu_fwd(x, w, bias, t=None, Bl=None):
    """-> (h, z): z = x w^T + bias (+ t Bl), h = gelu(z) in one launch (z kept for the GELU's backward)."""
    _need_cuda(x, w, bias, t, Bl)
    lib = _lib.load()
    _ensure_gemm_workspace(x.device)
    x, ldx = _rowmajor2d(x)
    w, ldw = _rowmajor2d(w)
    M, Kd = x.shape
    N = w.shape[0]
    r = ldt = ldb = 0
    if t is not None:
        t, ldt = _rowmajor2d(t)
        Bl, ldb = _rowmajor2d(Bl)
        r = t.shape[1]
    z = torch.empty((M, N), device=x.device, dtype=BF16)
    h = torch.empty((M, N), device=x.device, dtype=BF16)
    prof = _gemm_profile
    start = prof.begin() if prof is not None else None
    _lib.check(lib.b200_gemm_gelu_fwd_bf16(_p(x), ldx, _p(w), ldw, Kd, _p(t), ldt, _p(Bl), ldb, r, _p(bias), _p(z), N, _p(h), N,
                                           M, N, _stream()), "gemm_gelu_fwd")
    if prof is not None:
        prof.end(start, 2.0 * M * N * (Kd + r), (M, N, Kd, r, 0, 0, 0))
    return h, z


def gemm_gelu_bwd(dy, w, z, u=None, A=None):
    """dz = gelu'(z) * (dy w + u A^T); w [K, N] (the second MLP linear's weight), z [M, N]."""
    _need_cuda(dy, w, z, u, A)
    lib = _lib.load()
    _ensure_gemm_workspace(dy.device)
    dy, lddy = _rowmajor2d(dy)
    w, ldw = _rowmajor2d(w)
    z, ldz = _rowmajor2d(z)
    M, Kd = dy.shape
    N = w.shape[1]
    assert w.shape[0] == Kd and z.shape == (M, N)
    r = ldu = lda = 0
    if u is not None:
        u, ldu = _rowmajor2d(u)
        A, lda = _rowmajor2d(A)
        r = u.shape[1]
    dz = torch.empty((M, N), device=dy.device, dtype=BF16)
    prof = _gemm_profile
    start = prof.begin() if prof is not None else None
    _lib.check(lib.b200_gemm_gelu_bwd_bf16(_p(dy), lddy, _p(w), ldw, Kd, _p(u), ldu, _p(A), lda, r, _p(z), ldz, _p(dz), N, M, N,
                                           _stream()), "gemm_gelu_bwd")
    if prof is not None:
        prof.end(start, 2.0 * M * N * (Kd + r), (M, N, Kd, r, 0, 1, 0))
    return dz


def gemm_dropout_residual(x, w, bias, residual, p, seed, seed_dev=None, t=None, Bl=None):
    """y = residual + dropout(x w^T + bias (+ t Bl)) in one launch; mask = dropout()'s for the same (seed, seed_dev)."""
    _need_cuda(x, w, bias, residual, t, Bl, seed_dev)
    lib = _lib.load()
    _ensure_gemm_workspace(x.device)
    x, ldx = _rowmajor2d(x)
    w, ldw = _rowmajor2d(w)
    residual, ldr = _rowmajor2d(residual)
    M, Kd = x.shape
    N = w.shape[0]
    r = ldt = ldb = 0
    if t is not None:
        t, ldt = _rowmajor2d(t)
        Bl, ldb = _rowmajor2d(Bl)
        r = t.shape[1]
    y = torch.empty((M, N), device=x.device, dtype=BF16)
    prof = _gemm_profile
    start = prof.begin() if prof is not None else None
    _lib.check(lib.b200_gemm_dropout_residual_bf16(_p(x), ldx, _p(w), ldw, Kd, _p(t), ldt, _p(Bl), ldb, r, _p(bias), _p(residual),
                                                   ldr, _p(y), N, M, N, float(p), int(seed), _p(seed_dev), _stream()),
               "gemm_dropout_residual")
    if prof is not None:
        prof.end(start, 2.0 * M * N * (Kd + r), (M, N, Kd, r, 0, 0, 0))
    return y


def gemm_grouped(problems):
    """problems: list of dicts(a, b, out, a_mn, b_mn, accumulate, trans_out, alpha) with N <= 64.
    One persistent launch for all of them (chunks of GEMM_GROUP_MAX)."""
    lib = _lib.load()
    prof = _gemm_profile
    for i0 in range(0, len(problems), _lib.GEMM_GROUP_MAX):
        chunk = problems[i0:i0 + _lib.GEMM_GROUP_MAX]
        arr = (_lib.GemmProblem * len(chunk))()
        flops = 0.0
        for q, pr in zip(arr, chunk):
            a, b, out = pr["a"], pr["b"], pr["out"]
            _need_cuda(a, b, out)
            a_mn, b_mn = bool(pr.get("a_mn")), bool(pr.get("b_mn"))
            M, K = (a.shape[1], a.shape[0]) if a_mn else a.shape
            N = b.shape[1] if b_mn else b.shape[0]
            q.A, q.lda, q.a_mn = a.data_ptr(), a.stride(0), int(a_mn)
            q.B, q.ldb, q.b_mn = b.data_ptr(), b.stride(0), int(b_mn)
            q.C, q.ldc = out.data_ptr(), out.stride(0)
            q.M, q.N, q.K = M, N, K
            q.out_f32 = int(out.dtype == torch.float32)
            q.accumulate = int(bool(pr.get("accumulate")))
            q.trans_out = int(bool(pr.get("trans_out")))
            q.alpha = float(pr.get("alpha", 1.0))
            flops += 2.0 * M * N * K
        start = prof.begin() if prof is not None else None
        if STREAMK:
            need = int(lib.b200_gemm_grouped_workspace_bytes(len(chunk), arr))
            ws = _streamk_workspace(chunk[0]["a"].device, need)
            _lib.check(lib.b200_gemm_grouped_streamk_bf16(len(chunk), arr, _p(ws), ws.numel(), _stream()),
                       "gemm_grouped_streamk")
        else:
            _lib.check(lib.b200_gemm_grouped_bf16(len(chunk), arr, _stream()), "gemm_grouped")
        if prof is not None:
            prof.end(start, flops, ("grouped", len(chunk)))


# B200_GEMM_STREAMK=0 selects the tile-scheduled grouped kernel (A/B measurements)
STREAMK = os.environ.get("B200_GEMM_STREAMK", "1") != "0"
_streamk_ws: dict = {}
_streamk_keep: list = []


def _streamk_workspace(device, nbytes: int):
    """Zero-filled scratch of the stream-K grouped GEMM, one per device (all b200llm work of a process
    is issued on one stream at a time: forward, autograd and graph capture follow each other). The
    kernel leaves it zero-filled, so it is allocated and cleared once — outside graph capture, TrainStep
    warms up eagerly — and only re-made when a larger group shows up; earlier buffers stay alive
    because captured graphs may hold their address."""
    ws = _streamk_ws.get(device.index)
    if ws is None or ws.numel() < nbytes:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("stream-K GEMM workspace must be sized by an eager call before graph capture")
        ws = torch.zeros(max(nbytes, 48 << 20), dtype=torch.uint8, device=device)
        _streamk_keep.append(ws)
        _streamk_ws[device.index] = ws
    return ws


def rmsnorm_fwd(x, w, eps=1e-5, res=None, bf16_steps=True):
    """-> y (and the residual sum when res is given). x [M, d] bf16 (or fp32), w fp32 [d]."""
    _need_cuda(x, w, res)
    lib = _lib.load()
    x2 = x.reshape(-1, x.shape[-1])
    M, d = x2.shape
    y = torch.empty_like(x2)
    if x.dtype == torch.float32:
        assert res is None
        _lib.check(lib.b200_rmsnorm_fwd_f32(_p(x2), _p(w), _p(y), M, d, eps, _stream()), "rmsnorm_fwd_f32")
        return y.view_as(x)
    s = None
    if res is not None:
        res = res.reshape(-1, d)
        s = torch.empty_like(x2)
    _lib.check(lib.b200_rmsnorm_fwd_bf16(_p(x2), _p(res), _p(w), _p(s), _p(y), M, d, eps, int(bf16_steps), _stream()),
               "rmsnorm_fwd")
    return (y.view_as(x), s.view_as(x)) if res is not None else y.view_as(x)


def rmsnorm_bwd(x, w, dy, eps=1e-5, dres=None, need_dw=False, dw_out=None):
    """dw_out: an fp32 [d] accumulator (e.g. the weight's main_grad) the kernel adds into instead of a fresh tensor."""
    _need_cuda(x, w, dy, dres, dw_out)
    lib = _lib.load()
    x2 = x.reshape(-1, x.shape[-1])
    M, d = x2.shape
    dx = torch.empty_like(x2)
    dw = dw_out if dw_out is not None else (torch.zeros(d, device=x.device, dtype=torch.float32) if need_dw else None)
    dy2 = dy.reshape(-1, d)
    if x.dtype == torch.float32:
        _lib.check(lib.b200_rmsnorm_bwd_f32(_p(x2), _p(w), _p(dy2), _p(dx), _p(dw), M, d, eps, _stream()), "rmsnorm_bwd_f32")
    else:
        _lib.check(lib.b200_rmsnorm_bwd_bf16(_p(x2), _p(w), _p(dy2), _p(dres), _p(dx), _p(dw), M, d, eps, _stream()),
                   "rmsnorm_bwd")
    return dx.view_as(x), dw


def layernorm_fwd(x, w, b, eps=1e-5):
    if x.dtype == torch.float32:
        from . import f32
        return f32.layernorm_fwd(x, w, b, eps)
    _need_cuda(x, w, b)
    lib = _lib.load()
    x2 = x.reshape(-1, x.shape[-1])
    M, d = x2.shape
    y = torch.empty_like(x2)
    mean = torch.empty(M, device=x.device, dtype=torch.float32)
    rstd = torch.empty(M, device=x.device, dtype=torch.float32)
    _lib.check(lib.b200_layernorm_fwd_bf16(_p(x2), _p(w), _p(b), _p(y), _p(mean), _p(rstd), M, d, eps, _stream()),
               "layernorm_fwd")
    return y.view_as(x), mean, rstd


def layernorm_bwd(x, w, dy, mean, rstd, need_dw=True, dw_out=None, db_out=None, dres=None):
    """dw_out / db_out: fp32 [d] accumulators (main_grad of weight / bias) the kernel adds into; dres: gradient of the
    residual stream, added to dx as it is written."""
    if x.dtype == torch.float32:
        from . import f32
        dx, dw, db = f32.layernorm_bwd(x, w, dy, mean, rstd, need_dw, dw_out, db_out)
        return (dx if dres is None else dx + dres), dw, db
    _need_cuda(x, w, dy, dw_out, db_out, dres)
    lib = _lib.load()
    x2 = x.reshape(-1, x.shape[-1])
    M, d = x2.shape
    dx = torch.empty_like(x2)
    if dw_out is not None:
        dw, db = dw_out, db_out
    else:
        dw = torch.zeros(d, device=x.device, dtype=torch.float32) if need_dw else None
        db = torch.zeros(d, device=x.device, dtype=torch.float32) if need_dw else None
    _lib.check(lib.b200_layernorm_bwd_res_bf16(_p(x2), _p(w), _p(dy.reshape(-1, d)),
                                               _p(dres.reshape(-1, d)) if dres is not None else None, _p(mean), _p(rstd),
                                               _p(dx), _p(dw), _p(db), M, d, _stream()), "layernorm_bwd")
    return dx.view_as(x), dw, db


def swiglu_fwd(a, b, bf16_steps=True):
    """a, b: row-major 2-D views [M, F] (possibly the halves of one buffer) -> h [M, F]."""
    _need_cuda(a, b)
    a, lda = _rowmajor2d(a)
    b, ldb = _rowmajor2d(b)
    M, F = a.shape
    h = torch.empty((M, F), device=a.device, dtype=BF16)
    _lib.check(_lib.load().b200_swiglu_fwd_bf16(_p(a), lda, _p(b), ldb, _p(h), M, F, int(bf16_steps), _stream()),
               "swiglu_fwd")
    return h


def swiglu_bwd(a, b, dh, da=None, db=None):
    _need_cuda(a, b, dh)
    a, lda = _rowmajor2d(a)
    b, ldb = _rowmajor2d(b)
    M, F = a.shape
    da = torch.empty((M, F), device=a.device, dtype=BF16) if da is None else da
    db = torch.empty((M, F), device=a.device, dtype=BF16) if db is None else db
    dh = dh.reshape(M, F)
    assert dh.is_contiguous()
    _lib.check(_lib.load().b200_swiglu_bwd_bf16(_p(a), lda, _p(b), ldb, _p(dh), _p(da), da.stride(0), _p(db),
                                                db.stride(0), M, F, _stream()), "swiglu_bwd")
    return da, db


def gelu_fwd(x):
    if x.dtype == torch.float32:
        from . import f32
        return f32.gelu_fwd(x)
    _need_cuda(x)
    y = torch.empty_like(x)
    _lib.check(_lib.load().b200_gelu_fwd_bf16(_p(x), _p(y), x.numel(), _stream()), "gelu_fwd")
    return y


def gelu_bwd(x, dy):
    if x.dtype == torch.float32:
        from . import f32
        return f32.gelu_bwd(x, dy)
    _need_cuda(x, dy)
    dx = torch.empty_like(x)
    _lib.check(_lib.load().b200_gelu_bwd_bf16(_p(x), _p(dy), _p(dx), x.numel(), _stream()), "gelu_bwd")
    return dx


def rope(x2d, cos, sin, B, T, nheads, hd, inverse=False, bf16_steps=True, out=None):
    """RoPE on the first nheads*hd columns of the [B*T, ld] view x2d; out=None -> new tensor,
    out=x2d -> in place."""
    _need_cuda(x2d, cos, sin)
    x2d, ld = _rowmajor2d(x2d)
    if out is None:
        out = torch.empty((x2d.shape[0], nheads * hd), device=x2d.device, dtype=BF16)
    out, ldo = _rowmajor2d(out)
    assert cos.dtype == BF16 and sin.dtype == BF16 and cos.shape[1] == hd and cos.shape[0] >= T
    _lib.check(_lib.load().b200_rope_bf16(_p(x2d), ld, _p(out), ldo, _p(cos), _p(sin), B, T, nheads, hd, int(inverse),
                                          int(bf16_steps), _stream()), "rope")
    return out


def rope_(x2d, cos, sin, B, T, nheads, hd, inverse=False, bf16_steps=True):
    return rope(x2d, cos, sin, B, T, nheads, hd, inverse, bf16_steps, out=x2d)


def embedding_fwd(idx, tok_w, pos_w=None):
    if tok_w.dtype == torch.float32:
        from . import f32
        return f32.embedding_fwd(idx, tok_w, pos_w)
    _need_cuda(idx, tok_w, pos_w)
    B, T = idx.shape
    d = tok_w.shape[1]
    out = torch.empty((B, T, d), device=tok_w.device, dtype=BF16)
    _lib.check(_lib.load().b200_embedding_fwd_bf16(_p(idx.contiguous()), _p(tok_w), _p(pos_w), _p(out), B * T, T, d,
                                                   tok_w.shape[0], _stream()), "embedding_fwd")
    return out


def embedding_bwd(idx, dout, vocab, ctx=None, dtok_out=None, dpos_out=None):
    """dtok_out / dpos_out: fp32 accumulators ([vocab, d] / [ctx, d], e.g. main_grad) to scatter-add into."""
    if dout.dtype == torch.float32:
        from . import f32
        return f32.embedding_bwd(idx, dout, vocab, ctx, dtok_out, dpos_out)
    _need_cuda(idx, dout, dtok_out, dpos_out)
    B, T = idx.shape
    d = dout.shape[-1]
    dtok = dtok_out if dtok_out is not None else torch.zeros((vocab, d), device=dout.device, dtype=torch.float32)
    dpos = dpos_out if dpos_out is not None else (torch.zeros((ctx, d), device=dout.device, dtype=torch.float32) if ctx else None)
    _lib.check(_lib.load().b200_embedding_bwd_f32acc(_p(idx.contiguous()), _p(dout.contiguous()), _p(dtok), _p(dpos),
                                                     B * T, T, d, vocab, _stream()), "embedding_bwd")
    return dtok, dpos


def cross_entropy_(logits2d, targets, write_grad=True):
    """Loss (fp32 scalar tensor); logits2d [M, V(+pad)] is overwritten with dloss/dlogits."""
    _need_cuda(logits2d, targets)
    lib = _lib.load()
    logits2d, ld = _rowmajor2d(logits2d)
    M = logits2d.shape[0]
    V = logits2d.shape[1]
    tg = targets.reshape(-1).contiguous()
    assert tg.numel() == M and tg.dtype == torch.int64
    nbytes = lib.b200_cross_entropy_workspace_bytes(M)
    ws = torch.empty(nbytes, device=logits2d.device, dtype=torch.uint8)
    loss = torch.empty(1, device=logits2d.device, dtype=torch.float32)
    _lib.check(lib.b200_cross_entropy_bf16(_p(logits2d), ld, _p(tg), _p(loss), M, V, int(write_grad), _p(ws), nbytes,
                                           _stream()), "cross_entropy")
    return loss


def compact_gather(x2d, targets, capacity):
    """Rows of x2d whose target != -100, packed into [capacity, d] (zero padded), their targets
    (-100 padded) and the row index list (int32, -1 padded)."""
    _need_cuda(x2d, targets)
    lib = _lib.load()
    x2d, ldx = _rowmajor2d(x2d)
    M, d = x2d.shape
    tg = targets.reshape(-1).contiguous()
    idx = torch.empty(capacity, device=x2d.device, dtype=torch.int32)
    meta = torch.empty(2, device=x2d.device, dtype=torch.int32)
    xc = torch.empty((capacity, d), device=x2d.device, dtype=BF16)
    tc = torch.empty(capacity, device=x2d.device, dtype=torch.int64)
    _lib.check(lib.b200_compact_rows(_p(tg), M, capacity, _p(idx), _p(meta), _stream()), "compact_rows")
    _lib.check(lib.b200_gather_rows_bf16(_p(x2d), ldx, _p(tg), _p(idx), _p(meta), _p(xc), _p(tc), capacity, d, _stream()),
               "gather_rows")
    return xc, tc, idx


def scatter_rows(src, idx, M):
    """[M, d] with src[j] at row idx[j] and zeros elsewhere."""
    _need_cuda(src, idx)
    C_, d = src.shape
    dst = torch.zeros((M, d), device=src.device, dtype=BF16)
    _lib.check(_lib.load().b200_scatter_rows_bf16(_p(src), _p(idx), _p(dst), dst.stride(0), C_, d, _stream()), "scatter_rows")
    return dst


def grad_norm(g_flat, max_norm=1.0, out=None, ws=None):
    """-> fp32 tensor [norm, clip_coef]; g_flat bf16 or fp32."""
    _need_cuda(g_flat)
    lib = _lib.load()
    n = g_flat.numel()
    nbytes = lib.b200_grad_norm_workspace_bytes(n)
    if ws is None:
        ws = torch.empty(nbytes, device=g_flat.device, dtype=torch.uint8)
    if out is None:
        out = torch.empty(2, device=g_flat.device, dtype=torch.float32)
    fn = lib.b200_grad_norm_f32 if g_flat.dtype == torch.float32 else lib.b200_grad_norm_bf16
    _lib.check(fn(_p(g_flat), n, float(max_norm), _p(out), _p(ws), nbytes, _stream()), "grad_norm")
    return out


def adam_hyper(lr, beta1, beta2, eps, weight_decay, step):
    """7 floats (pinned host tensor) for adamw_'s dev_hyper argument."""
    host = torch.empty(7, dtype=torch.float32, pin_memory=torch.cuda.is_available())
    _lib.check(_lib.load().b200_adam_hyper(host.data_ptr(), lr, beta1, beta2, eps, weight_decay, int(step)), "adam_hyper")
    return host


def adamw_(p, g, m, v, step, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.1, grad_scale=None,
           dev_hyper=None):
    _need_cuda(p, g, m, v, grad_scale, dev_hyper)
    lib = _lib.load()
    if p.dtype == torch.float32:
        rc = lib.b200_adamw_f32(_p(p), _p(g), _p(m), _p(v), p.numel(), lr, beta1, beta2, eps, weight_decay, int(step),
                                _p(dev_hyper), _p(grad_scale), _stream())
    else:
        rc = lib.b200_adamw_bf16(_p(p), _p(g), int(g.dtype == torch.float32), _p(m), _p(v), p.numel(), lr, beta1, beta2,
                                 eps, weight_decay, int(step), _p(dev_hyper), _p(grad_scale), _stream())
    _lib.check(rc, "adamw")


def adamw_master_(p32, g32, m32, v32, p_bf16, step, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.1,
                  grad_scale=None, dev_hyper=None):
    """AdamW on fp32 master weights (fp32 gradients and moments) that also rewrites the bf16 compute copy."""
    _need_cuda(p32, g32, m32, v32, p_bf16, grad_scale, dev_hyper)
    assert p32.dtype == g32.dtype == m32.dtype == v32.dtype == torch.float32 and p_bf16.dtype == BF16
    assert p_bf16.numel() == p32.numel()
    _lib.check(_lib.load().b200_adamw_master_f32(_p(p32), _p(g32), _p(m32), _p(v32), _p(p_bf16), p32.numel(), lr, beta1,
                                                 beta2, eps, weight_decay, int(step), _p(dev_hyper), _p(grad_scale),
                                                 _stream()), "adamw_master")


def cast(x, dtype):
    _need_cuda(x)
    lib = _lib.load()
    out = torch.empty(x.shape, device=x.device, dtype=dtype)
    xc = x.contiguous()
    if x.dtype == torch.float32 and dtype == BF16:
        _lib.check(lib.b200_cast_f32_to_bf16(_p(xc), _p(out), x.numel(), _stream()), "cast")
    elif x.dtype == BF16 and dtype == torch.float32:
        _lib.check(lib.b200_cast_bf16_to_f32(_p(xc), _p(out), x.numel(), _stream()), "cast")
    else:
        raise RuntimeError(f"unsupported cast {x.dtype} -> {dtype}")
    return out


def transpose(x, alpha=1.0, out=None):
    _need_cuda(x)
    x, ldi = _rowmajor2d(x)
    rows, cols = x.shape
    if out is None:
        out = torch.empty((cols, rows), device=x.device, dtype=BF16)
    _lib.check(_lib.load().b200_transpose_bf16(_p(x), ldi, _p(out), out.stride(0), rows, cols, float(alpha), _stream()),
               "transpose")
    return out


def add(a, b):
    _need_cuda(a, b)
    out = torch.empty_like(a)
    _lib.check(_lib.load().b200_add_bf16(_p(a), _p(b), _p(out), a.numel(), _stream()), "add")
    return out


def colsum(x2d, like=None, out=None):
    """Column sums of a [M, N] bf16 matrix (fp32 accumulate) -> [N] in like.dtype (default fp32); out: an fp32 [N]
    accumulator to add into (returned as is)."""
    _need_cuda(x2d, out)
    x2d, ld = _rowmajor2d(x2d)
    M, N = x2d.shape
    if out is not None:
        _lib.check(_lib.load().b200_colsum_bf16(_p(x2d), ld, _p(out), M, N, _stream()), "colsum")
        return out
    out = torch.zeros(N, device=x2d.device, dtype=torch.float32)
    _lib.check(_lib.load().b200_colsum_bf16(_p(x2d), ld, _p(out), M, N, _stream()), "colsum")
    if like is not None and like.dtype == BF16:
        return cast(out, BF16)
    return out


def dropout(x, p, seed, residual=None, seed_dev=None):
    """(residual +) keep * x / (1 - p) with the counter-based mask of (seed + seed_dev[0], element index);
    seed_dev: optional int64 device tensor read by the kernel when it runs (per-step offset under graph replay)."""
    _need_cuda(x, residual, seed_dev)
    xc = x.contiguous()
    y = torch.empty_like(xc)
    rc = None if residual is None else residual.contiguous()
    _lib.check(_lib.load().b200_dropout_bf16(_p(xc), _p(rc), _p(y), xc.numel(), float(p), int(seed), _p(seed_dev),
                                             _stream()), "dropout")
    return y


def scale_dev_(x, scale):
    """x *= scale[0] in place (scale: fp32 device scalar; the kernel returns at once when it is exactly 1)."""
    _need_cuda(x, scale)
    assert x.is_contiguous() and scale.dtype == torch.float32 and x.dtype == BF16
    _lib.check(_lib.load().b200_scale_bf16_dev(_p(x), x.numel(), _p(scale), _stream()), "scale_dev")
    return x


def dropout_mask(n, p, seed, device="cuda"):
    out = torch.empty(n, device=device, dtype=torch.uint8)
    _lib.check(_lib.load().b200_dropout_mask_u8(_p(out), n, float(p), int(seed), _stream()), "dropout_mask")
    return out


def attention_dropout_mask(B, H, T, p, seed, device="cuda"):
    out = torch.empty((B, H, T, T), device=device, dtype=torch.uint8)
    _lib.check(_lib.load().b200_attention_dropout_mask_u8(_p(out), B, H, T, float(p), int(seed), _stream()),
               "attention_dropout_mask")
    return out


def attention_fwd(q, k, v, B, T, h, hkv, hd, out=None, dropout_p=0.0, seed=0, seed_dev=None):
    """q/k/v: 2-D row-major views [B*T, *] whose first h*hd (hkv*hd) columns are the heads."""
    _need_cuda(q, k, v)
    q, ldq = _rowmajor2d(q)
    k, ldk = _rowmajor2d(k)
    v, ldv = _rowmajor2d(v)
    o = torch.empty((B * T, h * hd), device=q.device, dtype=BF16) if out is None else out
    lse = torch.empty((B, h, T), device=q.device, dtype=torch.float32)
    _lib.check(_lib.load().b200_attention_fwd_bf16(_p(q), ldq, _p(k), ldk, _p(v), ldv, _p(o), o.stride(0), _p(lse), B, T,
                                                   h, hkv, hd, 1.0 / math.sqrt(hd), float(dropout_p), int(seed), _p(seed_dev),
                                                   _stream()), "attention_fwd")
    return o, lse


def attention_bwd(q, k, v, o, do, lse, B, T, h, hkv, hd, dq=None, dk=None, dv=None, dropout_p=0.0, seed=0,
                  seed_dev=None, rope=None):
    """rope = (cos, sin): q / k were rotated before the attention kernels (projection epilogue); dq / dk then come out
    multiplied by the transposed rotation (gradients of the un-rotated projections)."""
    _need_cuda(q, k, v, o, do, lse)
    q, ldq = _rowmajor2d(q)
    k, ldk = _rowmajor2d(k)
    v, ldv = _rowmajor2d(v)
    o, ldo = _rowmajor2d(o)
    do, lddo = _rowmajor2d(do)
    dev = q.device
    dq = torch.empty((B * T, h * hd), device=dev, dtype=BF16) if dq is None else dq
    dk = torch.empty((B * T, hkv * hd), device=dev, dtype=BF16) if dk is None else dk
    dv = torch.empty((B * T, hkv * hd), device=dev, dtype=BF16) if dv is None else dv
    lib = _lib.load()
    nbytes = lib.b200_attention_bwd_workspace_bytes(B, T, h, hd)
    ws = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    if rope is not None:
        assert dropout_p == 0.0
        cos, sin = rope
        assert cos.dtype == BF16 and cos.is_contiguous() and sin.is_contiguous() and cos.shape[1] == hd and cos.shape[0] >= T
        _lib.check(lib.b200_attention_bwd_rope_bf16(_p(q), ldq, _p(k), ldk, _p(v), ldv, _p(o), ldo, _p(do), lddo, _p(lse),
                                                    _p(ws), nbytes, _p(dq), dq.stride(0), _p(dk), dk.stride(0), _p(dv),
                                                    dv.stride(0), B, T, h, hkv, hd, 1.0 / math.sqrt(hd), _p(cos), _p(sin),
                                                    _stream()), "attention_bwd_rope")
        return dq, dk, dv
    _lib.check(lib.b200_attention_bwd_bf16(_p(q), ldq, _p(k), ldk, _p(v), ldv, _p(o), ldo, _p(do), lddo, _p(lse),
                                           _p(ws), nbytes, _p(dq), dq.stride(0), _p(dk), dk.stride(0), _p(dv),
                                           dv.stride(0), B, T, h, hkv, hd, 1.0 / math.sqrt(hd), float(dropout_p),
                                           int(seed), _p(seed_dev), _stream()), "attention_bwd")
    return dq, dk, dv


def attention_decode(q, k_cache, v_cache, T, h, hkv, hd, out=None):
    """One query token per sequence against the first T positions of a KV cache.
    q [B, h*hd]; k_cache, v_cache [B, t_max, hkv*hd] (keys already rotated) -> o [B, h*hd]."""
    _need_cuda(q, k_cache, v_cache)
    q, ldq = _rowmajor2d(q)
    B, t_max, width = k_cache.shape
    assert v_cache.shape == k_cache.shape and k_cache.is_contiguous() and v_cache.is_contiguous() and width == hkv * hd
    o = torch.empty((B, h * hd), device=q.device, dtype=BF16) if out is None else out
    _lib.check(_lib.load().b200_attention_decode_bf16(_p(q), ldq, _p(k_cache), _p(v_cache), width, t_max, _p(o),
                                                      o.stride(0), B, T, h, hkv, hd, 1.0 / math.sqrt(hd), _stream()),
               "attention_decode")
    return o
