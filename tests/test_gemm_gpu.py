"""GPU: the tcgen05 GEMM through the C-ABI against the oracle's linear / LoRA-linear
(oracle/reference_ops.py) evaluated in fp32 on the same bf16 inputs. bf16 tolerance 2e-2
relative (north star); in practice the fp32-accumulating kernel is ~3e-3 from the fp32 result."""
import pytest
import torch

from conftest import rel_err

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]

BF16_TOL = 2e-2


@pytest.fixture(scope="module")
def K():
    from building_llm_from_scratch_b200 import kernels
    return kernels


def rnd(*shape, scale=1.0, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(*shape, generator=g, device="cuda") * scale).to(torch.bfloat16)


@pytest.mark.parametrize("M,N,K_", [(128, 256, 64), (128, 64, 64), (128, 128, 128), (256, 512, 256),
                                   (8192, 2048, 2048), (300, 200, 136), (1000, 50257 // 16, 192), (64, 3072, 2048)])
def test_gemm_nt(K, M, N, K_):
    from oracle import reference_ops as R
    a, w = rnd(M, K_, seed=1), rnd(N, K_, scale=0.05, seed=2)
    c = K.gemm_ex(a, w)
    ref = R.linear(a.float(), w.float())
    assert c.shape == (M, N) and c.dtype == torch.bfloat16
    assert rel_err(c.float(), ref) < BF16_TOL


def test_gemm_nt_exact_small_integers(K):
    # integer-valued inputs: every product and partial sum is exact in fp32 -> bit-exact result
    g = torch.Generator(device="cuda").manual_seed(5)
    a = torch.randint(-3, 4, (256, 320), generator=g, device="cuda").to(torch.bfloat16)
    w = torch.randint(-3, 4, (384, 320), generator=g, device="cuda").to(torch.bfloat16)
    c = K.gemm_ex(a, w, out_dtype=torch.float32)
    assert torch.equal(c, a.float() @ w.float().t())


def test_gemm_lora_dual_k(K):
    """LinearWithLoRA forward (lora.py:46): x W^T + s (x A) B in one TMEM accumulation."""
    from oracle import reference_ops as R
    M, d_in, d_out, r, s = 512, 256, 384, 8, 2.0
    x, W = rnd(M, d_in, seed=1), rnd(d_out, d_in, scale=0.05, seed=2)
    A, B = rnd(d_in, r, scale=0.3, seed=3), rnd(r, d_out, scale=0.05, seed=4)
    sd = {"l.linear.weight": W.float(), "l.lora.A": A.float(), "l.lora.B": B.float()}
    ref = R.lora_linear(x.float(), sd, "l", s)
    At = K.transpose(A)                       # [r, in]  (K-major "B operand" of t = x A)
    t = K.gemm_ex(x, At, alpha=s)             # s * (x A)   [M, r]
    Bt = K.transpose(B)                       # [out, r]
    y = K.gemm_ex(x, W, a2=t, b2=Bt)
    assert rel_err(y.float(), ref) < BF16_TOL
    # and with r = 64 plus residual + bias + gelu epilogues
    r = 64
    A, B = rnd(d_in, r, scale=0.1, seed=5), rnd(r, d_out, scale=0.05, seed=6)
    bias, res = rnd(d_out, seed=7), rnd(M, d_out, seed=8)
    t = K.gemm_ex(x, K.transpose(A), alpha=s)
    y = K.gemm_ex(x, W, a2=t, b2=K.transpose(B), bias=bias, residual=res, gelu=True)
    lin = x.float() @ W.float().t() + t.float() @ B.float() + bias.float()
    ref = R.gelu_erf(lin) + res.float()
    assert rel_err(y.float(), ref) < BF16_TOL


@pytest.mark.parametrize("M,N,K_", [(256, 256, 128), (1024, 2048, 512), (200, 264, 72)])
def test_gemm_mn_major_operands(K, M, N, K_):
    a, b = rnd(M, K_, seed=1), rnd(N, K_, scale=0.1, seed=2)
    ref = a.float() @ b.float().t()
    at, bt = a.t().contiguous(), b.t().contiguous()   # stored [K, M], [K, N]
    assert rel_err(K.gemm_ex(a, bt, b1_mn=True, M=M, N=N).float(), ref) < BF16_TOL      # NN (dX = dY W)
    assert rel_err(K.gemm_ex(at, bt, a1_mn=True, b1_mn=True).float(), ref) < BF16_TOL   # TN (dW = dY^T X)
    assert rel_err(K.gemm_ex(at, b, a1_mn=True).float(), ref) < BF16_TOL


def test_gemm_splitk_accumulate_and_transposed_store(K):
    """LoRA weight gradients: dA = x^T u [in, r], dB^T = dY^T t stored as dB [r, out]; fp32
    atomics across split-K, accumulating on top of existing contents."""
    Mtok, d_in, d_out, r = 4096, 512, 768, 64
    x, u = rnd(Mtok, d_in, seed=1), rnd(Mtok, r, scale=0.1, seed=2)
    dY, t = rnd(Mtok, d_out, scale=0.1, seed=3), rnd(Mtok, r, seed=4)
    dA = torch.ones(d_in, r, device="cuda", dtype=torch.float32)
    K.gemm_ex(x, u, a1_mn=True, b1_mn=True, out=dA, accumulate=True)
    assert rel_err(dA, 1 + x.float().t() @ u.float()) < 1e-3
    dB = torch.zeros(r, d_out, device="cuda", dtype=torch.float32)
    K.gemm_ex(dY, t, a1_mn=True, b1_mn=True, out=dB, accumulate=True, trans_out=True, alpha=0.5)
    assert rel_err(dB, 0.5 * t.float().t() @ dY.float()) < 1e-3


def test_gemm_rejects_cpu_and_bad_args(K):
    with pytest.raises(RuntimeError):
        K.gemm_ex(torch.zeros(8, 8, dtype=torch.bfloat16), torch.zeros(8, 8, dtype=torch.bfloat16))
    from building_llm_from_scratch_b200 import _lib
    with pytest.raises(_lib.B200Error):
        K.gemm_ex(rnd(16, 12), rnd(16, 12))   # K = 12 breaks the 16-byte row pitch TMA needs


def test_gemm_grouped_matches_individual_launches(K):
    """Several LoRA weight-gradient problems of different sizes in one grouped launch."""
    Mtok = 2048
    probs, refs = [], []
    for i, (d_in, d_out) in enumerate([(512, 768), (256, 2048), (1024, 264), (2048, 512)]):
        x, u = rnd(Mtok, d_in, seed=10 + i), rnd(Mtok, 64, scale=0.1, seed=20 + i)
        dY, t = rnd(Mtok, d_out, scale=0.1, seed=30 + i), rnd(Mtok, 64, seed=40 + i)
        dA = torch.full((d_in, 64), 0.5, device="cuda", dtype=torch.float32)
        dB = torch.zeros(64, d_out, device="cuda", dtype=torch.float32)
        probs.append(dict(a=x, b=u, out=dA, a_mn=True, b_mn=True, accumulate=True))
        probs.append(dict(a=dY, b=t, out=dB, a_mn=True, b_mn=True, accumulate=True, trans_out=True))
        refs.append((dA, 0.5 + x.float().t() @ u.float()))
        refs.append((dB, t.float().t() @ dY.float()))
    # plus a plain bf16 (non-accumulating) rank-r projection t = x A, A read as an MN-major operand
    x, A = rnd(Mtok, 512, seed=50), rnd(512, 64, scale=0.1, seed=51)
    tt = torch.empty(Mtok, 64, device="cuda", dtype=torch.bfloat16)
    probs.append(dict(a=x, b=A, out=tt, b_mn=True, alpha=0.5))
    refs.append((tt, 0.5 * x.float() @ A.float()))
    K.gemm_grouped(probs)
    for got, ref in refs:
        assert rel_err(got.float(), ref) < (1e-3 if got.dtype == torch.float32 else BF16_TOL)


@pytest.mark.parametrize("streamk", [True, False])
def test_gemm_grouped_schedulers_agree_with_fp32(K, streamk, monkeypatch, request):
    """Both schedulers of the grouped launch (stream-K with shared-tile fix-up, tile-scheduled) on the
    shapes that stress the fix-up: K ranges cut across CTAs (vocab-sized K), ragged M, rank 8,
    fp32 and transposed stores; the stream-K scratch must be all zero again afterwards."""
    monkeypatch.setattr(K, "STREAMK", streamk)
    from building_llm_from_scratch_b200 import _lib
    _lib.gemm_set_option(_lib.GEMM_OPT_STREAMK, 1)        # otherwise the host may pick whole tiles
    request.addfinalizer(lambda: _lib.gemm_set_option(_lib.GEMM_OPT_STREAMK, -1))
    probs, refs = [], []
    # u = s * dY B^T of a vocab projection: 3 row tiles (ragged) x 501 K blocks
    dl, Bh = rnd(300, 32064, scale=0.05, seed=60), rnd(64, 32064, scale=0.1, seed=61)
    u = torch.empty(300, 64, device="cuda", dtype=torch.bfloat16)
    probs.append(dict(a=dl, b=Bh, out=u, alpha=0.5))
    refs.append((u, 0.5 * dl.float() @ Bh.float().t()))
    # t = x A, rank 8
    x, A8 = rnd(1000, 2048, seed=62), rnd(2048, 8, scale=0.1, seed=63)
    t8 = torch.empty(1000, 8, device="cuda", dtype=torch.bfloat16)
    probs.append(dict(a=x, b=A8, out=t8, b_mn=True))
    refs.append((t8, x.float() @ A8.float()))
    # fp32 store without accumulation (first weight-gradient contribution), transposed
    dY, t = rnd(4096, 640, scale=0.1, seed=64), rnd(4096, 64, seed=65)
    dB = torch.full((64, 640), 7.0, device="cuda", dtype=torch.float32)
    probs.append(dict(a=dY, b=t, out=dB, a_mn=True, b_mn=True, trans_out=True))
    refs.append((dB, t.float().t() @ dY.float()))
    # accumulation on top of existing contents
    dA = torch.ones(2048, 8, device="cuda", dtype=torch.float32)
    u8 = rnd(1000, 8, scale=0.1, seed=66)
    probs.append(dict(a=x, b=u8, out=dA, a_mn=True, b_mn=True, accumulate=True))
    refs.append((dA, 1 + x.float().t() @ u8.float()))
    K.gemm_grouped(probs)
    K.gemm_grouped(probs[:3])          # second launch over the same scratch (dA left out: it accumulates)
    for got, ref in refs:
        assert rel_err(got.float(), ref) < (1e-3 if got.dtype == torch.float32 else BF16_TOL)
    if streamk:
        ws = K._streamk_ws[torch.cuda.current_device()]
        assert int(torch.count_nonzero(ws)) == 0


@pytest.mark.parametrize("force", ["1", ""])
def test_skinny_gemm_ex_routes_through_streamk_exactly(K, force, request):
    """Integer-valued inputs: a single skinny product (one problem, K split over all SMs when forced,
    scheduler's own choice otherwise) is bit-exact."""
    from building_llm_from_scratch_b200 import _lib
    if force:
        _lib.gemm_set_option(_lib.GEMM_OPT_STREAMK, 1)
        request.addfinalizer(lambda: _lib.gemm_set_option(_lib.GEMM_OPT_STREAMK, -1))
    before = _lib.gemm_path_counts()
    g = torch.Generator(device="cuda").manual_seed(5)
    a = torch.randint(-3, 4, (8192, 2048), generator=g, device="cuda").to(torch.bfloat16)
    b = torch.randint(-3, 4, (64, 2048), generator=g, device="cuda").to(torch.bfloat16)
    got = K.gemm_ex(a, b, alpha=0.125)
    ref = (0.125 * (a.float() @ b.float().t())).to(torch.bfloat16)
    assert torch.equal(got, ref)
    if force:
        assert _lib.gemm_path_counts()["streamk"] == before["streamk"] + 1


@pytest.mark.parametrize("M,N,K_,r", [(8192, 2048, 4096, 64), (8192, 3072, 8192, 0), (8000, 2000, 4104, 8),
                                     (8192, 2048, 2048, 64), (3584, 2048, 8192, 0)])
def test_gemm_tail_split_wave(K, M, N, K_, r, request):
    """Shapes whose last wave of 256 x 256 tiles is at most half full (256 or 384 tiles on 74 CTA pairs) and
    whose K range is deep enough to split (the K = 2048 case is below the threshold and runs unsplit), and one whose
    last wave is just over half full (112 tiles: 38 tail tiles go out as 114 thirds in two rounds):
    the tail tiles are cut along K and summed through the registered workspace. Same results as with
    tail splitting off (bias + residual + LoRA K extension in the epilogue of the fix-up), exact for
    integer-valued inputs, arrival counters zero afterwards."""
    a, b = rnd(M, K_, seed=1), rnd(N, K_, scale=0.05, seed=2)
    bias, res = rnd(N, seed=3), rnd(M, N, seed=4)
    kw = dict(a2=rnd(M, r, seed=5), b2=rnd(N, r, scale=0.1, seed=6)) if r else {}
    from building_llm_from_scratch_b200 import _lib
    before = _lib.gemm_path_counts()
    got = K.gemm_ex(a, b, bias=bias, residual=res, **kw)
    split = _lib.gemm_path_counts()["pair_tail_split"] - before["pair_tail_split"]
    assert split == (1 if K_ >= 4096 else 0)               # deep K ranges are split, K = 2048 is not
    ref = a.float() @ b.float().t() + bias.float() + res.float()
    if r:
        ref = ref + kw["a2"].float() @ kw["b2"].float().t()
    assert rel_err(got.float(), ref) < BF16_TOL
    _lib.gemm_set_option(_lib.GEMM_OPT_TAIL_SPLIT, 0)
    request.addfinalizer(lambda: _lib.gemm_set_option(_lib.GEMM_OPT_TAIL_SPLIT, -1))
    plain = K.gemm_ex(a, b, bias=bias, residual=res, **kw)
    _lib.gemm_set_option(_lib.GEMM_OPT_TAIL_SPLIT, -1)
    assert rel_err(got.float(), plain.float()) < 4e-3          # only the fp32 summation order differs
    g = torch.Generator(device="cuda").manual_seed(9)
    ai = torch.randint(-2, 3, (M, K_), generator=g, device="cuda").to(torch.bfloat16)
    bi = torch.randint(-2, 3, (N, K_), generator=g, device="cuda").to(torch.bfloat16)
    exact = K.gemm_ex(ai, bi, alpha=0.25)
    assert torch.equal(exact, (0.25 * (ai.float() @ bi.float().t())).to(torch.bfloat16))
    assert int(torch.count_nonzero(K._gemm_ws[torch.cuda.current_device()][:4096])) == 0    # arrival counters


@pytest.mark.parametrize("M,Ks,N,share,a_mn", [(8192, (2048,), 64, False, False), (8192, (2048, 2048, 2048), 64, True, False),
                                               (8192, (2048, 2048), 64, True, False), (8192, (2048, 512, 512), 64, False, False),
                                               (8192, (8192,), 64, False, False), (8000, (2000,), 8, False, False),
                                               (8192, (128,), 64, False, False), (8192, (4096, 4096), 64, False, True),
                                               (4096, (1280, 1280), 8, True, False)])
def test_gemm_grouped_on_k_splitting_cta_pairs(K, M, Ks, N, share, a_mn, request):
    """Non-accumulating skinny groups that fill the machine as CTA pairs (rank-64 LoRA products of the bench shapes: t = x A for
    q | k | v sharing x, u = dY B of three different dY, K = 8192, ragged M / K / N, MN-major A): each pair splits K and
    exchanges partial sums through distributed shared memory. Same results as stream-K / tile scheduling up to the fp32
    summation order, fp32 and transposed outputs included."""
    from building_llm_from_scratch_b200 import _lib
    probs, refs = [], []
    x_shared = rnd(M, Ks[0], seed=1)
    for i, K_ in enumerate(Ks):
        a = x_shared if share else rnd(M, K_, seed=10 + i)
        b = rnd(K_, N, scale=0.05, seed=20 + i)                       # stored [K, N]: b_mn
        a_arg = a.t().contiguous() if a_mn else a
        out_dtype = torch.float32 if i == 1 else torch.bfloat16
        trans = i == 2
        out = torch.empty((N, M) if trans else (M, N), device="cuda", dtype=out_dtype)
        probs.append(dict(a=a_arg, b=b, out=out, a_mn=a_mn, b_mn=True, alpha=0.5, trans_out=trans))
        refs.append(0.5 * (a.float() @ b.float()))
    before = _lib.gemm_path_counts()["skinny_pair"]
    K.gemm_grouped(probs)
    assert _lib.gemm_path_counts()["skinny_pair"] == before + 1
    got = [(pr["out"].t() if pr["trans_out"] else pr["out"]).float().clone() for pr in probs]
    for gt, ref in zip(got, refs):
        assert rel_err(gt, ref) < BF16_TOL
    _lib.gemm_set_option(_lib.GEMM_OPT_SKINNY_PAIR, 0)
    request.addfinalizer(lambda: _lib.gemm_set_option(_lib.GEMM_OPT_SKINNY_PAIR, -1))
    for pr in probs:
        pr["out"].zero_()
    K.gemm_grouped(probs)
    assert _lib.gemm_path_counts()["skinny_pair"] == before + 1
    for gt, pr in zip(got, probs):
        other = (pr["out"].t() if pr["trans_out"] else pr["out"]).float()
        assert rel_err(gt, other) < 4e-3


@pytest.mark.parametrize("M,outs,K_,r", [(8192, (2048, 512, 512), 2048, 64), (1024, (768, 256), 512, 8),
                                        (300, (256, 256, 256, 512), 128, 16)])
def test_gemm_lora_packed_segments(K, M, outs, K_, r):
    """q | k | v style packed forward: one GEMM, each 256-column tile adds its own segment's LoRA update."""
    x = rnd(M, K_, seed=1)
    ws = [rnd(o, K_, scale=0.05, seed=10 + i) for i, o in enumerate(outs)]
    ts = [rnd(M, r, seed=20 + i) for i in range(len(outs))]
    Bs = [rnd(r, o, scale=0.1, seed=30 + i) for i, o in enumerate(outs)]
    starts = [0]
    for o in outs[:-1]:
        starts.append(starts[-1] + o)
    got = K.gemm_lora_packed(x, torch.cat(ws, 0), torch.cat(ts, 1), torch.cat(Bs, 1), r, starts)
    ref = torch.cat([x.float() @ w.float().t() + t.float() @ B.float() for w, t, B in zip(ws, ts, Bs)], dim=1)
    assert rel_err(got.float(), ref) < BF16_TOL
    for i, o in enumerate(outs):       # per segment too: a wrong t block would hide in the global norm
        sl = slice(starts[i], starts[i] + o)
        assert rel_err(got[:, sl].float(), ref[:, sl]) < BF16_TOL, i


@pytest.mark.parametrize("M,F,K_,r", [(8192, 8192, 2048, 64), (512, 768, 256, 8), (1000, 1024, 512, 0)])
def test_gemm_swiglu_backward_epilogue(K, M, F, K_, r):
    """(da | db) from the epilogue of dh = dY W (+ u A^T) against the two-kernel path (GEMM, then swiglu_bwd)
    and against autograd of the oracle's SwiGLU on the same dh."""
    from oracle import reference_ops as R
    dy, w = rnd(M, K_, scale=0.1, seed=1), rnd(K_, F, scale=0.05, seed=2)
    ab = rnd(M, 2 * F, seed=3)
    u = rnd(M, r, scale=0.1, seed=4) if r else None
    A = rnd(F, r, scale=0.1, seed=5) if r else None
    got = K.gemm_swiglu_bwd(dy, w, u, A, ab)
    dh = dy.float() @ w.float()
    if r:
        dh = dh + u.float() @ A.float().t()
    a = ab[:, :F].float().requires_grad_(True)
    b = ab[:, F:].float().requires_grad_(True)
    (torch.nn.functional.silu(a) * b).backward(dh)
    assert rel_err(got[:, :F].float(), a.grad) < BF16_TOL
    assert rel_err(got[:, F:].float(), b.grad) < BF16_TOL
    dh16 = K.gemm_ex(dy, w, b1_mn=True, a2=u, b2=A) if r else K.gemm_ex(dy, w, b1_mn=True)
    da, db = K.swiglu_bwd(ab[:, :F], ab[:, F:], dh16)
    assert rel_err(got[:, :F].float(), da.float()) < 1e-2 and rel_err(got[:, F:].float(), db.float()) < 1e-2


def _interleave_rows(w1, w2, blk=128):
    F, Kd = w1.shape
    return torch.stack([w1.view(F // blk, blk, Kd), w2.view(F // blk, blk, Kd)], dim=1).reshape(2 * F, Kd).contiguous()


def _interleave_lora_b(b1, b2, blk=128):
    r, F = b1.shape
    out = torch.zeros(2 * r, F // blk, 2, blk, device=b1.device, dtype=b1.dtype)
    out[:r, :, 0, :] = b1.view(r, F // blk, blk)
    out[r:, :, 1, :] = b2.view(r, F // blk, blk)
    return out.view(2 * r, 2 * F)


@pytest.mark.parametrize("M,F,K_,r", [(8192, 8192, 2048, 64), (512, 384, 256, 8), (1000, 1024, 512, 0)])
def test_gemm_swiglu_forward_epilogue(K, M, F, K_, r):
    """(a | b) and h = silu(a) * b from one GEMM over row-interleaved fc1 / fc2 weights against the separate
    path: two (LoRA) linears and the SwiGLU kernel (same bf16 roundings; the LoRA blocks are accumulated in one more
    K block, so agreement is to rounding, not bit-exact)."""
    x = rnd(M, K_, seed=1)
    w1, w2 = rnd(F, K_, scale=0.05, seed=2), rnd(F, K_, scale=0.05, seed=3)
    kw1 = kw2 = {}
    t2 = bil = None
    if r:
        t1, t2_ = rnd(M, r, seed=4), rnd(M, r, seed=5)
        b1, b2 = rnd(r, F, scale=0.1, seed=6), rnd(r, F, scale=0.1, seed=7)
        kw1, kw2 = dict(a2=t1, b2=b1, b2_mn=True), dict(a2=t2_, b2=b2, b2_mn=True)
        t2, bil = torch.cat([t1, t2_], dim=1), _interleave_lora_b(b1, b2)
    ab, h = K.gemm_swiglu_fwd(x, _interleave_rows(w1, w2), t2, bil)
    a_ref, b_ref = K.gemm_ex(x, w1, **kw1), K.gemm_ex(x, w2, **kw2)
    h_ref = K.swiglu_fwd(a_ref, b_ref)
    assert rel_err(ab[:, :F].float(), a_ref.float()) < 2e-3 and rel_err(ab[:, F:].float(), b_ref.float()) < 2e-3
    assert rel_err(h.float(), h_ref.float()) < 1e-2
    fa = x.float() @ w1.float().t() + (t1.float() @ b1.float() if r else 0)
    assert rel_err(ab[:, :F].float(), fa) < BF16_TOL


# ------------------------------------------------------------------ GPT-2 epilogue fusions
def _gelu(x):
    return 0.5 * x * (1.0 + torch.erf(x / 2 ** 0.5))


def _gelu_grad(x):
    return 0.5 * (1.0 + torch.erf(x / 2 ** 0.5)) + x * torch.exp(-0.5 * x * x) / (2 * torch.pi) ** 0.5


@pytest.mark.parametrize("M,N,K_,r", [(4096, 6400, 1600, 0), (512, 5120, 1280, 8), (300, 320, 64, 0)])
def test_gemm_gelu_forward_epilogue(K, M, N, K_, r):
    """z = x W^T + b (+ t B) and h = gelu(z) from one launch (Models/GPT2/GPT2.py:58-62); h is the GELU of the bf16 z."""
    x, w, b = rnd(M, K_, seed=1), rnd(N, K_, scale=0.05, seed=2), rnd(N, seed=3)
    t, Bl = (rnd(M, r, seed=4), rnd(r, N, scale=0.1, seed=5)) if r else (None, None)
    h, z = K.gemm_gelu_fwd(x, w, b, t, Bl)
    ref = x.float() @ w.float().t() + b.float()
    if r:
        ref = ref + t.float() @ Bl.float()
    assert rel_err(z.float(), ref) < BF16_TOL
    assert rel_err(h.float(), _gelu(z.float())) < 4e-3        # exact up to the final bf16 rounding
    assert rel_err(h.float(), _gelu(ref)) < BF16_TOL


@pytest.mark.parametrize("M,N,K_,r", [(4096, 6400, 1600, 0), (512, 5120, 1280, 8), (384, 1600, 256, 0)])
def test_gemm_gelu_backward_epilogue(K, M, N, K_, r):
    """dz = gelu'(z) * (dy W + u A^T) in the epilogue of the input-gradient GEMM; N = 1600 / 6400 have a partly
    filled last 256-column tile."""
    dy, w, z = rnd(M, K_, scale=0.1, seed=1), rnd(K_, N, scale=0.05, seed=2), rnd(M, N, scale=1.5, seed=3)
    u, A = (rnd(M, r, seed=4), rnd(N, r, scale=0.1, seed=5)) if r else (None, None)
    dz = K.gemm_gelu_bwd(dy, w, z, u, A)
    dh = dy.float() @ w.float()
    if r:
        dh = dh + u.float() @ A.float().t()
    assert rel_err(dz.float(), dh * _gelu_grad(z.float())) < BF16_TOL
    assert rel_err(dz.float(), K.gelu_bwd(z, dh.to(torch.bfloat16)).float()) < BF16_TOL


@pytest.mark.parametrize("M,N,K_,r", [(4096, 1600, 1600, 0), (4096, 1600, 6400, 0), (512, 1280, 1280, 8)])
def test_gemm_dropout_residual_epilogue_draws_the_standalone_mask(K, M, N, K_, r, request):
    """y = res + dropout(x W^T + b): the fused epilogue and the GEMM followed by the dropout kernel agree bit for bit
    (same counter-based mask, same bf16 roundings), also with a device-side seed offset; about p of the linear's
    outputs are dropped."""
    p, seed = 0.1, 1234567
    x, w, b, res = rnd(M, K_, seed=1), rnd(N, K_, scale=0.05, seed=2), rnd(N, seed=3), rnd(M, N, seed=4)
    t, Bl = (rnd(M, r, seed=5), rnd(r, N, scale=0.1, seed=6)) if r else (None, None)
    from building_llm_from_scratch_b200 import _lib
    _lib.gemm_set_option(_lib.GEMM_OPT_TAIL_SPLIT, 0)        # the fused epilogue never K-splits its tail wave; a split
    request.addfinalizer(lambda: _lib.gemm_set_option(_lib.GEMM_OPT_TAIL_SPLIT, -1))   # changes the fp32 summation order
    lin = K.gemm_ex(x, w, a2=t, b2=Bl, b2_mn=True, bias=b)
    _lib.gemm_set_option(_lib.GEMM_OPT_TAIL_SPLIT, -1)
    for off in (None, torch.tensor([77], device="cuda", dtype=torch.int64)):
        fused = K.gemm_dropout_residual(x, w, b, res, p, seed, off, t, Bl)
        plain = K.dropout(lin, p, seed, res, seed_dev=off)
        assert torch.equal(fused, plain)
        dropped = (fused.float() == res.float()).float().mean()
        assert abs(float(dropped) - p) < 0.01
    assert not torch.equal(fused, K.gemm_dropout_residual(x, w, b, res, p, seed, None, t, Bl))     # the offset matters


# ------------------------------------------------------------------ RoPE in the q|k|v projection's epilogue
def _rope_tables(T, hd, base=7812.5):
    inv = 1.0 / (base ** (torch.arange(0, hd, 2).float() / hd))
    ang = torch.arange(T)[:, None] * inv[None, :]
    ang = torch.cat([ang, ang], dim=1)
    return torch.cos(ang).to(torch.bfloat16).cuda(), torch.sin(ang).to(torch.bfloat16).cuda()


@pytest.mark.parametrize("B,T,h,hkv,hd,d,r", [(4, 2048, 32, 8, 64, 2048, 64), (2, 512, 8, 4, 128, 1024, 0), (1, 300, 4, 4, 64, 256, 8)])
def test_gemm_qkv_rope_epilogue_equals_projection_then_rope_kernel(K, B, T, h, hkv, hd, d, r):
    """compute_rope (common_components.py:6-35) in the epilogue of the packed q|k|v GEMM: bit-identical to the packed GEMM
    followed by the stand-alone rotation kernel (same bf16 roundings); the v columns are untouched."""
    M = B * T
    outs = (h * hd, hkv * hd, hkv * hd)
    x = rnd(M, d, seed=1)
    w_cat = rnd(sum(outs), d, scale=0.05, seed=2)
    cos, sin = _rope_tables(T, hd)
    starts = [0, outs[0], outs[0] + outs[1]]
    if r:
        t_cat, b_cat = rnd(M, 3 * r, seed=3), rnd(r, sum(outs), scale=0.1, seed=4)
        plain = K.gemm_lora_packed(x, w_cat, t_cat, b_cat, r, starts)
        fused = K.gemm_qkv_rope(x, w_cat, t_cat, b_cat, r, starts, cos, sin, T, hd, (h + hkv) * hd)
    else:
        plain = K.gemm_ex(x, w_cat)
        fused = K.gemm_qkv_rope(x, w_cat, None, None, 0, starts, cos, sin, T, hd, (h + hkv) * hd)
    ref = plain.clone()
    K.rope_(ref, cos, sin, B, T, h + hkv, hd)
    assert torch.equal(fused, ref)
    assert not torch.equal(fused[:, :outs[0]], plain[:, :outs[0]])


@pytest.mark.parametrize("B,T,h,hkv,hd", [(2, 512, 8, 2, 64), (1, 640, 4, 2, 128)])
def test_attention_backward_with_rope_transpose_in_the_epilogues(K, B, T, h, hkv, hd):
    """b200_attention_bwd_rope_bf16 = attention backward followed by the inverse rotation of dq | dk (compute_rope under
    autograd); the fused form rotates the fp32 accumulators, so it is compared at a few bf16 ulps."""
    qkv = rnd(B * T, (h + 2 * hkv) * hd, seed=1)
    nq, nk = h * hd, hkv * hd
    q, k, v = qkv[:, :nq], qkv[:, nq:nq + nk], qkv[:, nq + nk:]
    do = rnd(B * T, nq, scale=0.1, seed=2)
    cos, sin = _rope_tables(T, hd)
    o, lse = K.attention_fwd(q, k, v, B, T, h, hkv, hd)
    dq, dk, dv = K.attention_bwd(q, k, v, o, do, lse, B, T, h, hkv, hd)
    K.rope_(dq, cos, sin, B, T, h, hd, inverse=True)
    K.rope_(dk, cos, sin, B, T, hkv, hd, inverse=True)
    fq, fk, fv = K.attention_bwd(q, k, v, o, do, lse, B, T, h, hkv, hd, rope=(cos, sin))
    assert torch.equal(fv, dv)
    assert rel_err(fq.float(), dq.float()) < 6e-3
    assert rel_err(fk.float(), dk.float()) < 6e-3
