"""GPU: bandwidth kernels + attention through the C-ABI against the oracle on seeded inputs and
against the committed fixtures recorded from the live reference. Tolerances: bf16 2e-2 relative,
fp32 1e-4 relative (north star); integer work bit-exact."""
import math

import pytest
import torch

from conftest import close, rel_err

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]

BF16_TOL = 2e-2
FP32_TOL = 1e-4


@pytest.fixture(scope="module")
def K():
    from building_llm_from_scratch_b200 import kernels
    return kernels


def rnd(*shape, scale=1.0, seed=0, dtype=torch.bfloat16):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(*shape, generator=g, device="cuda") * scale).to(dtype)


# ------------------------------------------------------------------ RMSNorm
@pytest.mark.parametrize("M,d", [(8192, 2048), (333, 4096), (7, 64), (100, 1600), (64, 8192), (50, 5120), (9, 16384)])
def test_rmsnorm_bf16_matches_reference_roundings(K, M, d):
    from oracle import reference_ops as R
    x, w = rnd(M, d, scale=2.0, seed=1), (1 + 0.1 * rnd(d, seed=2, dtype=torch.float32))
    y = K.rmsnorm_fwd(x, w)
    ref = R.rmsnorm(x, w)
    assert rel_err(y.float(), ref.float()) < 2e-3
    # the bf16_steps mode reproduces the eager roundings: almost every element identical
    assert (y != ref).float().mean() < 0.02
    xr = x.clone().float().requires_grad_(True)
    gy = rnd(M, d, seed=3)
    R.rmsnorm(xr, w).backward(gy.float())
    dx, _ = K.rmsnorm_bwd(x, w, gy)
    assert rel_err(dx.float(), xr.grad) < BF16_TOL


def test_rmsnorm_golden_and_dw_and_residual(K, golden):
    g = golden("llama_modules.pt")["rmsnorm/torch.bfloat16"]
    x, w, gy = g["x"].cuda(), g["w"].cuda(), g["gy"].cuda()
    # the fixture was recorded by CPU eager, whose bf16 mean rounds once more than CUDA eager
    # (which the kernel's bf16_steps mode follows): identical up to 1 ulp on a few elements
    y = K.rmsnorm_fwd(x, w).cpu()
    assert rel_err(y.float(), g["y"].float()) < 5e-3   # a row whose rstd differs by 1 bf16 ulp shifts whole
    dx, dw = K.rmsnorm_bwd(x, w, gy, need_dw=True)
    assert rel_err(dx.cpu().float(), g["dx"].float()) < BF16_TOL
    assert rel_err(dw.cpu(), g["dw"].float()) < BF16_TOL
    f = golden("llama_modules.pt")["rmsnorm/torch.float32"]
    y32 = K.rmsnorm_fwd(f["x"].cuda(), f["w"].cuda())
    assert rel_err(y32.cpu(), f["y"]) < FP32_TOL
    dx32, dw32 = K.rmsnorm_bwd(f["x"].cuda(), f["w"].cuda(), f["gy"].cuda(), need_dw=True)
    assert rel_err(dx32.cpu(), f["dx"]) < FP32_TOL and rel_err(dw32.cpu(), f["dw"]) < FP32_TOL
    # fused residual add (Llama3.py:178,180): s = x + r in bf16, y = norm(s)
    from oracle import reference_ops as R
    a, r = rnd(300, 2048, seed=4), rnd(300, 2048, seed=5)
    wt = 1 + 0.1 * rnd(2048, seed=6, dtype=torch.float32)
    y, s = K.rmsnorm_fwd(a, wt, res=r)
    assert torch.equal(s, a + r)
    assert rel_err(y.float(), R.rmsnorm(a + r, wt).float()) < 2e-3
    dres = rnd(300, 2048, seed=7)
    dx, _ = K.rmsnorm_bwd(s, wt, gy := rnd(300, 2048, seed=8), dres=dres)
    dx0, _ = K.rmsnorm_bwd(s, wt, gy)
    assert rel_err(dx.float(), dx0.float() + dres.float()) < 1e-2


@pytest.mark.parametrize("M,d", [(257, 1024), (130, 4096), (35, 8192), (10, 16384)])
def test_rmsnorm_wide_rows_weight_gradient_and_residual_stream(K, M, d):
    """Rows covered by several cooperating warps: weight gradient, residual-stream gradient and fused
    residual add at every width class, ragged row counts."""
    from oracle import reference_ops as R
    x, r = rnd(M, d, seed=11), rnd(M, d, seed=12)
    w = 1 + 0.1 * rnd(d, seed=13, dtype=torch.float32)
    y, s = K.rmsnorm_fwd(x, w, res=r)
    assert torch.equal(s, x + r)
    assert rel_err(y.float(), R.rmsnorm(x + r, w).float()) < 2e-3
    gy, dres = rnd(M, d, seed=14), rnd(M, d, seed=15)
    xr = x.float().requires_grad_(True)
    wr = w.clone().requires_grad_(True)
    R.rmsnorm(xr, wr).backward(gy.float())
    dx, dw = K.rmsnorm_bwd(x, w, gy, dres=dres, need_dw=True)
    assert rel_err(dx.float(), xr.grad + dres.float()) < BF16_TOL
    assert rel_err(dw, wr.grad) < BF16_TOL


# ------------------------------------------------------------------ LayerNorm / GELU
@pytest.mark.parametrize("M,d", [(4096, 1600), (100, 768), (33, 1280)])
def test_layernorm(K, M, d):
    from oracle import reference_ops as R
    x, w, b = rnd(M, d, scale=1.5, seed=1), 1 + 0.1 * rnd(d, seed=2), 0.1 * rnd(d, seed=3)
    y, mean, rstd = K.layernorm_fwd(x, w, b)
    xr = x.float().requires_grad_(True)
    wr, br = w.float().requires_grad_(True), b.float().requires_grad_(True)
    ref = R.layernorm(xr, wr, br)
    assert rel_err(y.float(), ref) < BF16_TOL
    gy = rnd(M, d, seed=4)
    ref.backward(gy.float())
    dx, dw, db = K.layernorm_bwd(x, w, gy, mean, rstd)
    assert rel_err(dx.float(), xr.grad) < BF16_TOL
    assert rel_err(dw, wr.grad) < BF16_TOL and rel_err(db, br.grad) < BF16_TOL
    # the residual-stream gradient added while dx is written (with and without the dw / db pass)
    dres = rnd(M, d, seed=5)
    for need_dw in (True, False):
        dx2, dw2, db2 = K.layernorm_bwd(x, w, gy, mean, rstd, need_dw=need_dw, dres=dres)
        assert rel_err(dx2.float(), xr.grad + dres.float()) < BF16_TOL
        if need_dw:
            assert torch.equal(dw2, dw) and torch.equal(db2, db) or rel_err(dw2, dw) < 1e-5


def test_gelu(K):
    from oracle import reference_ops as R
    x = rnd(1000, 3077, scale=2.0, seed=1).reshape(-1)
    xr = x.float().requires_grad_(True)
    ref = R.gelu_erf(xr)
    assert rel_err(K.gelu_fwd(x).float(), ref) < BF16_TOL
    gy = rnd(x.numel(), seed=2)
    ref.backward(gy.float())
    assert rel_err(K.gelu_bwd(x, gy).float(), xr.grad) < BF16_TOL


# ------------------------------------------------------------------ SwiGLU / RoPE / embedding / casts
def test_swiglu(K):
    M, F = 1000, 8192
    ab = rnd(M, 2 * F, scale=1.5, seed=1)
    a, b = ab[:, :F], ab[:, F:]
    ref_bf16 = (a * torch.sigmoid(a)) * b        # the reference's eager bf16 op sequence
    h = K.swiglu_fwd(a, b)
    assert (h != ref_bf16).float().mean() < 0.02 and rel_err(h.float(), ref_bf16.float()) < 2e-3
    ar, br = a.float().requires_grad_(True), b.float().requires_grad_(True)
    gy = rnd(M, F, seed=2)
    ((ar * torch.sigmoid(ar)) * br).backward(gy.float())
    da, db = K.swiglu_bwd(a, b, gy)
    assert rel_err(da.float(), ar.grad) < BF16_TOL and rel_err(db.float(), br.grad) < BF16_TOL


def test_rope_golden_and_layout(K, golden):
    from oracle import reference_ops as R
    g = golden("llama_modules.pt")["rope/torch.bfloat16"]
    x = g["x"].cuda()                                   # [B, H, T, hd]
    B, H, T, hd = x.shape
    flat = x.transpose(1, 2).reshape(B * T, H * hd).contiguous()
    K.rope_(flat, g["cos"].cuda(), g["sin"].cuda(), B, T, H, hd)
    got = flat.view(B, T, H, hd).transpose(1, 2)
    assert torch.equal(got.cpu(), g["y"])
    gflat = g["gy"].cuda().transpose(1, 2).reshape(B * T, H * hd).contiguous()
    K.rope_(gflat, g["cos"].cuda(), g["sin"].cuda(), B, T, H, hd, inverse=True)
    assert rel_err(gflat.view(B, T, H, hd).transpose(1, 2).cpu().float(), g["dx"].float()) < BF16_TOL
    # big: q and k heads inside a packed qkv row, v untouched
    B, T, h, kv, hd = 2, 512, 32, 8, 64
    cos, sin = R.precompute_rope_params(hd, 7812.5, 2048, None)
    cos, sin = cos.bfloat16().cuda(), sin.bfloat16().cuda()
    qkv = rnd(B * T, (h + 2 * kv) * hd, seed=3)
    orig = qkv.clone()
    K.rope_(qkv, cos, sin, B, T, h + kv, hd)
    q = orig[:, :h * hd].view(B, T, h, hd).transpose(1, 2)
    assert torch.equal(qkv[:, :h * hd].view(B, T, h, hd).transpose(1, 2), R.rope(q, cos, sin))
    assert torch.equal(qkv[:, (h + kv) * hd:], orig[:, (h + kv) * hd:])


@pytest.mark.parametrize("M,N,ld", [(4096, 1600, 1600), (4096, 6400, 6400), (300, 264, 272), (1000, 50257, 50264), (77, 8, 8),
                                    (5000, 2048, 6144)])
def test_colsum_bias_gradient(K, M, N, ld):
    """out[n] (+)= sum_m x[m, n] (the bias gradient of nn.Linear): tiled kernel for 16-byte rows, first form otherwise
    (N = 50257); accumulates into an existing fp32 buffer; column slices of a wider buffer."""
    buf = rnd(M, ld, seed=1)
    x = buf[:, :N]
    ref = x.float().sum(0)
    got = K.colsum(x)
    assert got.dtype == torch.float32 and rel_err(got, ref) < 1e-5
    acc = torch.full((N,), 2.0, device="cuda")
    K.colsum(x, out=acc)
    assert rel_err(acc, ref + 2.0) < 1e-5


def test_embedding_and_casts(K):
    V, d, B, T = 1000, 256, 3, 37
    w, pw = rnd(V, d, seed=1), rnd(64, d, seed=2)
    idx = torch.randint(0, V, (B, T), device="cuda")
    assert torch.equal(K.embedding_fwd(idx, w), w[idx])
    assert torch.equal(K.embedding_fwd(idx, w, pw), (w[idx].float() + pw[:T].float()[None]).bfloat16())
    dout = rnd(B, T, d, seed=3)
    dtok, dpos = K.embedding_bwd(idx, dout, V, 64)
    ref = torch.zeros(V, d, device="cuda").index_add_(0, idx.reshape(-1), dout.reshape(-1, d).float())
    assert rel_err(dtok, ref) < 1e-5
    assert rel_err(dpos[:T], dout.float().sum(0)) < 1e-5
    x = rnd(100003, seed=4, dtype=torch.float32)
    assert torch.equal(K.cast(x, torch.bfloat16), x.bfloat16())
    assert torch.equal(K.cast(x.bfloat16(), torch.float32), x.bfloat16().float())
    m = rnd(70, 130, seed=5)
    assert torch.equal(K.transpose(m), m.t().contiguous())
    assert torch.equal(K.add(m, m), m + m)


# ------------------------------------------------------------------ cross-entropy / optimiser
@pytest.mark.parametrize("M,V,ld", [(12, 37, 40), (512, 50257, 50264), (256, 128256, 128256)])
def test_cross_entropy(K, M, V, ld):
    from oracle import reference_ops as R
    buf = rnd(M, ld, scale=3.0, seed=1)
    logits = buf[:, :V]
    tg = torch.randint(0, V, (M,), device="cuda")
    tg[torch.rand(M, device="cuda") < 0.5] = -100
    tg[0] = 5
    lr = logits.float().clone().requires_grad_(True)
    ref = R.cross_entropy(lr, tg)
    ref.backward()
    loss = K.cross_entropy_(logits, tg)
    assert rel_err(loss, ref) < 1e-4
    assert rel_err(logits.float(), lr.grad) < BF16_TOL
    assert torch.all(logits[tg == -100] == 0)


def test_cross_entropy_golden_and_all_ignored(K, golden):
    g = golden("schedule_ce.pt")
    buf = torch.zeros(12, 40, device="cuda", dtype=torch.bfloat16)
    buf[:, :37] = g["logits"].bfloat16().cuda()
    loss = K.cross_entropy_(buf[:, :37], g["targets"].cuda())
    assert rel_err(loss.cpu(), g["loss_bf16"].float()) < BF16_TOL
    assert rel_err(buf[:, :37].cpu().float(), g["dlogits_bf16"].float()) < BF16_TOL
    nan = K.cross_entropy_(buf[:, :37], torch.full((12,), -100, device="cuda"))
    assert math.isnan(float(nan))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_adamw_and_grad_norm_golden(K, golden, dtype):
    g = golden("optimizer.pt")[str(dtype)]
    p = g["p0"].cuda().clone()
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    for step, ((gp, gq), st, nrm, lr) in enumerate(zip(g["grads"], g["states"], g["norms"], g["lrs"]), start=1):
        flat = torch.cat([gp.reshape(-1), gq.reshape(-1)]).cuda().bfloat16()
        out = K.grad_norm(flat, 1.0)
        assert rel_err(out[0].cpu(), nrm.float()) < 5e-3
        coef = out[1:2] if dtype == torch.bfloat16 else torch.clamp(1.0 / (nrm.float().cuda() + 1e-6), max=1.0).reshape(1)
        K.adamw_(p, gp.cuda(), m, v, step, lr, grad_scale=coef)
        tol = FP32_TOL if dtype == torch.float32 else BF16_TOL
        assert rel_err(p.cpu().float(), st["p"].float()) < tol
        assert rel_err(m.cpu().float(), st["m_p"].float()) < tol
        assert rel_err(v.cpu().float(), st["v_p"].float()) < tol


def test_adamw_fp32_master_weights_golden(K, golden):
    """bf16_hybrid (datautils/mixed_precision.py:24-28): fp32 master parameters and moments follow the reference's
    fp32 AdamW + clip fixture to 1e-4, and the bf16 compute copy is exactly the rounded master."""
    g = golden("optimizer.pt")[str(torch.float32)]
    p = g["p0"].cuda().clone()
    lo = p.to(torch.bfloat16)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    for step, ((gp, gq), st, nrm, lr) in enumerate(zip(g["grads"], g["states"], g["norms"], g["lrs"]), start=1):
        coef = torch.clamp(1.0 / (nrm.float().cuda() + 1e-6), max=1.0).reshape(1)
        K.adamw_master_(p, gp.cuda().float(), m, v, lo, step, lr, grad_scale=coef)
        assert rel_err(p.cpu(), st["p"].float()) < FP32_TOL
        assert rel_err(m.cpu(), st["m_p"].float()) < FP32_TOL
        assert rel_err(v.cpu(), st["v_p"].float()) < FP32_TOL
        assert torch.equal(lo, p.to(torch.bfloat16))


def test_b200adamw_master_weights_mode():
    """optim.B200AdamW(master_weights=True) on bf16 parameters: tiny updates that vanish in bf16 state still
    accumulate in the fp32 master, and the parameters the model reads are the rounded master."""
    from building_llm_from_scratch_b200.datautils.mixed_precision import bf16_hybrid_policy, bf16_policy, optimizer_kwargs
    from building_llm_from_scratch_b200.optim import B200AdamW
    assert optimizer_kwargs(bf16_hybrid_policy) == {"master_weights": True}
    assert optimizer_kwargs(bf16_policy) == {"master_weights": False}
    torch.manual_seed(0)
    w = torch.nn.Parameter(torch.ones(4096, device="cuda", dtype=torch.bfloat16))
    opt = B200AdamW([w], lr=1e-5, weight_decay=0.0, **optimizer_kwargs(bf16_hybrid_policy))
    grp = opt._groups[torch.bfloat16]
    assert grp["master"].dtype == grp["m"].dtype == torch.float32
    for _ in range(50):
        opt.zero_grad()
        w.main_grad.fill_(1.0)
        opt.step()
    torch.cuda.synchronize()
    # 50 steps of -lr * sign(g): the master moved by 5e-4 although one step (1e-5) is far below bf16's 2^-8 spacing at 1
    assert abs(float(grp["master"][0]) - (1.0 - 50e-5)) < 2e-6
    assert torch.equal(w.data, grp["master"][:4096].to(torch.bfloat16))
    sd = opt.state_dict()
    opt2 = B200AdamW([torch.nn.Parameter(torch.ones(4096, device="cuda", dtype=torch.bfloat16))], lr=1e-5,
                     weight_decay=0.0, master_weights=True)
    opt2.load_state_dict(sd)
    assert torch.equal(opt2._groups[torch.bfloat16]["master"], grp["master"])


# ------------------------------------------------------------------ attention
def ref_attention(q, k, v, B, T, h, hkv, hd):
    """oracle causal attention on [B*T, heads*hd] layouts, fp32."""
    from oracle import reference_ops as R
    qh = q.float().view(B, T, h, hd).transpose(1, 2)
    kh = k.float().view(B, T, hkv, hd).transpose(1, 2).repeat_interleave(h // hkv, dim=1)
    vh = v.float().view(B, T, hkv, hd).transpose(1, 2).repeat_interleave(h // hkv, dim=1)
    return R.causal_attention(qh, kh, vh).transpose(1, 2).reshape(B * T, h * hd)


@pytest.mark.parametrize("B,T,h,hkv,hd", [(2, 256, 4, 2, 64), (1, 100, 2, 2, 64), (2, 517, 8, 2, 64),
                                          (1, 2048, 4, 1, 64), (2, 192, 4, 2, 128), (1, 77, 2, 1, 128),
                                          (2, 517, 4, 4, 128), (1, 2048, 8, 2, 128)])
def test_attention_fwd_bwd(K, B, T, h, hkv, hd):
    qkv = rnd(B * T, (h + 2 * hkv) * hd, seed=1)
    q, k, v = qkv[:, :h * hd], qkv[:, h * hd:(h + hkv) * hd], qkv[:, (h + hkv) * hd:]
    o, lse = K.attention_fwd(q, k, v, B, T, h, hkv, hd)
    qr, kr, vr = (t.float().clone().requires_grad_(True) for t in (q, k, v))
    ref = ref_attention(qr, kr, vr, B, T, h, hkv, hd)
    assert rel_err(o.float(), ref) < BF16_TOL
    do = rnd(B * T, h * hd, seed=2)
    ref.backward(do.float())
    dq, dk, dv = K.attention_bwd(q, k, v, o, do, lse, B, T, h, hkv, hd)
    assert rel_err(dq.float(), qr.grad) < BF16_TOL
    assert rel_err(dk.float(), kr.grad) < BF16_TOL
    assert rel_err(dv.float(), vr.grad) < BF16_TOL


@pytest.mark.parametrize("hd", [64, 128])
def test_attention_causality_property(K, hd):
    """Size-independent property: output row t must not depend on tokens > t."""
    B, T, h, hkv = 1, 1024, 4, 2
    qkv = rnd(B * T, (h + 2 * hkv) * hd, seed=3)
    q, k, v = qkv[:, :h * hd], qkv[:, h * hd:(h + hkv) * hd], qkv[:, (h + hkv) * hd:]
    o1, _ = K.attention_fwd(q, k, v, B, T, h, hkv, hd)
    qkv2 = qkv.clone()
    qkv2[600:] = rnd(T - 600, qkv.shape[1], seed=4)
    o2, _ = K.attention_fwd(qkv2[:, :h * hd], qkv2[:, h * hd:(h + hkv) * hd], qkv2[:, (h + hkv) * hd:], B, T, h, hkv, hd)
    assert torch.equal(o1[:600], o2[:600])


# ------------------------------------------------------------------ dropout
def test_dropout_kernel_statistics_and_backward_mask(K):
    n, p, seed = 1 << 20, 0.1, 1234
    x = rnd(n, seed=1)
    y = K.dropout(x, p, seed)
    mask = K.dropout_mask(n, p, seed).bool()
    assert abs(float((~mask).float().mean()) - p) < 3e-3                 # drop rate
    assert torch.equal(y[~mask], torch.zeros_like(y[~mask]))
    assert rel_err(y[mask].float(), x[mask].float() / (1 - p)) < 5e-3
    assert torch.equal(K.dropout(x, p, seed), y)                           # same seed -> same mask (backward)
    assert not torch.equal(K.dropout_mask(n, p, seed + 1).bool(), mask)
    r = rnd(n, seed=2)
    assert rel_err(K.dropout(x, p, seed, residual=r).float(), y.float() + r.float()) < 5e-3


@pytest.mark.parametrize("hd", [64, 128])
def test_attention_dropout_matches_oracle_with_same_mask(K, hd):
    """softmax -> dropout -> @V (GPT2.py:42-46) with the kernel's own keep mask fed to the oracle."""
    B, T, h, hkv, p, seed = 2, 200, 4, 2, 0.2, 99
    qkv = rnd(B * T, (h + 2 * hkv) * hd, seed=1)
    q, k, v = qkv[:, :h * hd], qkv[:, h * hd:(h + hkv) * hd], qkv[:, (h + hkv) * hd:]
    keep = K.attention_dropout_mask(B, h, T, p, seed).float()
    assert abs(float(1 - keep.mean()) - p) < 5e-3
    qr, kr, vr = (t.float().clone().requires_grad_(True) for t in (q, k, v))
    qh = qr.view(B, T, h, hd).transpose(1, 2)
    kh = kr.view(B, T, hkv, hd).transpose(1, 2).repeat_interleave(h // hkv, dim=1)
    vh = vr.view(B, T, hkv, hd).transpose(1, 2).repeat_interleave(h // hkv, dim=1)
    sc = (qh @ kh.transpose(2, 3)) / hd ** 0.5
    sc = sc.masked_fill(torch.triu(torch.ones(T, T, device="cuda"), 1).bool(), -torch.inf)
    w = torch.softmax(sc, -1) * keep / (1 - p)
    ref = (w @ vh).transpose(1, 2).reshape(B * T, h * hd)
    o, lse = K.attention_fwd(q, k, v, B, T, h, hkv, hd, dropout_p=p, seed=seed)
    assert rel_err(o.float(), ref) < BF16_TOL
    do = rnd(B * T, h * hd, seed=2)
    ref.backward(do.float())
    dq, dk, dv = K.attention_bwd(q, k, v, o, do, lse, B, T, h, hkv, hd, dropout_p=p, seed=seed)
    assert rel_err(dq.float(), qr.grad) < BF16_TOL
    assert rel_err(dk.float(), kr.grad) < BF16_TOL
    assert rel_err(dv.float(), vr.grad) < BF16_TOL


@pytest.mark.parametrize("B,T,h,hkv,hd,tmax", [(2, 300, 8, 2, 64, 512), (1, 2048, 32, 8, 64, 2048), (3, 77, 4, 4, 128, 100),
                                               (2, 1000, 8, 1, 128, 1024)])
def test_attention_decode_against_oracle(K, B, T, h, hkv, hd, tmax):
    """One query token per sequence against a KV cache = the last row of the oracle's causal attention."""
    from oracle import reference_ops as R
    kc, vc = rnd(B, tmax, hkv * hd, seed=1), rnd(B, tmax, hkv * hd, seed=2)
    q = rnd(B, h * hd, seed=3)
    o = K.attention_decode(q, kc, vc, T, h, hkv, hd)
    qh = q.float().view(B, 1, h, hd).transpose(1, 2)                                  # [B, h, 1, hd]
    kh = kc[:, :T].float().view(B, T, hkv, hd).transpose(1, 2).repeat_interleave(h // hkv, dim=1)
    vh = vc[:, :T].float().view(B, T, hkv, hd).transpose(1, 2).repeat_interleave(h // hkv, dim=1)
    w = torch.softmax(qh @ kh.transpose(2, 3) / hd ** 0.5, dim=-1)
    ref = (w @ vh).transpose(1, 2).reshape(B, h * hd)
    assert rel_err(o.float(), ref) < BF16_TOL
