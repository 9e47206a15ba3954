"""GPU: the drop-in modules (same class names / state_dict keys as the reference) against fixtures
recorded from the live reference modules, and against the oracle on the same weights.
bf16 tolerance 2e-2 relative per op (north star); end-to-end gradients through a whole bf16 model
are compared at 5e-2; the 200-step loss curve must track the reference's within 1 % (smoothed over
8 steps, plus 0.02 abs for the bf16 loss read-out)."""
import copy

import pytest
import torch

from conftest import close, rel_err

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]
BF16_TOL = 2e-2


def to_cuda(sd, dtype=None):
    out = {}
    for k, v in sd.items():
        v = v.cuda()
        if dtype is not None and v.is_floating_point() and not k.endswith("norm1.weight") and not k.endswith(
                "norm2.weight") and not k.endswith("final_norm.weight"):
            v = v.to(dtype)
        out[k] = v
    return out


def test_llama_modules_against_reference_fixtures(golden):
    from building_llm_from_scratch_b200.Models.Llama import Llama3 as L
    from building_llm_from_scratch_b200.Models.Llama.common_components import FeedForward, RMSNorm, compute_rope
    g = golden("llama_modules.pt")
    # GroupedQueryAttention (bf16)
    f = g["gqa/torch.bfloat16"]
    L.SharedBuffers._buffers.clear()
    fc = dict(factor=32.0, low_freq_factor=1.0, high_freq_factor=4.0, original_context_length=8192)
    att = L.GroupedQueryAttention(256, 256, 64, 4, 2, rope_base=31250.0, rope_config=fc, dtype=torch.bfloat16).cuda()
    att.load_state_dict(f["sd"])
    assert torch.equal(att.cos.cpu(), f["sd"]["cos"]) and torch.equal(att.sin.cpu(), f["sd"]["sin"])
    x = f["x"].cuda().requires_grad_(True)
    y = att(x)
    y.backward(f["gy"].cuda())
    assert rel_err(y.cpu().float(), f["y"].float()) < BF16_TOL
    assert rel_err(x.grad.cpu().float(), f["dx"].float()) < BF16_TOL
    for k, v in f["grads"].items():
        assert rel_err(dict(att.named_parameters())[k].grad.cpu().float(), v.float()) < BF16_TOL, k
    # FeedForward
    f = g["ffn/torch.bfloat16"]
    ff = FeedForward(dict(emb_dim=64, hidden_dim=160, dtype=torch.bfloat16)).cuda()
    ff.load_state_dict(f["sd"])
    x = f["x"].cuda().requires_grad_(True)
    y = ff(x)
    y.backward(f["gy"].cuda())
    assert rel_err(y.cpu().float(), f["y"].float()) < BF16_TOL
    assert rel_err(x.grad.cpu().float(), f["dx"].float()) < BF16_TOL
    for k, v in f["grads"].items():
        assert rel_err(dict(ff.named_parameters())[k].grad.cpu().float(), v.float()) < BF16_TOL, k
    # RMSNorm module + compute_rope function
    f = g["rmsnorm/torch.bfloat16"]
    n = RMSNorm(64).cuda()
    n.load_state_dict({"weight": f["w"]})
    x = f["x"].cuda().requires_grad_(True)
    y = n(x)
    y.backward(f["gy"].cuda())
    assert rel_err(y.cpu().float(), f["y"].float()) < 5e-3
    assert rel_err(x.grad.cpu().float(), f["dx"].float()) < BF16_TOL
    assert rel_err(n.weight.grad.cpu(), f["dw"].float()) < BF16_TOL
    f = g["rope/torch.bfloat16"]
    assert torch.equal(compute_rope(f["x"].cuda(), f["cos"].cuda(), f["sin"].cuda()).cpu(), f["y"])


def build_llama(g, fused_opt):
    from building_llm_from_scratch_b200.Models.Llama import Llama3 as L
    from building_llm_from_scratch_b200.lora import replace_linear_with_lora
    from building_llm_from_scratch_b200.optim import B200AdamW
    L.SharedBuffers._buffers.clear()
    model = L.Llama3Model(copy.deepcopy(g["cfg"]))
    for p in model.parameters():
        p.requires_grad = False
    replace_linear_with_lora(model, rank=g["rank"], alpha=g["alpha"], dtype=torch.bfloat16)
    model.load_state_dict(g["sd"])
    model.cuda().train()
    opt = B200AdamW(model.parameters(), lr=5e-4, weight_decay=0.1) if fused_opt else None
    return model, opt


@pytest.mark.parametrize("fused_opt", [False, True])
def test_llama3_lora_model_against_reference_fixture(golden, fused_opt):
    g = golden("llama_model.pt")
    model, opt = build_llama(g, fused_opt)
    x, y = g["x"].cuda(), g["y"].cuda()
    with torch.no_grad():
        model.eval()
        logits = model(x)
        model.train()
    assert logits.shape == g["logits"].shape and logits.dtype == torch.bfloat16
    assert rel_err(logits.cpu().float(), g["logits"].float()) < BF16_TOL
    if opt is not None:
        opt.zero_grad()
    loss = model.forward_loss(x, y)
    loss.backward()
    from building_llm_from_scratch_b200 import ops as _ops
    _ops.flush_weight_grads()      # weight-gradient GEMMs into main_grad are queued and grouped
    assert rel_err(loss.cpu(), g["loss"].float()) < BF16_TOL
    named = dict(model.named_parameters())
    assert set(g["grads"]) == {k for k, p in named.items() if p.requires_grad}
    for k, v in g["grads"].items():
        got = named[k].main_grad if fused_opt else named[k].grad
        assert got is not None, k
        assert close(got.cpu().float(), v.float(), 5e-2, atol=1e-6), (k, rel_err(got.cpu().float(), v.float()))
    # the materialised-logits path (evaluate_model / generate) gives the same loss
    from building_llm_from_scratch_b200 import ops
    loss2 = ops.cross_entropy(model(x), y)[0]
    assert rel_err(loss2.cpu(), loss.detach().cpu()) < 1e-3


def test_llama3_loss_curve_tracks_reference(golden):
    from building_llm_from_scratch_b200.train import Trainer
    g = golden("llama_model.pt")
    c = g["curve"]
    model, opt = build_llama(g, True)
    tr = Trainer(model, opt, g["cfg"], [], None, "/tmp", warmup_steps=c["warmup"], initial_lr=1e-5, min_lr=1e-6,
                 device="cuda")
    tr.set_schedule(c["total"], peak_lr=c["peak_lr"])
    losses = []
    for s in range(c["total"]):
        xb, yb = c["batches"][s % len(c["batches"])]
        with torch.no_grad():
            losses.append(float(model.forward_loss(xb.cuda(), yb.cuda())))
        tr.train_batch(xb, yb)
    assert tr.track_lrs == pytest.approx(c["lrs"], rel=1e-12)
    ref = c["losses"]
    assert len(ref) == 200 and ref[-1] < ref[0] - 0.5, "fixture should actually learn"
    # per-step: loose (bf16 weights / bf16 Adam state drift differently step to step);
    # smoothed over one pass through the 8 batches: the north-star 1 % bound
    for s, (a, b) in enumerate(zip(losses, ref)):
        assert abs(a - b) <= 0.03 * abs(b) + 0.05, (s, a, b)
    w = 8
    for s in range(0, 200 - w + 1, w):
        a = sum(losses[s:s + w]) / w
        b = sum(ref[s:s + w]) / w
        assert abs(a - b) <= 0.01 * abs(b) + 0.02, (s, a, b)


@pytest.mark.parametrize("which", ["model", "model_lora"])
def test_gpt2_model_against_reference_fixture(golden, which):
    from building_llm_from_scratch_b200.Models.GPT2.GPT2 import GPTModel
    from building_llm_from_scratch_b200.lora import replace_linear_with_lora
    g = golden("gpt2.pt")[which]
    model = GPTModel(copy.deepcopy(g["cfg"]))
    if which == "model_lora":
        for p in model.parameters():
            p.requires_grad = False
        replace_linear_with_lora(model, rank=g["rank"], alpha=g["alpha"], dtype=torch.float32)
    model.load_state_dict(g["sd"])
    model = model.cuda().to(torch.bfloat16).train()     # harness cast, SURVEY.md R1/R2
    x, y = g["x"].cuda(), g["y"].cuda()
    logits = model(x)
    assert rel_err(logits.detach().cpu().float(), g["logits"]) < BF16_TOL
    loss = model.forward_loss(x, y)
    loss.backward()
    assert rel_err(loss.cpu(), g["loss"]) < BF16_TOL
    named = dict(model.named_parameters())
    for k, v in g["grads"].items():
        assert named[k].grad is not None, k
        assert close(named[k].grad.cpu().float(), v, 6e-2, atol=2e-5), (k, rel_err(named[k].grad.cpu().float(), v))


def test_activation_checkpointing_matches_plain(golden):
    g = golden("llama_model.pt")
    model, _ = build_llama(g, False)
    x, y = g["x"].cuda(), g["y"].cuda()
    model.forward_loss(x, y).backward()
    plain = {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}
    model.zero_grad()
    model.use_actv_ckpt = True
    model.forward_loss(x, y).backward()
    for k, p in model.named_parameters():
        if p.grad is not None:
            assert close(p.grad.float(), plain[k].float(), 1e-3, atol=1e-6), k


@pytest.mark.parametrize("lora", [True, False])
def test_activation_checkpointing_keeps_the_main_grad_path(golden, lora):
    """Under torch.utils.checkpoint (use_reentrant=False) backward receives DETACHED copies of the saved weights — no
    main_grad, no gradient-ready subscription. The ops keep the Parameters they were called with, so a checkpointed
    block still accumulates straight into the optimizer's flat fp32 buffer: nothing comes back as .grad, the flat
    gradients equal the un-checkpointed run's, and every parameter is reported ready (what B200DDP's overlap needs)."""
    from building_llm_from_scratch_b200 import ops
    from building_llm_from_scratch_b200.optim import B200AdamW
    g = golden("llama_model.pt")
    model, _ = build_llama(g, False)
    if not lora:                                       # full fine-tuning of the same modules
        for p in model.parameters():
            p.requires_grad = p.dtype == torch.bfloat16
    opt = B200AdamW(model.parameters(), lr=5e-4, weight_decay=0.1)
    x, y = g["x"].cuda(), g["y"].cuda()

    def run(ckpt):
        model.use_actv_ckpt = ckpt
        opt.zero_grad()
        seen = []
        ops.register_grad_ready(opt._trainable, lambda ps: seen.extend(id(q) for q in ps))
        model.forward_loss(x, y).backward()
        ops.flush_weight_grads()
        ops.unregister_grad_ready(opt._trainable)
        stray = [k for k, p in model.named_parameters() if p.grad is not None]
        return [gr.clone() for gr in opt.flat_grads], set(seen), stray

    plain, seen_plain, stray_plain = run(False)
    ck, seen_ck, stray_ck = run(True)
    assert stray_plain == [] and stray_ck == [], (stray_plain, stray_ck)
    want = {id(p) for p in opt._trainable}
    assert seen_plain == want and seen_ck == want
    for a, b in zip(plain, ck):
        assert float(a.abs().max()) > 0 and close(b, a, 1e-3, atol=1e-6)


def test_ops_reject_cpu_tensors():
    from building_llm_from_scratch_b200 import ops
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.rmsnorm(torch.zeros(4, 64, dtype=torch.bfloat16), torch.ones(64), 1e-5)


def test_row_compacted_vocab_projection_is_exact(golden):
    """Projecting only the non-ignored target rows gives the same loss and the same gradients."""
    from building_llm_from_scratch_b200 import ops
    g = golden("llama_model.pt")
    x, y = g["x"].cuda(), g["y"].cuda()
    n_valid = int((g["y"] != -100).sum())
    assert 0 < n_valid < y.numel()
    results = []
    for cap in (0, (n_valid + 15) // 16 * 16):
        model, _ = build_llama(g, False)
        loss = model.forward_loss(x, y, capacity=cap)
        loss.backward()
        results.append((loss.detach().clone(), {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}))
    (l0, g0), (l1, g1) = results
    assert rel_err(l1, l0) < 1e-5
    assert set(g0) == set(g1)
    for k in g0:
        assert close(g1[k].float(), g0[k].float(), 2e-3, atol=1e-6), (k, rel_err(g1[k].float(), g0[k].float()))
    # an undersized capacity must not silently drop rows: the loss turns into NaN
    model, _ = build_llama(g, False)
    bad = model.forward_loss(x, y, capacity=16)
    assert torch.isnan(bad)


def test_llama2_modules_against_reference_fixtures(golden):
    """Llama-2 MultiHeadAttention, TransformerBlock and Llama2Model (Models/Llama/Llama2.py:61-190: fp32 RoPE tables,
    kv groups == heads) against outputs AND gradients recorded from the reference's own modules."""
    from building_llm_from_scratch_b200.Models.Llama import Llama2 as L2
    g = golden("llama2.pt")
    f = g["mha"]
    att = L2.MultiHeadAttention(256, 256, f["ctx"], f["n_heads"], dtype=torch.bfloat16).cuda()
    att.load_state_dict(f["sd"])
    assert att.cos.dtype == torch.float32 and torch.equal(att.cos.cpu(), f["sd"]["cos"])
    x = f["x"].cuda().requires_grad_(True)
    y = att(x)
    y.backward(f["gy"].cuda())
    assert rel_err(y.cpu().float(), f["y"].float()) < BF16_TOL
    assert rel_err(x.grad.cpu().float(), f["dx"].float()) < BF16_TOL
    for k, v in f["grads"].items():
        assert rel_err(dict(att.named_parameters())[k].grad.cpu().float(), v.float()) < BF16_TOL, k
    f = g["block"]
    blk = L2.TransformerBlock(f["cfg"]).cuda()
    blk.load_state_dict(f["sd"])
    x = f["x"].cuda().requires_grad_(True)
    y = blk(x)
    y.backward(f["gy"].cuda())
    assert rel_err(y.cpu().float(), f["y"].float()) < BF16_TOL
    assert rel_err(x.grad.cpu().float(), f["dx"].float()) < BF16_TOL
    for k, v in f["grads"].items():
        assert rel_err(dict(blk.named_parameters())[k].grad.cpu().float(), v.float()) < BF16_TOL, k
    f = g["model"]
    model = L2.Llama2Model(copy.deepcopy(f["cfg"])).cuda()
    model.load_state_dict(f["sd"])
    model.train()
    logits = model(f["x"].cuda())
    assert rel_err(logits.cpu().float(), f["logits"].float()) < BF16_TOL
    loss = model.forward_loss(f["x"].cuda(), f["y"].cuda())
    assert abs(float(loss) - float(f["loss"])) <= BF16_TOL * abs(float(f["loss"]))
    loss.backward()
    params = dict(model.named_parameters())
    assert set(f["grads"]) == set(params)
    for k, v in f["grads"].items():
        assert close(params[k].grad.cpu().float(), v.float(), 5e-2, atol=1e-5), (k, rel_err(params[k].grad.cpu().float(), v.float()))


def test_gpt2_trains_with_dropout(golden):
    """GPT-2 at its default drop_rate 0.1 (embedding, residual and attention-weight dropout through the
    b200llm kernels): a few optimizer steps run and the loss goes down; eval mode is deterministic."""
    from building_llm_from_scratch_b200.Models.GPT2.GPT2 import GPTModel
    from building_llm_from_scratch_b200.optim import B200AdamW
    from building_llm_from_scratch_b200.train import Trainer
    g = golden("gpt2.pt")["model"]
    cfg = dict(copy.deepcopy(g["cfg"]), drop_rate=0.1)
    torch.manual_seed(0)
    model = GPTModel(cfg)
    model.load_state_dict(g["sd"])
    model = model.cuda().to(torch.bfloat16).train()
    opt = B200AdamW(model.parameters(), lr=1e-3, weight_decay=0.1)
    tr = Trainer(model, opt, cfg, [], None, "/tmp", warmup_steps=2, device="cuda")
    tr.set_schedule(40)
    x, y = g["x"], g["y"]
    losses = [float(tr.train_batch(x, y)) for _ in range(12)]
    assert all(l == l for l in losses) and losses[-1] < losses[0] - 0.3
    model.eval()
    with torch.no_grad():
        a, b2 = model(x.cuda()), model(x.cuda())
    assert torch.equal(a, b2)


def test_trainstep_graph_replay_matches_eager_trainer(golden):
    """engine.TrainStep (CUDA-graph replay, row-compacted vocab projection, device-side LR / bias
    corrections) follows the same loss trajectory as Trainer.train_batch run eagerly."""
    from building_llm_from_scratch_b200.engine import TrainStep
    from building_llm_from_scratch_b200.train import Trainer
    g = golden("llama_model.pt")
    c = g["curve"]

    def make():
        model, opt = build_llama(g, True)
        tr = Trainer(model, opt, g["cfg"], [], None, "/tmp", warmup_steps=c["warmup"], initial_lr=1e-5, min_lr=1e-6,
                     device="cuda")
        tr.set_schedule(c["total"], peak_lr=c["peak_lr"])
        return model, tr

    _, tr_a = make()
    eager = [float(tr_a.train_batch(*c["batches"][s % len(c["batches"])])) for s in range(12)]
    _, tr_b = make()
    step = TrainStep(tr_b, tuple(c["batches"][0][0].shape), use_graph=True)
    step.bucket_rows = 16
    step.capture(warmup=2)
    graph = []
    for s in range(12):
        xb, yb = c["batches"][s % len(c["batches"])]
        graph.append(float(step.run(xb.pin_memory(), yb.pin_memory())))
    assert step.bucket_graphs, "the row-compacted graphs should have been used"
    assert tr_b.track_lrs == tr_a.track_lrs
    for s, (a, b) in enumerate(zip(graph, eager)):
        assert abs(a - b) <= 0.01 * abs(b) + 0.02, (s, a, b)
    assert graph[-1] < graph[0] - 0.05


def test_generate_and_evaluate_paths(golden, tmp_path):
    """generate.py drop-in (greedy + top-k sampling) and Trainer.evaluate_model / save_checkpoint."""
    from building_llm_from_scratch_b200.generate import generate
    from building_llm_from_scratch_b200.train import Trainer
    from oracle import reference_ops as R
    g = golden("llama_model.pt")
    model, _ = build_llama(g, False)
    x = g["x"][:, :20].cuda()
    out = generate(model, x, max_new_tokens=6, context_size=64, temperature=0.0)
    assert out.shape == (2, 26) and torch.equal(out[:, :20], x)
    # every greedy token is the arg-max of the oracle's last-position logits up to bf16 near-ties
    sd = {k: v.cuda() for k, v in g["sd"].items()}
    for t in range(20, 26):
        ref = R.llama3_forward(out[:, :t], sd, g["cfg"], g["alpha"] / g["rank"])[:, -1, :].float()
        chosen = ref.gather(1, out[:, t:t + 1])
        assert torch.all(chosen >= ref.max(dim=-1, keepdim=True).values - 0.08)
    assert rel_err(model.forward_last(x).float(), model(x)[:, -1, :].float()) < 1e-2
    torch.manual_seed(0)
    s1 = generate(model, x, max_new_tokens=5, context_size=64, temperature=1.0, top_k=5)
    assert s1.shape == (2, 25)
    tr = Trainer(model, torch.optim.AdamW([p for p in model.parameters() if p.requires_grad], lr=1e-4), g["cfg"], [], None,
                 tmp_path, device="cuda")
    loader = [g["curve"]["batches"][0], g["curve"]["batches"][1]]
    tl, vl = tr.evaluate_model(loader, loader, eval_iter=2)
    assert abs(tl - vl) < 1e-6 and 5.0 < tl < 8.0 and model.training
    tr.save_checkpoint("model_pg_test.pth")
    saved = torch.load(tmp_path / "model_pg_test.pth", weights_only=False)
    assert "trf_blocks.0.att.mask" in saved and "out_head.lora.A" in saved


def test_packed_projections_match_separate_ones(golden, monkeypatch):
    """q|k|v and fc1|fc2 as one packed projection (single RoPE launch, one backward dX GEMM against the
    row-concatenated frozen weights) against the one-GEMM-per-linear path on the same LoRA model."""
    g = golden("llama_model.pt")
    x, y = g["x"].cuda(), g["y"].cuda()
    results = []
    for packed in ("1", "0"):
        monkeypatch.setenv("B200_PACKED", packed)
        model, opt = build_llama(g, fused_opt=True)
        opt.zero_grad()
        h = model.tok_emb(x).detach().requires_grad_(True)        # exposes dX of the first block
        out = model.trf_blocks[0](h)
        out.float().square().mean().backward()
        loss = model.forward_loss(x, y)
        loss.backward()
        opt.flush_pending_grads()
        results.append((out.detach().float(), h.grad.float(), float(loss.detach()),
                        {k: p.main_grad.clone() for k, p in model.named_parameters() if p.requires_grad}))
    (o1, d1, l1, g1), (o0, d0, l0, g0) = results
    assert rel_err(o1, o0) < 1e-2 and rel_err(d1, d0) < BF16_TOL
    assert abs(l1 - l0) < 1e-2 * abs(l0)
    for k in g0:
        assert close(g1[k], g0[k], 5e-2, atol=1e-6), (k, rel_err(g1[k], g0[k]))
    # a frozen weight written to (load_state_dict) invalidates the cached row-concatenation
    from building_llm_from_scratch_b200 import ops
    att = model.trf_blocks[0].att
    ws = [att.W_query.linear.weight, att.W_key.linear.weight, att.W_value.linear.weight]
    c1 = ops._packed_weight(ws)
    with torch.no_grad():
        ws[1].mul_(2.0)
    c2 = ops._packed_weight(ws)
    assert torch.equal(c2, torch.cat(ws, 0)) and not torch.equal(c1, c2)


def test_kv_cache_generation_matches_full_window(golden):
    """Prefill + one-token decode steps against the KV cache give the logits of re-running the whole window
    (what the reference's generate.py does), for the Llama-3 LoRA model and for GPT-2."""
    from building_llm_from_scratch_b200.generate import generate
    g = golden("llama_model.pt")
    model, _ = build_llama(g, False)
    model.eval()
    x = g["x"][:, :24].cuda()
    cache = model.new_kv_cache(x.shape[0], 40)
    seq = x[:, :17]
    lg = model.forward_cached(seq, cache)
    with torch.no_grad():
        assert rel_err(lg.float(), model.forward_last(seq).float()) < BF16_TOL
        for t in range(17, 24):
            seq = x[:, :t + 1]
            lg = model.forward_cached(seq[:, -1:], cache)
            full = model.forward_last(seq).float()
            assert rel_err(lg.float(), full) < BF16_TOL, t
    assert cache.pos == 24
    a = generate(model, x[:, :12], max_new_tokens=8, context_size=64, use_cache=True)
    b = generate(model, x[:, :12], max_new_tokens=8, context_size=64, use_cache=False)
    assert a.shape == b.shape and float((a == b).float().mean()) > 0.9      # bf16 near-ties may pick another token
    # GPT-2 (learned positions, biases, LayerNorm)
    from building_llm_from_scratch_b200.Models.GPT2.GPT2 import GPTModel
    gg = golden("gpt2.pt")["model"]
    m2 = GPTModel(copy.deepcopy(gg["cfg"]))
    m2.load_state_dict(gg["sd"])
    m2 = m2.cuda().to(torch.bfloat16).eval()
    xs = gg["x"].cuda()[:, :16]
    cache = m2.new_kv_cache(xs.shape[0], 32)
    lg = m2.forward_cached(xs[:, :9], cache)
    with torch.no_grad():
        assert rel_err(lg.float(), m2.forward_last(xs[:, :9]).float()) < BF16_TOL
        for t in range(9, 16):
            lg = m2.forward_cached(xs[:, t:t + 1], cache)
            assert rel_err(lg.float(), m2.forward_last(xs[:, :t + 1]).float()) < BF16_TOL, t


def test_training_state_resume(golden, tmp_path):
    """save_training_state / load_training_state: a run resumed from the snapshot continues with the same
    step count, learning rates and (up to fp32 summation order in the weight-gradient atomics) parameters."""
    from building_llm_from_scratch_b200.train import Trainer
    g = golden("llama_model.pt")
    batches = g["curve"]["batches"]

    def fresh():
        model, opt = build_llama(g, True)
        tr = Trainer(model, opt, g["cfg"], [], None, tmp_path, warmup_steps=2, device="cuda")
        tr.set_schedule(50, peak_lr=1e-3)
        return model, opt, tr
    model, opt, tr = fresh()
    for i in range(3):
        tr.train_batch(*batches[i])
    tr.save_training_state("state.pt").join()
    for i in range(3, 5):
        tr.train_batch(*batches[i])
    want = {k: p.detach().float().clone() for k, p in model.named_parameters() if p.requires_grad}
    lrs = list(tr.track_lrs)
    model2, opt2, tr2 = fresh()
    tr2.load_training_state("state.pt")
    assert tr2.global_step == 2 and opt2._step == 3 and tr2.track_lrs == lrs[:3]
    for i in range(3, 5):
        tr2.train_batch(*batches[i])
    assert tr2.track_lrs == lrs
    for k, p in model2.named_parameters():
        if p.requires_grad:
            assert rel_err(p.detach().float(), want[k]) < 1e-2, k
    # a torch.optim.AdamW state is refused rather than misread
    with pytest.raises(ValueError):
        opt2.load_state_dict({"state": {}, "param_groups": []})


# ======================================================================================
# Real dimensions: one block + head of the bench workload (and of the other BASELINE configs) against the
# oracle evaluated on the GPU from the same state_dict, with the routing that only these sizes trigger.
# ======================================================================================
def _oracle_sd(model):
    """state_dict of a drop-in model as leaf tensors the oracle can differentiate (LoRA A / B and any trainable
    weight require grad; the mask buffers are not needed)."""
    sd = {}
    params = dict(model.named_parameters())
    for k, v in model.state_dict().items():
        if k.endswith("mask"):
            continue
        t = v.detach().clone()
        if k in params and params[k].requires_grad:
            t.requires_grad_(True)
        sd[k] = t
    return sd


def test_llama32_1b_block_and_head_at_full_dimensions():
    """d 2048, 32 heads / 8 kv groups, ffn 8192, V 128 256, B4 x T2048 (the bench's micro-batch), LoRA r=64 alpha=32 with perturbed B, ~60 % of
    the targets ignored (so the row-compacted vocab projection runs): block output, input gradient, loss and all
    16 LoRA gradients against the oracle; and the launches must have taken the packed q|k|v, SwiGLU-epilogue,
    K-split tail and stream-K paths this shape is meant to trigger."""
    from building_llm_from_scratch_b200 import _lib, ops
    from building_llm_from_scratch_b200.Models.Llama import Llama3 as L
    from building_llm_from_scratch_b200.Models.Llama.config import get_config_llama
    from building_llm_from_scratch_b200.lora import LinearWithLoRA, head_operands, replace_linear_with_lora
    from oracle import reference_ops as R
    B, T, r, alpha = 4, 2048, 64, 32
    cfg = get_config_llama("1B", "llama3_2", target_context_length=T)
    cfg["n_layers"] = 1
    torch.manual_seed(123)
    L.SharedBuffers._buffers.clear()
    with torch.device("cuda"):
        model = L.Llama3Model(copy.deepcopy(cfg))
    for p in model.parameters():
        p.requires_grad = False
    with torch.device("cuda"):
        replace_linear_with_lora(model, rank=r, alpha=alpha, dtype=torch.bfloat16)
    gen = torch.Generator(device="cuda").manual_seed(5)
    with torch.no_grad():
        for m in model.modules():
            if isinstance(m, LinearWithLoRA):
                m.lora.B.copy_((torch.randn(m.lora.B.shape, generator=gen, device="cuda") * 0.02).to(torch.bfloat16))
                m.linear.weight.copy_((torch.randn(m.linear.weight.shape, generator=gen, device="cuda") * 0.02).to(torch.bfloat16))
    model.train()
    sd = _oracle_sd(model)
    x0 = (torch.randn(B, T, cfg["emb_dim"], generator=gen, device="cuda") * 0.5).to(torch.bfloat16)
    tg = torch.randint(0, cfg["vocab_size"], (B, T), generator=gen, device="cuda")
    tg[torch.rand(B, T, generator=gen, device="cuda") < 0.6] = -100
    n_valid = int((tg != -100).sum())
    capacity = (n_valid + 255) // 256 * 256
    assert 0 < capacity < B * T

    before = _lib.gemm_path_counts()
    xa = x0.clone().requires_grad_(True)
    y = model.trf_blocks(xa)
    hs = model.final_norm(y).to(torch.bfloat16)
    w, A, Bm, s = head_operands(model.out_head)
    loss = ops.linear_cross_entropy(hs, w, A, Bm, s, tg, capacity, True)[0]
    loss.backward()
    ops.flush_weight_grads()
    torch.cuda.synchronize()
    took = {k: v - before[k] for k, v in _lib.gemm_path_counts().items()}
    assert took["packed_segments"] >= 1, took          # q|k|v as one GEMM with per-segment LoRA blocks
    assert took["swiglu_fwd"] == 1 and took["swiglu_bwd"] == 1, took
    assert took["pair_tail_split"] >= 1, took          # fc3: 256 tiles on 74 CTA pairs with an 8192-deep K
    assert took["streamk"] >= 1, took                  # u = dlogits B^T over the 128 256-wide K range

    xb = x0.clone().requires_grad_(True)
    scal = alpha / r
    yo = R.llama3_block(xb, sd, "trf_blocks.0", cfg, scal)
    ho = R.rmsnorm(yo, sd["final_norm.weight"]).to(torch.bfloat16)
    lo = R.cross_entropy(R.lora_linear(ho, sd, "out_head", scal), tg)
    lo.backward()
    assert rel_err(y.float(), yo.float()) < BF16_TOL
    assert abs(float(loss) - float(lo)) <= BF16_TOL * abs(float(lo))
    assert rel_err(xa.grad.float(), xb.grad.float()) < BF16_TOL
    n = 0
    for k, p in model.named_parameters():
        if p.requires_grad:
            assert p.grad is not None and sd[k].grad is not None, k
            assert rel_err(p.grad.float(), sd[k].grad.float()) < 5e-2, k
            n += 1
    assert n == 16


def test_gpt2_1p5b_block_at_full_dimensions():
    """One GPT2-1.5B block as BASELINE configs[3] trains it (d 1600, 25 heads, biases on out_proj / MLP, full
    fine-tuning, B4 x T1024, dropout off for parity): output, input gradient and every weight / bias gradient
    against the oracle. 1600 = 6.25 tiles of 256: every GEMM has an N tail, and 25 heads is not a power of two."""
    from building_llm_from_scratch_b200.Models.GPT2 import GPT2 as G
    from building_llm_from_scratch_b200.Models.GPT2.config import get_config_gpt2
    from oracle import reference_ops as R
    cfg = copy.deepcopy(get_config_gpt2("1.5B"))
    cfg["drop_rate"] = 0.0
    B, T = 4, 1024
    torch.manual_seed(1)
    with torch.device("cuda"):
        blk = G.TransformerBlock(cfg)
    blk = blk.to(torch.bfloat16).train()
    gen = torch.Generator(device="cuda").manual_seed(9)
    with torch.no_grad():
        for k, p in blk.named_parameters():
            if k.endswith("bias"):
                p.copy_((torch.randn(p.shape, generator=gen, device="cuda") * 0.1).to(torch.bfloat16))
    sd = {"blocks.0." + k: v for k, v in _oracle_sd(blk).items()}
    x0 = (torch.randn(B, T, cfg["emb_dim"], generator=gen, device="cuda") * 0.5).to(torch.bfloat16)
    gy = (torch.randn(B, T, cfg["emb_dim"], generator=gen, device="cuda") * 0.1).to(torch.bfloat16)
    xa = x0.clone().requires_grad_(True)
    y = blk(xa)
    y.backward(gy)
    xb = x0.clone().requires_grad_(True)
    yo = R.gpt2_block(xb, sd, "blocks.0", cfg)
    yo.backward(gy)
    assert rel_err(y.float(), yo.float()) < BF16_TOL
    assert rel_err(xa.grad.float(), xb.grad.float()) < BF16_TOL
    n = 0
    for k, p in blk.named_parameters():
        ref = sd["blocks.0." + k].grad
        assert p.grad is not None and ref is not None, k
        assert close(p.grad.float(), ref.float(), 5e-2, atol=2e-4), (k, rel_err(p.grad.float(), ref.float()))
        n += 1
    assert n == 13       # 4 attention weights + out_proj bias, 2 MLP weights + biases, 2 x (LayerNorm weight, bias)


def test_attention_hd128_t4096_against_oracle():
    """BASELINE configs[4]'s attention shape (Llama-3-8B: 32 heads / 8 kv groups of width 128, T 4096, B 1),
    forward and backward against the oracle's materialised softmax(QK^T)V."""
    from building_llm_from_scratch_b200 import kernels as K
    from oracle import reference_ops as R
    B, T, h, hkv, hd = 1, 4096, 32, 8, 128
    gen = torch.Generator(device="cuda").manual_seed(3)
    qkv = (torch.randn(B * T, (h + 2 * hkv) * hd, generator=gen, device="cuda") * 0.5).to(torch.bfloat16)
    q, k, v = qkv[:, :h * hd], qkv[:, h * hd:(h + hkv) * hd], qkv[:, (h + hkv) * hd:]
    do = (torch.randn(B * T, h * hd, generator=gen, device="cuda") * 0.1).to(torch.bfloat16)
    o, lse = K.attention_fwd(q, k, v, B, T, h, hkv, hd)
    dq, dk, dv = K.attention_bwd(q, k, v, o, do, lse, B, T, h, hkv, hd)
    qr, kr, vr = (t.float().clone().requires_grad_(True) for t in (q, k, v))
    qh = qr.view(B, T, h, hd).transpose(1, 2)
    kh = kr.view(B, T, hkv, hd).transpose(1, 2).repeat_interleave(h // hkv, dim=1)
    vh = vr.view(B, T, hkv, hd).transpose(1, 2).repeat_interleave(h // hkv, dim=1)
    ref = R.causal_attention(qh, kh, vh).transpose(1, 2).reshape(B * T, h * hd)
    ref.backward(do.float())
    assert rel_err(o.float(), ref) < BF16_TOL
    assert rel_err(dq.float(), qr.grad) < BF16_TOL
    assert rel_err(dk.float(), kr.grad) < BF16_TOL
    assert rel_err(dv.float(), vr.grad) < BF16_TOL


def _mk(shape, gen, scale=1.0, grad=True):
    t = (torch.randn(*shape, generator=gen, device="cuda") * scale).to(torch.bfloat16)
    return t.requires_grad_(True) if grad else t


@pytest.mark.parametrize("lora", [False, True])
def test_gpt2_fused_block_ops_match_their_compositions(lora):
    """ops.linear_dropout_residual and ops.gelu_mlp (GELU / dropout + residual / GELU-backward in GEMM epilogues) against
    the same computation from separate ops (linear, gelu, dropout with the same seeds) at GPT2-774M's MLP shape:
    outputs and every gradient."""
    from building_llm_from_scratch_b200 import ops
    gen = torch.Generator(device="cuda").manual_seed(17)
    M, d, F, r, s, p = 1024, 1280, 5120, 8, 2.0, 0.1
    x0, res0 = _mk((M, d), gen, 0.5), _mk((M, d), gen, 0.5)
    ws = dict(w1=_mk((F, d), gen, 0.03), b1=_mk((F,), gen, 0.1), w2=_mk((d, F), gen, 0.03), b2=_mk((d,), gen, 0.1))
    if lora:
        ws.update(A1=_mk((d, r), gen, 0.1), B1=_mk((r, F), gen, 0.1), A2=_mk((F, r), gen, 0.1), B2=_mk((r, d), gen, 0.1))
    else:
        ws.update(A1=None, B1=None, A2=None, B2=None)
    gy = _mk((M, d), gen, 0.1, grad=False)

    def run(fused):
        leaves = {k: (v.detach().clone().requires_grad_(True) if v is not None else None) for k, v in ws.items()}
        x, res = x0.detach().clone().requires_grad_(True), res0.detach().clone().requires_grad_(True)
        L = leaves
        if fused:
            y = ops.gelu_mlp(x, L["w1"], L["b1"], L["A1"], L["B1"], L["w2"], L["b2"], L["A2"], L["B2"], s, res, p, 99)[0]
            y2 = ops.linear_dropout_residual(y, L["w2"][:, :d].contiguous().detach(), None, None, None, 0.0, res, p, 7)[0]
        else:
            z = ops.lora_linear(x, L["w1"], L["b1"], L["A1"], L["B1"], s, None)[0]
            h = ops.gelu(z)
            o = ops.lora_linear(h, L["w2"], L["b2"], L["A2"], L["B2"], s, None)[0]
            y = ops.dropout(o, res, p, 99)
            o2 = ops.lora_linear(y, L["w2"][:, :d].contiguous().detach(), None, None, None, 0.0, None)[0]
            y2 = ops.dropout(o2, res, p, 7)
        y2.backward(gy)
        grads = {k: v.grad for k, v in L.items() if v is not None}
        grads.update(x=x.grad, res=res.grad)
        return y2.detach(), grads
    yf, gf = run(True)
    yu, gu = run(False)
    assert rel_err(yf.float(), yu.float()) < 4e-3
    for k in gu:
        assert gf[k] is not None, k
        assert close(gf[k].float(), gu[k].float(), BF16_TOL, atol=1e-5), (k, rel_err(gf[k].float(), gu[k].float()))


def test_gpt2_dropout_masks_change_between_graph_replays(golden):
    """drop_rate 0.1 under CUDA-graph replay: the site seeds are frozen into the captured kernel arguments, the
    per-step counter the kernels read from device memory is not — the same batch replayed twice sees two different
    masks (different losses), and the counter advances once per step."""
    from building_llm_from_scratch_b200 import ops
    from building_llm_from_scratch_b200.Models.GPT2.GPT2 import GPTModel
    from building_llm_from_scratch_b200.engine import TrainStep
    from building_llm_from_scratch_b200.optim import B200AdamW
    from building_llm_from_scratch_b200.train import Trainer
    cfg = dict(vocab_size=512, context_length=128, emb_dim=256, n_heads=4, n_layers=2, drop_rate=0.1, qkv_bias=False)
    torch.manual_seed(0)
    model = GPTModel(cfg).cuda().to(torch.bfloat16).train()
    opt = B200AdamW(model.parameters(), lr=0.0, weight_decay=0.0)          # lr 0: only the masks can change the loss
    tr = Trainer(model, opt, cfg, [], None, "/tmp", device="cuda")
    tr.set_schedule(100, peak_lr=0.0)
    tr.initial_lr = tr.min_lr = 0.0
    tr.lr_increment = 0.0
    step = TrainStep(tr, (4, 128), use_graph=True)
    step.compact_head = False
    step.capture(warmup=1)
    x = torch.randint(0, 512, (4, 128), device="cuda")
    y = torch.roll(x, -1, 1)
    c0 = int(ops.dropout_step_tensor("cuda"))
    losses = [float(step.run(x, y).clone()) for _ in range(4)]
    assert int(ops.dropout_step_tensor("cuda")) == c0 + 4
    assert len(set(losses)) == 4, losses
    assert max(losses) - min(losses) < 0.5
