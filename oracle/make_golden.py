"""Generate tests/golden/*.pt from the LIVE reference (run in the build container only).

    python oracle/make_golden.py [/root/reference]

The reference has no golden vectors of its own, so the known-answer fixtures
are produced by importing its modules unmodified from the read-only mount, feeding them seeded
inputs and recording inputs + outputs + gradients. Fixtures carry their inputs and weights, so
replaying them does not depend on RNG streams. /root/reference does not exist on the GPU box;
nothing at test time reads it.
"""
from __future__ import annotations

import contextlib
import copy
import io
import os
import sys
import types

import torch

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

# matplotlib is absent in this image and utils.py imports it: stub it.
for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.ticker"):
    if name not in sys.modules:
        mod = types.ModuleType(name)
        mod.MaxNLocator = object
        sys.modules[name] = mod
sys.path.insert(0, REF)

from Models.GPT2.GPT2 import GPTModel  # noqa: E402
from Models.GPT2.GPT2 import FeedForward as GPTFeedForward  # noqa: E402
from Models.GPT2.GPT2 import MultiHeadAttention  # noqa: E402
from Models.Llama.common_components import FeedForward, RMSNorm, compute_rope  # noqa: E402
from Models.Llama.Llama3 import (GroupedQueryAttention, Llama3Model, SharedBuffers,  # noqa: E402
                                 precompute_rope_params)
from Models.Llama import Llama2 as ref_llama2  # noqa: E402
from datautils.dataloader_instruction_finetune import custom_collate_fn  # noqa: E402
from datautils.dataset import DatasetPT  # noqa: E402
from lora import LinearWithLoRA, replace_linear_with_lora  # noqa: E402
import train as ref_train  # noqa: E402


def save(name, obj):
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, name)
    torch.save(obj, path)
    print(f"wrote {path} ({os.path.getsize(path) / 1024:.1f} KiB)")


def grads_of(module):
    return {k: p.grad.clone() for k, p in module.named_parameters() if p.grad is not None}


def perturb_lora_B(model, gen, std=0.02):
    # B starts at zero => dA == 0 on step 0; perturb so dA is exercised.
    for m in model.modules():
        if isinstance(m, LinearWithLoRA):
            with torch.no_grad():
                m.lora.B.copy_((torch.randn(m.lora.B.shape, generator=gen) * std).to(m.lora.B.dtype))


def data_pipeline():
    cases = [
        dict(batch=[(2, [10, 11, 12, 13, 14]), (3, [20, 21, 22]), (1, [30])], kw={}),
        dict(batch=[(2, [10, 50256, 12, 50256, 14]), (1, [20, 21])], kw={}),
        dict(batch=[(2, [10, 11, 12, 13, 14, 15, 16]), (3, [20, 21, 22])], kw=dict(allowed_max_length=4)),
        dict(batch=[(0, [10, 11, 12]), (1, [20, 21, 22])], kw={}),
        dict(batch=[(1, [10, 11, 12])], kw={}),
        dict(batch=[(4, [7, 8, 9]), (2, [50256, 50256, 5])], kw=dict(allowed_max_length=8)),
    ]
    gen = torch.Generator().manual_seed(7)
    for _ in range(6):
        n = int(torch.randint(1, 6, (1,), generator=gen))
        batch = []
        for _ in range(n):
            ln = int(torch.randint(1, 40, (1,), generator=gen))
            toks = torch.randint(0, 50300, (ln,), generator=gen).tolist()
            batch.append((int(torch.randint(0, ln + 2, (1,), generator=gen)), toks))
        kw = {} if torch.rand(1, generator=gen) < 0.5 else dict(allowed_max_length=int(torch.randint(2, 30, (1,), generator=gen)))
        cases.append(dict(batch=batch, kw=kw))
    for c in cases:
        x, y = custom_collate_fn(copy.deepcopy(c["batch"]), **c["kw"])
        c["inputs"], c["targets"] = x, y

    class Tok:  # DatasetPT only calls tokenizer.encode(txt, allowed_special=...)
        def __init__(self, ids):
            self.ids = ids

        def encode(self, txt, allowed_special=None):
            return list(self.ids)

    shifts = []
    for n, L, s in [(10, 4, 4), (9, 4, 4), (8, 4, 4), (5, 4, 4), (4, 4, 4), (10, 4, 2), (33, 8, 8), (17, 16, 16)]:
        ids = list(range(100, 100 + n))
        ds = DatasetPT("x", Tok(ids), L, s)
        shifts.append(dict(tokens=ids, max_length=L, stride=s,
                           inputs=[t.tolist() for t in ds.input_ids],
                           targets=[t.tolist() for t in ds.target_ids]))
    save("data_pipeline.pt", dict(collate=cases, shift=shifts))


def schedule_and_ce():
    # LR schedule through the reference Trainer (train.py:94-107) with a dummy model
    lin = torch.nn.Linear(4, 4)
    opt = torch.optim.AdamW(lin.parameters(), lr=5e-4, weight_decay=0.1)
    tr = ref_train.Trainer(lin, opt, {}, [], None, "/tmp", warmup_steps=10, initial_lr=1e-5, min_lr=1e-6)
    tr.peak_lr, tr.total_training_steps = 5e-4, 40
    tr.lr_increment = (tr.peak_lr - tr.initial_lr) / tr.warmup_steps
    tr.calc_loss_batch = lambda a, b: (lin(a) ** 2).mean()
    for _ in range(40):
        tr.train_batch(torch.ones(2, 4), None)
    gen = torch.Generator().manual_seed(3)
    logits = torch.randn(12, 37, generator=gen) * 3
    targets = torch.randint(0, 37, (12,), generator=gen)
    targets[[1, 4, 5, 9]] = -100
    lg = logits.clone().requires_grad_(True)
    loss = torch.nn.functional.cross_entropy(lg, targets)
    loss.backward()
    lb = logits.bfloat16().clone().requires_grad_(True)
    loss_b = torch.nn.functional.cross_entropy(lb, targets)
    loss_b.backward()
    all_ign = torch.nn.functional.cross_entropy(logits, torch.full((12,), -100))
    save("schedule_ce.pt", dict(lrs=tr.track_lrs, peak_lr=5e-4, total=40, warmup=10, initial_lr=1e-5, min_lr=1e-6,
                                logits=logits, targets=targets, loss=loss.detach(), dlogits=lg.grad,
                                loss_bf16=loss_b.detach(), dlogits_bf16=lb.grad, loss_all_ignored=all_ign))


def optimizer():
    out = {}
    for dtype in (torch.float32, torch.bfloat16):
        gen = torch.Generator().manual_seed(11)
        p = torch.nn.Parameter((torch.randn(257, generator=gen) * 0.5).to(dtype))
        q = torch.nn.Parameter((torch.randn(64, 3, generator=gen) * 0.1).to(dtype))
        p0, q0 = p.detach().clone(), q.detach().clone()
        opt = torch.optim.AdamW([p, q], lr=5e-4, weight_decay=0.1)
        grads, states, norms = [], [], []
        for step in range(5):
            gp = (torch.randn(257, generator=gen) * (3.0 if step % 2 == 0 else 0.01)).to(dtype)
            gq = (torch.randn(64, 3, generator=gen) * (3.0 if step % 2 == 0 else 0.01)).to(dtype)
            p.grad, q.grad = gp.clone(), gq.clone()
            for g in opt.param_groups:
                g["lr"] = 5e-4 * (step + 1) / 5
            nrm = torch.nn.utils.clip_grad_norm_([p, q], max_norm=1.0)
            opt.step()
            grads.append((gp, gq))
            norms.append(nrm.detach().clone())
            states.append(dict(p=p.detach().clone(), q=q.detach().clone(), clipped_gp=p.grad.clone(),
                               m_p=opt.state[p]["exp_avg"].clone(), v_p=opt.state[p]["exp_avg_sq"].clone()))
        out[str(dtype)] = dict(p0=p0, q0=q0, grads=grads, states=states, norms=norms,
                               lrs=[5e-4 * (s + 1) / 5 for s in range(5)])
    save("optimizer.pt", out)


def llama_modules():
    out = {}
    gen = torch.Generator().manual_seed(21)
    for dtype in (torch.float32, torch.bfloat16):
        tag = str(dtype)
        # RMSNorm
        x = (torch.randn(2, 5, 64, generator=gen) * 2).to(dtype).requires_grad_(True)
        norm = RMSNorm(64)
        with torch.no_grad():
            norm.weight.copy_(1 + 0.1 * torch.randn(64, generator=gen))
        y = norm(x)
        gy = torch.randn(y.shape, generator=gen).to(dtype)
        y.backward(gy)
        out[f"rmsnorm/{tag}"] = dict(x=x.detach(), w=norm.weight.detach().clone(), y=y.detach(), gy=gy,
                                     dx=x.grad.clone(), dw=norm.weight.grad.clone())
        # RoPE with Llama-3.2 frequency smoothing
        fc = dict(factor=32.0, low_freq_factor=1.0, high_freq_factor=4.0, original_context_length=8192)
        cos, sin = precompute_rope_params(16, theta_base=31250.0, context_length=32, freq_config=fc)
        cos, sin = cos.to(dtype), sin.to(dtype)
        xr = torch.randn(2, 4, 12, 16, generator=gen).to(dtype).requires_grad_(True)
        yr = compute_rope(xr, cos, sin)
        gr = torch.randn(yr.shape, generator=gen).to(dtype)
        yr.backward(gr)
        out[f"rope/{tag}"] = dict(x=xr.detach(), cos=cos, sin=sin, y=yr.detach(), gy=gr, dx=xr.grad.clone(),
                                  theta=31250.0, freq_config=fc, ctx=32, head_dim=16)
        # SwiGLU feed-forward
        cfg = dict(emb_dim=64, hidden_dim=160, dtype=dtype)
        ff = FeedForward(cfg)
        xf = torch.randn(2, 7, 64, generator=gen).to(dtype).requires_grad_(True)
        yf = ff(xf)
        gf = torch.randn(yf.shape, generator=gen).to(dtype)
        yf.backward(gf)
        out[f"ffn/{tag}"] = dict(x=xf.detach(), sd={k: v.clone() for k, v in ff.state_dict().items()}, y=yf.detach(),
                                 gy=gf, dx=xf.grad.clone(), grads=grads_of(ff))
        # GQA attention (hd 64 so the CUDA path can replay it; 4 heads, 2 kv groups)
        SharedBuffers._buffers.clear()
        att = GroupedQueryAttention(256, 256, 64, 4, 2, rope_base=31250.0, rope_config=fc, dtype=dtype)
        xa = torch.randn(2, 40, 256, generator=gen).to(dtype).requires_grad_(True)
        ya = att(xa)
        ga = torch.randn(ya.shape, generator=gen).to(dtype)
        ya.backward(ga)
        sd = {k: v.clone() for k, v in att.state_dict().items() if k != "mask"}
        out[f"gqa/{tag}"] = dict(x=xa.detach(), sd=sd, y=ya.detach(), gy=ga, dx=xa.grad.clone(), grads=grads_of(att),
                                 n_heads=4, n_kv_groups=2)
    save("llama_modules.pt", out)


TINY_LLAMA = dict(vocab_size=512, context_length=64, emb_dim=128, n_heads=2, n_layers=2, hidden_dim=256,
                  n_kv_groups=1, rope_base=31250.0, dtype=torch.bfloat16,
                  rope_freq=dict(factor=32.0, low_freq_factor=1.0, high_freq_factor=4.0, original_context_length=8192))


def alpaca_batch(gen, B, T, vocab):
    """Alpaca-shaped synthetic batch through the reference collate."""
    items = []
    for i in range(B):
        ln = T + 5 if i == 0 else int(torch.randint(T // 3, T, (1,), generator=gen))
        toks = torch.randint(0, vocab, (ln,), generator=gen).tolist()
        items.append((int(torch.randint(2, T // 4, (1,), generator=gen)), toks))
    # tiny vocabularies cannot hold the default pad id 50256; the masking logic is unchanged
    pad = 50256 if vocab > 50256 else vocab - 1
    return custom_collate_fn(items, pad_token_id=pad, allowed_max_length=T)


def llama_model():
    gen = torch.Generator().manual_seed(31)
    torch.manual_seed(123)
    SharedBuffers._buffers.clear()
    cfg = copy.deepcopy(TINY_LLAMA)
    model = Llama3Model(cfg)
    for p in model.parameters():
        p.requires_grad = False
    replace_linear_with_lora(model, rank=8, alpha=16, dtype=torch.bfloat16)
    perturb_lora_B(model, gen)
    x, y = alpaca_batch(gen, 2, 64, cfg["vocab_size"])
    sd0 = {k: v.clone() for k, v in model.state_dict().items() if not k.endswith(".mask")}
    logits = model(x)
    loss = torch.nn.functional.cross_entropy(logits.flatten(0, 1), y.flatten())
    loss.backward()
    out = dict(cfg=cfg, rank=8, alpha=16, sd=sd0, x=x, y=y, logits=logits.detach(), loss=loss.detach(),
               grads=grads_of(model))
    # 200 reference Trainer steps (the north star's loss-curve horizon) cycling over 8 batches
    opt = torch.optim.AdamW(model.parameters(), lr=5e-4, weight_decay=0.1)
    tr = ref_train.Trainer(model, opt, cfg, [], None, "/tmp", warmup_steps=10, initial_lr=1e-5, min_lr=1e-6)
    tr.peak_lr, tr.total_training_steps = 5e-4, 200
    tr.lr_increment = (tr.peak_lr - tr.initial_lr) / tr.warmup_steps
    batches = [alpaca_batch(gen, 2, 64, cfg["vocab_size"]) for _ in range(8)]
    losses = []
    model.train()
    for s in range(200):
        xb, yb = batches[s % 8]
        with torch.no_grad():
            losses.append(float(torch.nn.functional.cross_entropy(model(xb).flatten(0, 1), yb.flatten())))
        tr.train_batch(xb, yb)
    out["curve"] = dict(batches=batches, losses=losses, lrs=tr.track_lrs, warmup=10, total=200, peak_lr=5e-4)
    save("llama_model.pt", out)


TINY_GPT = dict(vocab_size=257, context_length=64, emb_dim=128, n_heads=2, n_layers=2, drop_rate=0.0, qkv_bias=True)


def gpt_modules_and_model():
    out = {}
    gen = torch.Generator().manual_seed(41)
    torch.manual_seed(123)
    mha = MultiHeadAttention(128, 128, 64, 0.0, 2, qkv_bias=True)
    xa = torch.randn(2, 48, 128, generator=gen).requires_grad_(True)
    ya = mha(xa)
    ga = torch.randn(ya.shape, generator=gen)
    ya.backward(ga)
    out["mha/fp32"] = dict(x=xa.detach(), sd={k: v.clone() for k, v in mha.state_dict().items() if k != "mask"},
                           y=ya.detach(), gy=ga, dx=xa.grad.clone(), grads=grads_of(mha), n_heads=2)
    ff = GPTFeedForward(dict(emb_dim=64))
    xf = torch.randn(3, 5, 64, generator=gen).requires_grad_(True)
    yf = ff(xf)
    gf = torch.randn(yf.shape, generator=gen)
    yf.backward(gf)
    out["ffn/fp32"] = dict(x=xf.detach(), sd={k: v.clone() for k, v in ff.state_dict().items()}, y=yf.detach(),
                           gy=gf, dx=xf.grad.clone(), grads=grads_of(ff))
    for lora in (False, True):
        torch.manual_seed(123)
        model = GPTModel(copy.deepcopy(TINY_GPT))
        if lora:
            for p in model.parameters():
                p.requires_grad = False
            replace_linear_with_lora(model, rank=8, alpha=16, dtype=torch.float32)
            perturb_lora_B(model, gen)
        x = torch.randint(0, 257, (2, 64), generator=gen)
        y = torch.roll(x, -1, dims=1)
        y[0, 10:20] = -100
        sd0 = {k: v.clone() for k, v in model.state_dict().items() if not k.endswith(".mask")}
        with contextlib.redirect_stdout(io.StringIO()):  # GPT2.py:123 prints on every forward
            logits = model(x)
        loss = torch.nn.functional.cross_entropy(logits.flatten(0, 1), y.flatten())
        loss.backward()
        out["model_lora" if lora else "model"] = dict(cfg=copy.deepcopy(TINY_GPT), rank=8, alpha=16, sd=sd0, x=x, y=y,
                                                      logits=logits.detach(), loss=loss.detach(), grads=grads_of(model))
    save("gpt2.pt", out)


def llama2_modules():
    """Llama-2 modules (Models/Llama/Llama2.py:61-190: MHA with fp32 RoPE tables, out-of-place masking, block, model)
    in bf16, with every gradient. The driver path for Llama-2 is broken upstream, the modules are not."""
    out = {}
    gen = torch.Generator().manual_seed(51)
    torch.manual_seed(123)
    mha = ref_llama2.MultiHeadAttention(256, 256, 64, 4, dtype=torch.bfloat16)
    xa = torch.randn(2, 40, 256, generator=gen).to(torch.bfloat16).requires_grad_(True)
    ya = mha(xa)
    ga = torch.randn(ya.shape, generator=gen).to(torch.bfloat16)
    ya.backward(ga)
    out["mha"] = dict(x=xa.detach(), sd={k: v.clone() for k, v in mha.state_dict().items() if k != "mask"},
                      y=ya.detach(), gy=ga, dx=xa.grad.clone(), grads=grads_of(mha), n_heads=4, ctx=64)
    cfg = dict(vocab_size=512, context_length=64, emb_dim=256, n_heads=4, n_layers=2, hidden_dim=512, dtype=torch.bfloat16)
    blk = ref_llama2.TransformerBlock(cfg)
    with torch.no_grad():
        blk.norm1.weight.copy_(1 + 0.1 * torch.randn(256, generator=gen))
        blk.norm2.weight.copy_(1 + 0.1 * torch.randn(256, generator=gen))
    xb = torch.randn(2, 48, 256, generator=gen).to(torch.bfloat16).requires_grad_(True)
    yb = blk(xb)
    gb = torch.randn(yb.shape, generator=gen).to(torch.bfloat16)
    yb.backward(gb)
    out["block"] = dict(cfg=cfg, x=xb.detach(), sd={k: v.clone() for k, v in blk.state_dict().items() if not k.endswith("mask")},
                        y=yb.detach(), gy=gb, dx=xb.grad.clone(), grads=grads_of(blk))
    cfg = dict(vocab_size=256, context_length=64, emb_dim=128, n_heads=2, n_layers=2, hidden_dim=256, dtype=torch.bfloat16)
    model = ref_llama2.Llama2Model(cfg)                  # full fine-tuning: every parameter gets a gradient
    x = torch.randint(0, 256, (2, 64), generator=gen)
    y = torch.roll(x, -1, dims=1)
    y[1, :20] = -100
    sd0 = {k: v.clone() for k, v in model.state_dict().items() if not k.endswith(".mask")}
    logits = model(x)
    loss = torch.nn.functional.cross_entropy(logits.flatten(0, 1), y.flatten())
    loss.backward()
    out["model"] = dict(cfg=cfg, sd=sd0, x=x, y=y, logits=logits.detach(), loss=loss.detach(), grads=grads_of(model))
    save("llama2.pt", out)


if __name__ == "__main__":
    torch.set_num_threads(8)
    if len(sys.argv) > 2 and sys.argv[2] == "llama2":       # add the Llama-2 fixtures without regenerating the rest
        llama2_modules()
        raise SystemExit(0)
    data_pipeline()
    schedule_and_ce()
    optimizer()
    llama_modules()
    llama_model()
    gpt_modules_and_model()
    llama2_modules()
