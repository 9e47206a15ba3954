"""smoke(): one tiny training step of the hot path on cuda:0, checked against the oracle."""
from __future__ import annotations

import copy

import torch


def run_smoke() -> None:
    if not torch.cuda.is_available():
        raise RuntimeError("smoke() needs a GPU (the b200llm path has no CPU fallback)")
    from oracle import reference_ops as R  # the oracle is the checker here, never the product path

    from . import _lib
    from .Models.Llama import Llama3 as L
    from .lora import LinearWithLoRA, replace_linear_with_lora
    from .optim import B200AdamW
    from .train import Trainer

    torch.cuda.set_device(0)
    _lib.check(_lib.load().b200_init(0), "init")
    cfg = dict(vocab_size=1024, context_length=256, emb_dim=256, n_heads=4, n_layers=2, hidden_dim=512,
               n_kv_groups=2, rope_base=31250.0, dtype=torch.bfloat16,
               rope_freq=dict(factor=32.0, low_freq_factor=1.0, high_freq_factor=4.0, original_context_length=8192))
    torch.manual_seed(123)
    L.SharedBuffers._buffers.clear()
    model = L.Llama3Model(copy.deepcopy(cfg))
    for p in model.parameters():
        p.requires_grad = False
    replace_linear_with_lora(model, rank=8, alpha=16, dtype=torch.bfloat16)
    gen = torch.Generator().manual_seed(1)
    for m in model.modules():
        if isinstance(m, LinearWithLoRA):
            with torch.no_grad():
                m.lora.B.copy_((torch.randn(m.lora.B.shape, generator=gen) * 0.02).to(torch.bfloat16))
    model.cuda().train()
    x = torch.randint(0, cfg["vocab_size"], (2, 256), generator=gen)
    y = torch.roll(x, -1, 1)
    y[:, :40] = -100
    sd = {k: v.detach().clone() for k, v in model.state_dict().items() if not k.endswith("mask")}
    launches0 = _lib.launch_count()
    opt = B200AdamW(model.parameters(), lr=5e-4, weight_decay=0.1)
    tr = Trainer(model, opt, cfg, [], None, "/tmp", device="cuda")
    tr.set_schedule(20)
    loss = float(tr.train_batch(x, y))
    torch.cuda.synchronize()
    ref_logits = R.llama3_forward(x.cuda(), sd, cfg, 16 / 8)
    ref = float(R.cross_entropy(ref_logits, y.cuda()))
    n = _lib.launch_count() - launches0
    if not (abs(loss - ref) <= 2e-2 * abs(ref)):
        raise AssertionError(f"smoke: loss {loss} vs oracle {ref}")
    if n <= 0:
        raise AssertionError("smoke: no b200llm kernels were launched")
    print(f"smoke ok: loss {loss:.4f} (oracle {ref:.4f}), {n} b200llm kernel launches")
