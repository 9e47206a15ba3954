"""The reference arm: the UNMODIFIED upstream modules, run through their own public API.

`baseline/_ref/` is a verbatim copy of the read-only reference tree made by
`__graft_entry__.build()` (git-ignored: it never enters this repo's history, but it travels to the
GPU box with the snapshot). This file only *drives* it — model construction, LoRA activation,
Trainer.train_batch — exactly the way the reference's own main.py / build_components.py do:

    Llama3Model(cfg)                                   Models/Llama/Llama3.py:185-204
    freeze + replace_linear_with_lora(r, alpha)        build_components.py:117-135, lora.py:49-65
    torch.optim.AdamW(model.parameters(), lr, wd=0.1)  build_components.py:258
    Trainer(...).train_batch(x, y)                     train.py:94-126 (zero_grad, LR, forward,
                                                       F.cross_entropy, backward, clip 1.0, step)

Harness-level overrides only where the reference cannot run the named config as written
(every Llama context is forced to 1024, so the cfg dict is built here with
context_length = seq_len and rope_base rescaled by the reference's own rescale_theta; R10:
matplotlib is absent, utils.py imports it, so an empty stub module is registered first).

Used by bench.py's `--impl reference` / `cpu_baseline` legs and by tools/reference_on_gpu.py.
Nothing in the product package imports this.
"""
from __future__ import annotations

import contextlib
import copy
import io
import os
import sys
import time
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_DIR, "Models", "Llama", "Llama3.py"))


_mods = None


def modules():
    """Import the reference's modules from baseline/_ref (once). They use top-level names (`Models`,
    `lora`, `train`, `utils`, ...), so the directory goes to the FRONT of sys.path for the import and
    the product package (building_llm_from_scratch_b200.*) is never shadowed: it is always imported by
    its package-qualified name."""
    global _mods
    if _mods is not None:
        return _mods
    if not available():
        raise RuntimeError(f"{REF_DIR} is missing: run __graft_entry__.build() where /root/reference is mounted")
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.ticker"):
        if name not in sys.modules:
            mod = types.ModuleType(name)
            mod.MaxNLocator = object
            sys.modules[name] = mod
    sys.path.insert(0, REF_DIR)
    try:
        from Models.Llama import Llama3 as ref_llama3
        from Models.Llama import config as ref_llama_cfg
        from Models.Llama.common_components import rescale_theta
        from Models.GPT2 import GPT2 as ref_gpt2
        from Models.GPT2 import config as ref_gpt2_cfg
        import lora as ref_lora
        import train as ref_train
    finally:
        sys.path.remove(REF_DIR)
    _mods = types.SimpleNamespace(llama3=ref_llama3, llama_cfg=ref_llama_cfg, rescale_theta=rescale_theta,
                                  gpt2=ref_gpt2, gpt2_cfg=ref_gpt2_cfg, lora=ref_lora, train=ref_train)
    return _mods


def llama32_1b_cfg(seq_len: int, n_layers=None):
    """LLAMA32_CONFIG_1B with context_length = seq_len (harness override of SURVEY.md R4)."""
    m = modules()
    cfg = copy.deepcopy(m.llama_cfg.LLAMA32_CONFIG_1B)
    cfg["rope_base"] = m.rescale_theta(cfg["rope_base"], cfg["context_length"], seq_len)
    cfg["context_length"] = seq_len
    if n_layers is not None:
        cfg["n_layers"] = n_layers
    return cfg


def build_llama_lora_trainer(cfg, rank, alpha, device="cpu", lr=5e-4, weight_decay=0.1, total_steps=100_000):
    """The reference's model + LoRA + AdamW + Trainer for an instruction-finetune run, on `device`."""
    m = modules()
    torch.manual_seed(123)                                     # utils.set_seed(123), main.py:58
    m.llama3.SharedBuffers._buffers.clear()
    with contextlib.redirect_stdout(io.StringIO()):
        model = m.llama3.Llama3Model(cfg)
    for p in model.parameters():                               # build_components.py:129-130
        p.requires_grad = False
    m.lora.replace_linear_with_lora(model, rank=rank, alpha=alpha, dtype=cfg["dtype"])
    model.to(device)
    model.train()
    opt = torch.optim.AdamW(model.parameters(), lr=lr, weight_decay=weight_decay)
    tr = m.train.Trainer(model, opt, cfg, [], None, "/tmp", device=device)
    tr.peak_lr = lr                                            # train.py:154-156
    tr.total_training_steps = total_steps
    tr.lr_increment = (tr.peak_lr - tr.initial_lr) / tr.warmup_steps
    return tr


def train_batch(tr, x, y):
    """One reference Trainer.train_batch. The reference logs and swallows exceptions inside the step
    (SURVEY.md R9), so success is checked by the step counter and by a finite loss."""
    before = tr.tokens_seen
    with contextlib.redirect_stdout(io.StringIO()):
        tr.train_batch(x, y)
    if tr.tokens_seen != before + x.numel():
        raise RuntimeError("reference Trainer.train_batch did not complete (its exception handler swallowed an error)")


def time_steps(tr, batches, warmup, steps, sync=None):
    """Wall-clock seconds of each of `steps` train_batch calls after `warmup` untimed ones."""
    out = []
    for i in range(warmup + steps):
        x, y = batches[i % len(batches)]
        if sync:
            sync()
        t0 = time.perf_counter()
        train_batch(tr, x, y)
        if sync:
            sync()
        if i >= warmup:
            out.append(time.perf_counter() - t0)
    return out
