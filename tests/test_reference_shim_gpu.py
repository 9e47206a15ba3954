"""GPU: the reference's own assembly code (build_components.py) + Trainer call sequence (main.py:120-137), started
through the shim, trains on the CUDA kernels. Subprocess for the same reason as the CPU test; uses baseline/_ref (the
copy that travels to the GPU box)."""
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = next((p for p in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference")
            if os.path.isfile(os.path.join(p, "Models", "Llama", "Llama3.py"))), None)

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600), pytest.mark.skipif(REF is None, reason="reference copy not available")]

SCRIPT = textwrap.dedent("""
    import argparse, sys, torch
    sys.path.insert(0, {root!r})
    from building_llm_from_scratch_b200 import reference_shim, _lib
    reference_shim.install({ref!r})
    import build_components as bc
    from train import Trainer                                   # main.py:10
    args = argparse.Namespace(model="GPT2", num_params="124M", data_type="bf16", load_weights=False, debug=False,
                              use_actv_ckpt=False, use_lora=True, lora_rank=8, lora_alpha=16, use_fsdp=False,
                              use_zero_opt=False, run_type="single_gpu", lr=5e-4)
    dev = torch.device("cuda")
    cfg = bc.build_config(args)
    cfg["drop_rate"] = 0.0
    model = bc.build_model(cfg, 0, dev, args).to(torch.bfloat16)
    opt = bc.build_optimizer(args, model)
    tr = Trainer(config=cfg, data_files=[], loaderObj=None, device=dev, model=model, optimizer=opt, save_dir="/tmp",
                 rank=0, warmup_steps=2, initial_lr=1e-5, min_lr=1e-6, eval_freq=10, save_ckpt_freq=100,
                 print_sample_iter=10, eval_iter=5)
    tr.set_schedule(20)
    g = torch.Generator().manual_seed(0)
    x = torch.randint(0, cfg["vocab_size"], (2, 256), generator=g)
    y = torch.roll(x, -1, 1)
    n0 = _lib.launch_count()
    losses = [float(tr.train_batch(x.to(dev), y.to(dev))) for _ in range(6)]
    assert _lib.launch_count() - n0 > 100, "the b200 kernels did not run"
    assert all(l == l for l in losses) and losses[-1] < losses[0], losses
    print("SHIM-GPU-OK", [round(l, 3) for l in losses])
""")


def test_reference_assembly_trains_on_the_b200_kernels():
    r = subprocess.run([sys.executable, "-c", SCRIPT.format(root=ROOT, ref=REF)], capture_output=True, text=True, timeout=500)
    assert r.returncode == 0 and "SHIM-GPU-OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
