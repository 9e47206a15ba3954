"""CPU: the reference's unmodified build_components.py / train.py assemble the b200 drop-ins once the shim is installed
(`building_llm_from_scratch_b200.reference_shim`). Runs in a subprocess — the shim rebinds names inside the reference's
top-level modules (`Models`, `lora`, `train`), which other tests import unpatched. Skipped where the reference checkout
is not mounted."""
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = next((p for p in ("/root/reference", os.path.join(ROOT, "baseline", "_ref"))
            if os.path.isfile(os.path.join(p, "Models", "Llama", "Llama3.py"))), None)

pytestmark = pytest.mark.skipif(REF is None, reason="reference checkout not available")

SCRIPT = textwrap.dedent("""
    import argparse, sys, torch
    sys.path.insert(0, {root!r})
    from building_llm_from_scratch_b200 import reference_shim
    done = reference_shim.install({ref!r})
    import build_components as bc, train, lora
    import Models.Llama.Llama3 as ref_l3, Models.GPT2.GPT2 as ref_gpt2
    from building_llm_from_scratch_b200.Models.GPT2.GPT2 import GPTModel
    from building_llm_from_scratch_b200.Models.Llama.Llama3 import Llama3Model
    from building_llm_from_scratch_b200.lora import LinearWithLoRA
    from building_llm_from_scratch_b200.optim import B200AdamW
    from building_llm_from_scratch_b200.train import Trainer
    assert ref_l3.__file__.startswith({ref!r}) and ref_l3.Llama3Model is Llama3Model
    assert hasattr(ref_l3, "Llama3Tokenizer")                  # the rest of the reference module is untouched
    assert train.Trainer is Trainer and lora.LinearWithLoRA is LinearWithLoRA
    # the reference's own assembly code (build_components.py:50-81, 190-236, 239-258), GPT-2 debug config + LoRA
    args = argparse.Namespace(model="GPT2", num_params="124M", data_type="bf16", load_weights=False, debug=True,
                              use_actv_ckpt=False, use_lora=True, lora_rank=8, lora_alpha=16, use_fsdp=False,
                              use_zero_opt=False, run_type="single_gpu", lr=5e-4)
    cfg = bc.build_config(args)
    model = bc.build_model(cfg, 0, torch.device("cpu"), args)
    assert isinstance(model, GPTModel)
    wrapped = [m for m in model.modules() if isinstance(m, LinearWithLoRA)]
    assert len(wrapped) == 2 * 6 + 1, len(wrapped)             # 4 attention + 2 MLP linears per block, + out_head
    trainable = [n for n, p in model.named_parameters() if p.requires_grad]
    assert trainable and all(".lora." in n for n in trainable)
    assert bc.DDP.__mro__[1].__name__ == "B200DDP"
    print("SHIM-OK", sorted(done))
""")


def test_reference_assembly_code_builds_the_b200_modules():
    r = subprocess.run([sys.executable, "-c", SCRIPT.format(root=ROOT, ref=REF)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "SHIM-OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_reference_cli_help_runs_through_the_shim():
    """args.py's parser (args.py:46-93) is reached through main.py started by the shim's CLI."""
    r = subprocess.run([sys.executable, "-m", "building_llm_from_scratch_b200.reference_shim", REF, "main.py", "--help"],
                       capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0 and "--lora_rank" in r.stdout and "--use_zero_opt" in r.stdout, r.stdout[-1500:] + r.stderr[-3000:]
