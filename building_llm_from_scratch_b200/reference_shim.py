"""Run the reference's own, unmodified scripts (main.py, build_components.py) on the b200 modules.

    python -m building_llm_from_scratch_b200.reference_shim /path/to/Building_LLM_from_scratch main.py --model llama3_2 ...

`install(ref_root)` puts the reference checkout at the front of `sys.path`, imports ITS modules and rebinds, inside
those modules' namespaces, the names of the decoder training step to this package's drop-ins:

    Models/Llama/Llama3.py   Llama3Model, TransformerBlock, GroupedQueryAttention, SharedBuffers, precompute_rope_params
    Models/Llama/Llama2.py   Llama2Model, TransformerBlock, MultiHeadAttention, precompute_rope_params
    Models/Llama/common_components.py   compute_rope, FeedForward, RMSNorm
    Models/GPT2/GPT2.py      GPTModel, TransformerBlock, MultiHeadAttention, FeedForward, GELU
    lora.py                  LoRALayer, LinearWithLoRA, replace_linear_with_lora
    train.py                 Trainer                      (main.py:10 `from train import Trainer`)
    generate.py              generate
    build_components.py      DDP -> B200DDP, build_optimizer -> B200AdamW (zero=args.use_zero_opt)   (:176, :239-258)

Everything else of the reference — argument parsing (args.py:46-93), config assembly, tokenizers, weight loading, data
loaders, logging, checkpoint cadence (main.py:58-157) — keeps running as written, which is the point: the flag face of
the reference selects the CUDA path without an edit to its files. Third-party modules the reference imports at module
level but that are absent from an offline image (tiktoken, huggingface_hub, matplotlib) are replaced by stubs that
raise on first use, so `--help`, model construction and training from pre-tokenised data work without them.
FSDP (`--use_fsdp`) is out of scope (broken upstream) and left untouched.
"""
from __future__ import annotations

import importlib
import os
import runpy
import sys
import types

_STUBS = ("tiktoken", "huggingface_hub", "matplotlib", "matplotlib.pyplot", "matplotlib.ticker")


def _stub_missing():
    for name in _STUBS:
        if name in sys.modules:
            continue
        try:
            importlib.import_module(name)
            continue
        except Exception:
            pass
        mod = types.ModuleType(name)

        def _missing(attr, _n=name):
            if attr.startswith("__"):
                raise AttributeError(attr)

            def fail(*a, **k):
                raise ModuleNotFoundError(f"{_n}.{attr} was called, but {_n} is not installed in this image "
                                          "(the reference shim only stubs it so the reference's modules import)")
            return fail
        mod.__getattr__ = lambda attr, _m=_missing: _m(attr)      # module-level __getattr__ (PEP 562)
        sys.modules[name] = mod


def install(ref_root: str) -> dict:
    """Rebind the hot-path names inside the reference's modules. Returns {reference module name: [names rebound]}."""
    ref_root = os.path.abspath(ref_root)
    if not os.path.isfile(os.path.join(ref_root, "Models", "Llama", "Llama3.py")):
        raise FileNotFoundError(f"{ref_root} is not a checkout of the reference (Models/Llama/Llama3.py not found)")
    _stub_missing()
    if ref_root not in sys.path:
        sys.path.insert(0, ref_root)

    from . import ddp as b_ddp, generate as b_generate, lora as b_lora, optim as b_optim, train as b_train
    from .Models.GPT2 import GPT2 as b_gpt2
    from .Models.Llama import Llama2 as b_l2, Llama3 as b_l3, common_components as b_cc

    table = {
        "Models.Llama.Llama3": (b_l3, ["Llama3Model", "TransformerBlock", "GroupedQueryAttention", "SharedBuffers",
                                       "precompute_rope_params"]),
        "Models.Llama.Llama2": (b_l2, ["Llama2Model", "TransformerBlock", "MultiHeadAttention", "precompute_rope_params"]),
        "Models.Llama.common_components": (b_cc, ["compute_rope", "FeedForward", "RMSNorm"]),
        "Models.GPT2.GPT2": (b_gpt2, ["GPTModel", "TransformerBlock", "MultiHeadAttention", "FeedForward", "GELU"]),
        "lora": (b_lora, ["LoRALayer", "LinearWithLoRA", "replace_linear_with_lora"]),
        "train": (b_train, ["Trainer"]),
        "generate": (b_generate, ["generate"]),
    }
    done = {}
    for ref_name, (ours, names) in table.items():
        ref_mod = importlib.import_module(ref_name)
        if not os.path.abspath(getattr(ref_mod, "__file__", "")).startswith(ref_root):
            raise RuntimeError(f"module {ref_name!r} resolved to {ref_mod.__file__}, not to the reference under {ref_root}")
        for n in names:
            if hasattr(ours, n):
                setattr(ref_mod, n, getattr(ours, n))
                done.setdefault(ref_name, []).append(n)

    bc = importlib.import_module("build_components")

    class DDP(b_ddp.B200DDP):
        """torch DDP's call shape (build_components.py:176 `DDP(model, device_ids=[rank])`)."""

        def __init__(self, module, device_ids=None, **kw):
            super().__init__(module)

    def build_optimizer(args, model):
        """build_components.py:239-258: AdamW(lr=args.lr, weight_decay=0.1), ZeRO-1 sharded under --use_zero_opt."""
        return b_optim.B200AdamW(model.parameters(), lr=args.lr, weight_decay=0.1, zero=bool(args.use_zero_opt))

    for n in ("GPTModel", "TransformerBlock"):
        setattr(bc, n, getattr(b_gpt2, n))
    bc.Llama2Model, bc.Llama3Model = b_l2.Llama2Model, b_l3.Llama3Model
    bc.DDP, bc.build_optimizer = DDP, build_optimizer
    done["build_components"] = ["GPTModel", "TransformerBlock", "Llama2Model", "Llama3Model", "DDP", "build_optimizer"]
    return done


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if len(argv) < 2:
        print(__doc__)
        print("usage: python -m building_llm_from_scratch_b200.reference_shim REF_ROOT SCRIPT [script args...]")
        return 2
    ref_root, script = argv[0], argv[1]
    install(ref_root)
    path = script if os.path.isabs(script) else os.path.join(os.path.abspath(ref_root), script)
    sys.argv = [path] + argv[2:]
    runpy.run_path(path, run_name="__main__")
    return 0


if __name__ == "__main__":
    sys.exit(main())
