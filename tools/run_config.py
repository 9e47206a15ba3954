"""Run a few training steps of one of BASELINE.json's other configs on 1 GPU (sanity + secondary
numbers; the headline bench is bench.py). Usage: python tools/run_config.py {gpt2-774m-lora|gpt2-1.5b|llama3-8b} [--layers N]"""
import argparse
import copy
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", choices=["gpt2-774m-lora", "gpt2-1.5b", "llama3-8b"])
    ap.add_argument("--layers", type=int, default=None)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--drop", type=float, default=0.0)
    ap.add_argument("--graph", action="store_true", help="replay the step as a CUDA graph (engine.TrainStep)")
    args = ap.parse_args()
    from building_llm_from_scratch_b200 import _lib
    from building_llm_from_scratch_b200.datautils.synthetic import gutenberg_batch
    from building_llm_from_scratch_b200.lora import replace_linear_with_lora
    from building_llm_from_scratch_b200.optim import B200AdamW
    from building_llm_from_scratch_b200.train import Trainer
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    torch.manual_seed(123)
    if args.config.startswith("gpt2"):
        from building_llm_from_scratch_b200.Models.GPT2.GPT2 import GPTModel
        from building_llm_from_scratch_b200.Models.GPT2.config import get_config_gpt2
        cfg = copy.deepcopy(get_config_gpt2("774M" if "774" in args.config else "1.5B"))
        cfg["drop_rate"] = args.drop
        if args.layers:
            cfg["n_layers"] = args.layers
        lora = "lora" in args.config
        with torch.device(dev):
            model = GPTModel(cfg, use_actv_ckpt=lora)
        if lora:
            for p in model.parameters():
                p.requires_grad = False
            with torch.device(dev):
                replace_linear_with_lora(model, rank=8, alpha=16, dtype=torch.float32)
        model = model.to(torch.bfloat16)            # harness cast
        T = 1024
    else:
        from building_llm_from_scratch_b200.Models.Llama.Llama3 import Llama3Model
        from building_llm_from_scratch_b200.Models.Llama.config import get_config_llama
        cfg = get_config_llama("8B", "llama3", target_context_length=4096)
        if args.layers:
            cfg["n_layers"] = args.layers
        with torch.device(dev):
            model = Llama3Model(cfg, use_actv_ckpt=True)
        T = 4096
    model.train()
    n_train = sum(p.numel() for p in model.parameters() if p.requires_grad)
    opt = B200AdamW(model.parameters(), lr=5e-4, weight_decay=0.1)
    tr = Trainer(model, opt, cfg, [], None, "/tmp", device=dev)
    tr.set_schedule(1000)
    gen = torch.Generator().manual_seed(7)
    x, y = gutenberg_batch(gen, args.batch, T, cfg["vocab_size"])
    losses = []
    n0 = _lib.launch_count()
    if args.graph:
        from building_llm_from_scratch_b200.engine import TrainStep
        step = TrainStep(tr, (args.batch, T), use_graph=True)
        step.capture(warmup=2)
        xd, yd = x.to(dev), y.to(dev)
        for s in range(args.steps + 2):
            if s == 2:
                torch.cuda.synchronize()
                t0 = time.perf_counter()
            losses.append(step.run(xd, yd).clone())
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / args.steps
        losses = [float(l) for l in losses]
    else:
        for s in range(args.steps + 2):
            if s == 2:
                torch.cuda.synchronize()
                t0 = time.perf_counter()
            losses.append(float(tr.train_batch(x, y)))
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / args.steps
    print(json.dumps({"config": args.config, "n_layers": cfg["n_layers"], "batch": args.batch, "seq": T,
                      "trainable_params": n_train, "mode": "cuda_graph" if args.graph else "eager", "ms_per_step": dt * 1e3, "tokens_per_s": args.batch * T / dt,
                      "losses": [round(l, 4) for l in losses], "b200llm_launches": _lib.launch_count() - n0,
                      "max_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}))


if __name__ == "__main__":
    main()
