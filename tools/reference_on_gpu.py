"""The practical bar: the UNMODIFIED reference modules on one B200, PyTorch eager
(cuBLASLt GEMMs + unfused ATen kernels), same Llama-3.2-1B LoRA r=64 workload and the same synthetic
Alpaca-shaped batches as bench.py, timed with CUDA events over Trainer.train_batch.

    python tools/reference_on_gpu.py [--batch 4] [--steps 5] [--warmup 2] > gpurun_out/ref_gpu.json
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--seq", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    args = ap.parse_args()
    from baseline import reference_arm as RA
    from bench import LORA_ALPHA, LORA_RANK, flops_per_token, load_peaks
    from building_llm_from_scratch_b200.datautils.synthetic import alpaca_batch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    cfg = RA.llama32_1b_cfg(args.seq)
    tr = RA.build_llama_lora_trainer(cfg, LORA_RANK, LORA_ALPHA, device=dev)
    gen = torch.Generator().manual_seed(1234)
    batches = [alpaca_batch(gen, args.batch, args.seq, cfg["vocab_size"]) for _ in range(4)]
    batches = [(x.pin_memory(), y.pin_memory()) for x, y in batches]
    for i in range(args.warmup):
        RA.train_batch(tr, *batches[i % 4])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        RA.train_batch(tr, *batches[i % 4])
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    tok_s = args.batch * args.seq / (ms / 1e3)
    peaks = load_peaks()
    fpt = flops_per_token(cfg, LORA_RANK, args.seq)
    print(json.dumps({
        "what": "unmodified reference modules (baseline/_ref), PyTorch eager on 1 x B200, Trainer.train_batch",
        "workload": f"Llama-3.2-1B LoRA r={LORA_RANK} bf16, micro-batch {args.batch} x seq {args.seq}, Alpaca-shaped synthetic",
        "ms_per_step": ms, "tokens_per_s": tok_s, "steps": args.steps, "warmup": args.warmup,
        "frac_of_burst_bf16_peak": tok_s * fpt / 1e12 / peaks["tf_burst"],
        "max_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30, "torch": torch.__version__}))


if __name__ == "__main__":
    main()
