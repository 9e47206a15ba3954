"""Training steps of BASELINE.json's other configs through the same step engine as bench.py, on 1 or N GPUs.

    python tools/run_config.py {gpt2-774m-lora|gpt2-1.5b|llama3-8b} [--steps K] [--warmup W] [--batch B] [--layers N]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/run_config.py gpt2-1.5b ...

One bench-format JSON line on rank 0: whole-job tokens/s (device-timed, CUDA events, barrier + synchronize on both
sides, max over ranks), fraction of the measured burst bf16 peak on the algorithmic FLOPs of SURVEY.md §8(d)
(activation-checkpoint recompute excluded), and for N > 1 the parameter-checksum self-check plus the exposed cost of
the gradient exchange (the same captured step with the exchange left out, timed the same way).
GPT-2 runs at the reference's drop_rate 0.1 (Models/GPT2/config.py) unless --drop says otherwise.
"""
import argparse
import copy
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def algorithmic_flops_per_token(model, seq_len, d, n_layers):
    """6 * P_lin (trainable linears) or 4 * P_lin + 6 * P_lora (frozen base) + causal attention."""
    from building_llm_from_scratch_b200.lora import LinearWithLoRA
    p_lin = p_lora = 0
    wrapped = {id(m.linear) for m in model.modules() if isinstance(m, LinearWithLoRA)}
    for m in model.modules():
        if isinstance(m, LinearWithLoRA):
            p_lin += m.linear.weight.numel()
            p_lora += m.lora.A.numel() + m.lora.B.numel()
        elif isinstance(m, torch.nn.Linear) and id(m) not in wrapped:
            p_lin += m.weight.numel()
    frozen = bool(wrapped)
    attn = 0.5 * (4 + 8) * seq_len * d * n_layers
    return (4 * p_lin + 6 * p_lora if frozen else 6 * p_lin) + attn


def build_config(name, dev, layers=None, drop=None):
    """-> (model on dev in train mode, cfg, seq_len, description) for one of BASELINE.json's other configs."""
    from building_llm_from_scratch_b200.lora import replace_linear_with_lora
    torch.manual_seed(123)
    if name.startswith("gpt2"):
        from building_llm_from_scratch_b200.Models.GPT2.GPT2 import GPTModel
        from building_llm_from_scratch_b200.Models.GPT2.config import get_config_gpt2
        cfg = copy.deepcopy(get_config_gpt2("774M" if "774" in name else "1.5B"))
        if drop is not None:
            cfg["drop_rate"] = drop
        if layers:
            cfg["n_layers"] = layers
        lora = "lora" in name
        with torch.device(dev):
            model = GPTModel(cfg, use_actv_ckpt=lora)      # configs[1]: LoRA r=8 + activation checkpointing
        if lora:
            for p in model.parameters():
                p.requires_grad = False
            with torch.device(dev):
                replace_linear_with_lora(model, rank=8, alpha=16, dtype=torch.float32)
        model = model.to(torch.bfloat16)                   # harness cast
        T = 1024
        what = ("GPT2-774M LoRA r=8 + activation checkpointing" if lora else "GPT2-1.5B full pretraining") + \
            f", bf16, drop_rate {cfg['drop_rate']}"
    else:
        from building_llm_from_scratch_b200.Models.Llama.Llama3 import Llama3Model
        from building_llm_from_scratch_b200.Models.Llama.config import get_config_llama
        cfg = get_config_llama("8B", "llama3", target_context_length=4096)
        if layers:
            cfg["n_layers"] = layers
        with torch.device(dev):
            model = Llama3Model(cfg, use_actv_ckpt=True)
        T = 4096
        what = "Llama-3-8B GQA full pretraining, bf16, activation checkpointing"
    model.train()
    return model, cfg, T, what


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", choices=["gpt2-774m-lora", "gpt2-1.5b", "llama3-8b"])
    ap.add_argument("--layers", type=int, default=None, help="debug: fewer blocks (not a result for the config)")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=4, help="micro-batch per GPU (args.py:53 default 4)")
    ap.add_argument("--drop", type=float, default=None, help="GPT-2 drop_rate (default: the config's 0.1)")
    ap.add_argument("--eager", action="store_true", help="no CUDA graph")
    ap.add_argument("--zero", action="store_true", help="ZeRO-1 optimizer-state sharding (--use_zero_opt)")
    ap.add_argument("--bucket-mb", type=float, default=64.0)
    ap.add_argument("--sm-limit", type=int, default=0, help="SMs the persistent kernels may fill (0 = all): leaves room for NCCL")
    args = ap.parse_args()

    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        opts = dist.ProcessGroupNCCL.Options()       # see bench.nccl_options
        opts.is_high_priority_stream = os.environ.get("B200_NCCL_HIGH_PRIORITY", "1") != "0"
        dist.init_process_group("nccl", device_id=dev, pg_options=opts)

    from building_llm_from_scratch_b200 import _lib
    from building_llm_from_scratch_b200.datautils.synthetic import gutenberg_batch
    from building_llm_from_scratch_b200.ddp import B200DDP
    from building_llm_from_scratch_b200.engine import TrainStep
    from building_llm_from_scratch_b200.optim import B200AdamW
    from building_llm_from_scratch_b200.train import Trainer
    _lib.check(_lib.load().b200_init(local_rank), "init")
    if args.sm_limit:
        _lib.check(_lib.load().b200_set_sm_limit(args.sm_limit), "set_sm_limit")
    model, cfg, T, what = build_config(args.config, dev, args.layers, args.drop)
    n_train = sum(p.numel() for p in model.parameters() if p.requires_grad)
    fpt = algorithmic_flops_per_token(model, T, cfg["emb_dim"], cfg["n_layers"])
    if world > 1:
        model = B200DDP(model, bucket_mb=args.bucket_mb)
    opt = B200AdamW(model.parameters(), lr=5e-4, weight_decay=0.1, zero=args.zero)
    tr = Trainer(model, opt, cfg, [], None, "/tmp", device=dev, rank=rank)
    tr.set_schedule(1000)
    gen = torch.Generator().manual_seed(7 + rank)
    batches = [tuple(t.to(dev) for t in gutenberg_batch(gen, args.batch, T, cfg["vocab_size"])) for _ in range(2)]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step, steps):
        losses = []
        for i in range(args.warmup):
            step.run(*batches[i % 2])
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            losses.append(step.run(*batches[i % 2]).clone())
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        return ms / steps, [float(l) for l in losses]

    n0 = _lib.launch_count()
    step = TrainStep(tr, (args.batch, T), use_graph=not args.eager)
    step.compact_head = False                              # Gutenberg-shaped: no ignored targets
    step.capture(warmup=2)
    ms, losses = timed(step, args.steps)
    identical = None
    if world > 1:
        from building_llm_from_scratch_b200.ddp import ranks_identical
        identical = ranks_identical([g["p"] for g in opt._groups.values()])
    comm_exposed = None
    if world > 1 and not args.eager and not args.zero:
        local = TrainStep(tr, (args.batch, T), use_graph=True)
        local.ddp = None
        local.compact_head = False
        local.capture(warmup=1)
        ms_local, _ = timed(local, args.steps)
        comm_exposed = ms - ms_local
    peaks_path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    burst = json.load(open(peaks_path))["bf16_tflops"] if os.path.exists(peaks_path) else 1590.0
    tok_s = args.batch * T * world / (ms / 1e3)
    if rank == 0:
        print(json.dumps({
            "config": args.config, "workload": f"{what}, micro-batch {args.batch} x seq {T} per GPU, Gutenberg-shaped synthetic",
            "n_gpus": world, "n_layers": cfg["n_layers"], "trainable_params": n_train,
            "mode": ("eager" if args.eager else "cuda_graph") + (" + ZeRO-1" if args.zero else ""),
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "tokens_per_s": tok_s,
            "algorithmic_gflops_per_token": fpt / 1e9,
            "frac_of_burst_bf16_peak_per_gpu": tok_s / world * fpt / 1e12 / burst, "burst_peak_tflops": burst,
            "ranks_identical": identical, "comm_exposed_ms": comm_exposed, "sm_limit": args.sm_limit,
            "allreduce_bytes_per_step": None if world == 1 else 4 * sum(g["g"].numel() for g in opt._groups.values()),
            "losses": [round(l, 4) for l in losses], "b200llm_launches_per_step": step.launches_per_step,
            "b200llm_launches_total": _lib.launch_count() - n0,
            "max_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}), flush=True)
    if world > 1:
        step.close()
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
