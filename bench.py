#!/usr/bin/env python
"""bench.py — train tokens/sec of the decoder training step (Llama-3.2-1B LoRA bf16, BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one Trainer.train_batch over one synthetic Alpaca-shaped batch (per-GPU micro-batch
B=4 x T=2048, the reference default, args.py:53): zero_grad, LR schedule, forward + fused vocab
projection/cross-entropy, backward, [data-parallel gradient mean], clip 1.0, AdamW.

Prints ONE JSON line (rank 0): `value` = whole-job tokens/s with inputs resident in HBM,
device-timed (CUDA events, max over ranks); `e2e` = the same metric through the public API
(engine.TrainStep.run) with pinned HOST batches (H2D of tokens+targets and D2H of the loss inside the
timed region); `roofline` for the dominant kernel (the tcgen05 GEMM) from a separately instrumented
pass; `cpu_baseline` = the oracle (CPU port of the reference modules) timed on this box's host cores
on a bounded sample. `--impl reference` times that CPU path alone.
"""
from __future__ import annotations

import argparse
import copy
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "train tokens/sec (Llama-3.2-1B LoRA bf16)"
UNIT = "tokens/s"
SEQ_LEN = 2048
MICRO_BATCH = 4
LORA_RANK, LORA_ALPHA = 64, 32          # args.py:89-90 defaults
PEAK_LR, WEIGHT_DECAY = 5e-4, 0.1       # args.py lr default / build_components.py:258


def llama32_1b_cfg(seq_len=SEQ_LEN, n_layers=None):
    from building_llm_from_scratch_b200.Models.Llama.config import get_config_llama
    cfg = get_config_llama("1B", "llama3_2", target_context_length=seq_len)   # SURVEY.md R4 override
    if n_layers is not None:
        cfg["n_layers"] = n_layers
    return cfg


def flops_per_token(cfg, rank, seq_len):
    """LoRA with frozen base = 4*P_lin + 6*P_lora + causal attention."""
    d, f, L, V = cfg["emb_dim"], cfg["hidden_dim"], cfg["n_layers"], cfg["vocab_size"]
    hd = d // cfg["n_heads"]
    kv = cfg["n_kv_groups"] * hd
    lin = [(d, d), (d, kv), (d, kv), (d, d), (d, f), (d, f), (f, d)]
    p_lin = L * sum(i * o for i, o in lin) + d * V
    p_lora = L * sum(rank * (i + o) for i, o in lin) + rank * (d + V)
    attn = 0.5 * (4 + 8) * seq_len * d * L
    return 4 * p_lin + 6 * p_lora + attn


def attn_flops_per_token(cfg, seq_len):
    return 0.5 * (4 + 8) * seq_len * cfg["emb_dim"] * cfg["n_layers"]


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return dict(tf_burst=p["bf16_tflops"], tf_sustained=p["bf16_tflops_sustained"], hbm=p["hbm_gbs"], src="measured")
    return dict(tf_burst=1590.0, tf_sustained=1400.0, hbm=6650.0, src="fallback")


def kernel_rooflines(device, peaks):
    """Every other kernel of the step against the roofline that bounds it, timed alone with CUDA events
    (graph replays, a 256 MB write before each call evicts the L2) at the shapes of the bench workload:
    HBM-bound kernels as algorithmic bytes / time vs the measured copy bandwidth, attention as causal
    FLOPs / time vs the burst bf16 peak. Not part of the timed steps."""
    from building_llm_from_scratch_b200 import kernels as K
    M, d, F, V = MICRO_BATCH * SEQ_LEN, 2048, 8192, 128256
    B, T, h, hkv, hd = MICRO_BATCH, SEQ_LEN, 32, 8, 64
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)
    bf = lambda *sh: (torch.randn(*sh, device=device) * 0.5).to(torch.bfloat16)

    def timeit(fn):
        """ms per call: a graph of 5 x (L2-evicting memset, fn) minus a graph of the 5 memsets alone, so
        neither the host's launch latency nor the memset is in the number."""
        fn()
        torch.cuda.synchronize()
        ga, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(ga):
            for _ in range(5):
                flush.zero_()
        with torch.cuda.graph(gb):
            for _ in range(5):
                flush.zero_()
                fn()

        def run(g):
            ts = []
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); g.replay(); e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            return sorted(ts)[2]
        ms = (run(gb) - run(ga)) / 5
        del ga, gb
        return ms

    rows = []

    def hbm(name, nbytes, fn):
        ms = timeit(fn)
        gbs = nbytes / ms / 1e6
        rows.append({"kernel": name, "bound": "hbm", "us": round(ms * 1e3, 1), "achieved": round(gbs, 1),
                     "peak": peaks["hbm"], "unit": "GB/s", "frac": round(gbs / peaks["hbm"], 3)})

    x, r, w = bf(M, d), bf(M, d), torch.ones(d, device=device)
    hbm("rmsnorm_fwd + residual add [8192x2048] (8 B/elt)", 8 * M * d, lambda: K.rmsnorm_fwd(x, w, res=r))
    hbm("rmsnorm_bwd [8192x2048] (6 B/elt)", 6 * M * d, lambda: K.rmsnorm_bwd(x, w, r))
    a, b, dh = bf(M, F), bf(M, F), bf(M, F)
    hbm("swiglu_fwd [8192x8192] (6 B/elt)", 6 * M * F, lambda: K.swiglu_fwd(a, b))
    da, db = torch.empty_like(a), torch.empty_like(a)
    hbm("swiglu_bwd [8192x8192] (10 B/elt)", 10 * M * F, lambda: K.swiglu_bwd(a, b, dh, da=da, db=db))
    del a, b, dh, da, db
    cos, sin = bf(T, hd), bf(T, hd)
    q = bf(M, h * hd)
    hbm("rope q [8192 x 32 heads x 64] in place (4 B/elt)", 4 * M * h * hd, lambda: K.rope_(q, cos, sin, B, T, h, hd))
    rows_ce = 4096
    logits = bf(rows_ce, V)
    tg = torch.randint(0, V, (rows_ce,), device=device)
    hbm("cross-entropy fwd+bwd in place [4096x128256] (6 B/logit)", 6 * rows_ce * V, lambda: K.cross_entropy_(logits, tg))
    del logits
    # rank-64 LoRA products (gemm_grouped / gemm_streamk): bytes of the streamed activations, read once
    A3 = [bf(d, 64) for _ in range(3)]
    t3 = [torch.empty(M, 64, device=device, dtype=torch.bfloat16) for _ in range(3)]
    hbm("rank-64 product t = x A for q|k|v [3 x 8192x64x2048, x shared] (x read once: 2 B/elt)", 2 * M * d,
        lambda: K.gemm_grouped([dict(a=x, b=A3[i], out=t3[i], b_mn=True) for i in range(3)]))
    hbm("rank-64 product t = x A, one projection [8192x64x2048]", 2 * M * d,
        lambda: K.gemm_grouped([dict(a=x, b=A3[0], out=t3[0], b_mn=True)]))
    hF, AF = bf(M, F), bf(F, 64)
    hbm("rank-64 product t = h A for fc3 [8192x64x8192]", 2 * M * F,
        lambda: K.gemm_grouped([dict(a=hF, b=AF, out=t3[0], b_mn=True)]))
    wg, wg_bytes = [], 0
    for d_in, d_out in [(d, d), (d, 512), (d, 512), (d, d), (d, F), (d, F), (F, d)]:
        xi, u, dy, t = (hF if d_in == F else x), bf(M, 64), (hF if d_out == F else bf(M, d_out)), bf(M, 64)
        wg.append(dict(a=xi, b=u, out=torch.zeros(d_in, 64, device=device), a_mn=True, b_mn=True, accumulate=True))
        wg.append(dict(a=dy, b=t, out=torch.zeros(64, d_out, device=device), a_mn=True, b_mn=True, accumulate=True, trans_out=True))
        wg_bytes += 2 * M * (d_in + d_out)
    hbm("rank-64 weight gradients dA, dB of one block [14 problems, one launch] (x and dY read once)", wg_bytes,
        lambda: K.gemm_grouped(wg))
    del hF, AF, wg, A3, t3
    n = 53_428_224
    p_, g_, m_, v_ = (torch.zeros(n, device=device, dtype=torch.bfloat16), torch.zeros(n, device=device),
                      torch.zeros(n, device=device, dtype=torch.bfloat16), torch.zeros(n, device=device, dtype=torch.bfloat16))
    hbm("adamw bf16 params, fp32 grads [53.4 M] (16 B/param)", 16 * n, lambda: K.adamw_(p_, g_, m_, v_, 1, 1e-4))
    hbm("grad-norm fp32 [53.4 M] (4 B/param)", 4 * n, lambda: K.grad_norm(g_))
    del p_, g_, m_, v_
    qkv = bf(M, (h + 2 * hkv) * hd)
    q, k, v = qkv[:, :h * hd], qkv[:, h * hd:(h + hkv) * hd], qkv[:, (h + hkv) * hd:]
    do = bf(M, h * hd)
    o, lse = K.attention_fwd(q, k, v, B, T, h, hkv, hd)
    fl = 2.0 * 2 * B * h * T * T * hd / 2
    for name, flops, fn in (("attention fwd B4 T2048 h32 kv8 hd64 (causal FLOPs)", fl, lambda: K.attention_fwd(q, k, v, B, T, h, hkv, hd)),
                            ("attention bwd (2.5 x fwd FLOPs)", 2.5 * fl, lambda: K.attention_bwd(q, k, v, o, do, lse, B, T, h, hkv, hd))):
        ms = timeit(fn)
        tf = flops / ms / 1e9
        rows.append({"kernel": name, "bound": "tensor+mufu", "us": round(ms * 1e3, 1), "achieved": round(tf, 1),
                     "peak": peaks["tf_burst"], "unit": "TFLOP/s", "frac": round(tf / peaks["tf_burst"], 3)})
    return rows


def gemm_traffic_from_profile():
    """DRAM bytes per launch of the dominant kernel (CTA-pair GEMM) from the committed ncu capture
    (profiles/r2_gemm_traffic_2layer_step.json, made by tools/ncu_launch_summary.py from one ncu pass over a 2-block
    step: dram__bytes_read.sum + dram__bytes_write.sum summed over the kernel's launches; the round-1 file when that one
    is absent); None if neither exists."""
    for rel in ("profiles/r2_gemm_traffic_2layer_step.json", "profiles/r1_gemm_traffic_2layer_step_v3.json"):
        try:
            d = json.load(open(os.path.join(ROOT, rel)))
            for name, v in d.items():
                if "2cta" in name and "<0>" in name:
                    return {"dram_bytes_per_launch": v["dram_bytes"] / v["launches"], "launches": v["launches"],
                            "tensor_active_pct": v.get("tensor_active_pct"), "source": rel}
        except Exception:
            continue
    return None


class ClockSampler(threading.Thread):
    """Samples SM clocks / throttle reasons with nvidia-smi during the timed region."""

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.gpu)], capture_output=True, text=True, timeout=5).stdout.strip()
                parts = [x.strip() for x in out.split(",")]
                self.samples.append(float(parts[0]))
                self.max_mhz = float(parts[1])
                for n, v in zip(names, parts[2:6]):
                    if v.lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:
                pass
            self._halt.wait(0.2)

    def stop(self):
        self._halt.set()
        self.join(timeout=3)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# --------------------------------------------------------------------------------------
# CPU baseline / reference arm: the UNMODIFIED reference (baseline/_ref) through its own Trainer.train_batch
# on the host cores. Only when that copy is missing (never on a box prepared by build()) does the oracle
# port stand in, and the line then says kind = "port".
# --------------------------------------------------------------------------------------
REF_MICRO_BATCH = 1       # bounded sample: one sequence of the workload per step (~4-7 s of CPU work per step)


def reference_cpu_steps(warmup, steps):
    """Time `steps` reference train_batch calls (after `warmup`) at micro-batch REF_MICRO_BATCH x SEQ_LEN, all 16
    blocks, LoRA r=64, bf16, every row through the vocab head (the reference has no other way) -> dict."""
    from building_llm_from_scratch_b200.datautils.synthetic import alpaca_batch
    torch.set_num_threads(os.cpu_count() or 1)
    cores = torch.get_num_threads()
    gen = torch.Generator().manual_seed(1234)
    from baseline import reference_arm as RA
    if RA.available():
        cfg = RA.llama32_1b_cfg(SEQ_LEN)
        batches = [alpaca_batch(gen, REF_MICRO_BATCH, SEQ_LEN, cfg["vocab_size"]) for _ in range(4)]
        tr = RA.build_llama_lora_trainer(cfg, LORA_RANK, LORA_ALPHA, device="cpu", lr=PEAK_LR, weight_decay=WEIGHT_DECAY)
        secs = RA.time_steps(tr, batches, warmup, steps)
        kind = "reference"
        what = ("unmodified reference (baseline/_ref): Llama3Model + replace_linear_with_lora + Trainer.train_batch "
                "(zero_grad, LR, forward, F.cross_entropy, backward, clip 1.0, AdamW)")
    else:
        from oracle import reference_ops as R
        cfg = llama32_1b_cfg(SEQ_LEN)
        batches = [alpaca_batch(gen, REF_MICRO_BATCH, SEQ_LEN, cfg["vocab_size"]) for _ in range(4)]
        sd = oracle_state(cfg, LORA_RANK)
        secs = []
        for i in range(warmup + steps):
            x, y = batches[i % len(batches)]
            t0 = time.perf_counter()
            loss = R.cross_entropy(R.llama3_forward(x, sd, cfg, LORA_ALPHA / LORA_RANK), y)
            loss.backward()
            if i >= warmup:
                secs.append(time.perf_counter() - t0)
            del loss
        kind = "port"
        what = "oracle port of the reference modules (baseline/_ref absent): forward + cross-entropy + backward, no optimizer step"
    mean = sum(secs) / len(secs)
    tokens = REF_MICRO_BATCH * SEQ_LEN
    return dict(value=tokens / mean, ms_per_step=1e3 * mean, kind=kind, cores=cores, secs=secs,
                sample=f"{what}; Llama-3.2-1B, all 16 blocks, LoRA r={LORA_RANK}, bf16, micro-batch {REF_MICRO_BATCH} x seq "
                       f"{SEQ_LEN} Alpaca-shaped, {len(secs)} timed steps after {warmup} warm-up on {cores} host threads, "
                       f"{mean:.2f} s/step (min {min(secs):.2f}, max {max(secs):.2f})")


def oracle_state(cfg, rank):
    """Random-init state_dict with the reference's key names (bf16, fp32 norm weights)."""
    d, f, V, L = cfg["emb_dim"], cfg["hidden_dim"], cfg["vocab_size"], cfg["n_layers"]
    hd = d // cfg["n_heads"]
    kv = cfg["n_kv_groups"] * hd
    from oracle import reference_ops as R
    cos, sin = R.precompute_rope_params(hd, cfg["rope_base"], cfg["context_length"], cfg["rope_freq"])
    sd = {"tok_emb.weight": (torch.randn(V, d) * 0.02).bfloat16(), "final_norm.weight": torch.ones(d)}

    def lin(prefix, i, o):
        sd[prefix + ".linear.weight"] = (torch.randn(o, i) * 0.02).bfloat16()
        sd[prefix + ".lora.A"] = ((torch.rand(i, rank) * 2 - 1) / rank ** 0.5).bfloat16().requires_grad_(True)
        sd[prefix + ".lora.B"] = (torch.randn(rank, o) * 0.01).bfloat16().requires_grad_(True)

    for l in range(L):
        p = f"trf_blocks.{l}"
        for name, (i, o) in {"att.W_query": (d, d), "att.W_key": (d, kv), "att.W_value": (d, kv), "att.out_proj": (d, d),
                             "ff.fc1": (d, f), "ff.fc2": (d, f), "ff.fc3": (f, d)}.items():
            lin(f"{p}.{name}", i, o)
        sd[f"{p}.norm1.weight"] = torch.ones(d)
        sd[f"{p}.norm2.weight"] = torch.ones(d)
        sd[f"{p}.att.cos"] = cos.bfloat16()
        sd[f"{p}.att.sin"] = sin.bfloat16()
    lin("out_head", d, V)
    return sd


def run_reference_arm(args):
    """`--impl reference`: rank 0 alone times the reference on the host cores; the other ranks exit."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = reference_cpu_steps(max(args.warmup, 1), args.steps)
    cfg = llama32_1b_cfg()
    config = workload_config(cfg, 1)
    config["workload"] += (f" -- reference arm: a bounded sample of that workload, micro-batch {REF_MICRO_BATCH} x seq {SEQ_LEN} "
                           "per step on the host CPU")
    config["global_batch"] = REF_MICRO_BATCH
    config["parallelism"] = "cpu"
    config["l2_policy"] = "n/a (host CPU)"
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": max(args.warmup, 1), "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "step_seconds": [round(t, 3) for t in r["secs"]]}
    print(json.dumps(line), flush=True)


def workload_config(cfg, n_gpus):
    return {"workload": f"Llama-3.2-1B LoRA r={LORA_RANK} alpha={LORA_ALPHA} instruction finetune, bf16, Alpaca-shaped "
                        f"synthetic batches, micro-batch {MICRO_BATCH} x seq {SEQ_LEN} per GPU (BASELINE.json configs[2])",
            "global_batch": MICRO_BATCH * n_gpus, "seq_len": SEQ_LEN, "parallelism": f"dp{n_gpus}",
            "l2_policy": "per-step working set (2.5 GB frozen weights + activations) far exceeds the 126 MB L2",
            "n_layers": cfg["n_layers"], "vocab_size": cfg["vocab_size"]}


# --------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------
def build_trainer(cfg, device, world_size):
    from building_llm_from_scratch_b200.Models.Llama import Llama3 as L
    from building_llm_from_scratch_b200.ddp import B200DDP
    from building_llm_from_scratch_b200.lora import replace_linear_with_lora
    from building_llm_from_scratch_b200.optim import B200AdamW
    from building_llm_from_scratch_b200.train import Trainer
    torch.manual_seed(123)                                       # utils.set_seed(123), main.py:58
    with torch.device(device):
        model = L.Llama3Model(copy.deepcopy(cfg))
    for p in model.parameters():                                 # activate_lora, build_components.py:117-135
        p.requires_grad = False
    with torch.device(device):
        replace_linear_with_lora(model, rank=LORA_RANK, alpha=LORA_ALPHA, dtype=cfg["dtype"])
    model.train()
    if world_size > 1:
        # bucket size of the gradient exchange (MB of fp32 gradients per all-reduce); B200_DDP_BUCKET_MB for A/B runs
        model = B200DDP(model, bucket_mb=float(os.environ.get("B200_DDP_BUCKET_MB", "64")))
    opt = B200AdamW(model.parameters(), lr=PEAK_LR, weight_decay=WEIGHT_DECAY)
    tr = Trainer(model, opt, cfg, [], None, "/tmp", device=device)
    tr.set_schedule(total_training_steps=100_000)
    return tr


def nccl_options(dist):
    """High-priority NCCL streams: the bucket all-reduces overlap backward, whose persistent GEMM kernels fill every SM;
    with default priority the collective's CTAs queue behind each newly launched GEMM and most of the exchange ends up
    after backward. B200_NCCL_HIGH_PRIORITY=0 restores the default."""
    opts = dist.ProcessGroupNCCL.Options()
    opts.is_high_priority_stream = os.environ.get("B200_NCCL_HIGH_PRIORITY", "1") != "0"
    return opts


def data_parallel_selfcheck(step, dev_batches, n_valid, world):
    """N > 1: (1) after two optimizer steps on per-rank batches the flat parameter buffers must be bit-identical on
    every rank (checksums all-gathered); (2) the exchanged gradient must equal the mean over ranks of the local
    gradients (the reference DDP's mean-of-means): local flat gradients are all-gathered and
    averaged in fp32 as the expected value, then compared with what the exchange leaves in the buffer."""
    import torch.distributed as dist
    for i in range(2):
        step.run(*dev_batches[i % len(dev_batches)], n_valid=n_valid[i % len(dev_batches)])
    torch.cuda.synchronize()
    from building_llm_from_scratch_b200.ddp import ranks_identical
    identical = ranks_identical([g["p"] for g in step.opt._groups.values()])
    # gradient mean: one eager forward/backward without the exchange, then the exchange on its own
    step.capacity = 0
    step._fwd_bwd(False)
    worst = 0.0
    for flat in step.opt.flat_grads:
        local = flat.clone()
        parts = [torch.empty_like(local) for _ in range(world)]
        dist.all_gather(parts, local)
        expect = torch.stack(parts).mean(0)
        del parts
        step.ddp.reduce_gradients([flat])
        worst = max(worst, float((flat - expect).abs().max() / expect.abs().max().clamp_min(1e-30)))
        del expect, local
    torch.cuda.synchronize()
    return {"ranks_identical": identical, "grad_mean_max_rel": worst,
            "checked": "parameter checksums after 2 steps; exchanged flat gradient vs the fp32 mean of the all-gathered local gradients"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--layers", type=int, default=None, help="debug: fewer blocks (result is then not a bench value)")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--full-head", action="store_true", help="project every row through the vocab head, ignored targets included")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    if args.impl == "reference":
        run_reference_arm(args)
        return

    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device, pg_options=nccl_options(dist))

    from building_llm_from_scratch_b200 import _lib, kernels
    from building_llm_from_scratch_b200.datautils.synthetic import alpaca_batch
    from building_llm_from_scratch_b200.engine import TrainStep
    _lib.check(_lib.load().b200_init(local_rank), "init")

    cfg = llama32_1b_cfg(SEQ_LEN, n_layers=args.layers)
    tr = build_trainer(cfg, device, world)
    gen = torch.Generator().manual_seed(1234 + rank)
    n_batches = 4
    host = [alpaca_batch(gen, MICRO_BATCH, SEQ_LEN, cfg["vocab_size"]) for _ in range(n_batches)]
    host = [(x.pin_memory(), y.pin_memory()) for x, y in host]
    dev = [(x.to(device), y.to(device)) for x, y in host]
    n_valid = [int((y != -100).sum()) for _, y in host]        # known from the host copy of each batch
    ignored = float(sum((y == -100).float().mean() for _, y in host) / n_batches)

    def make_step(**kw):
        st = TrainStep(tr, (MICRO_BATCH, SEQ_LEN), use_graph=not args.no_graph, **kw)
        st.compact_head = not args.full_head
        st.capture(warmup=2)
        for n in n_valid:
            st.prepare(n)
        return st

    step = main_step = make_step()

    # ---- data-parallel self-check (N > 1), before anything is timed: numbers, not just an exit code
    dp_check = None
    if world > 1:
        dp_check = data_parallel_selfcheck(step, dev, n_valid, world)
        if not dp_check["ranks_identical"] and step.capture_collective:
            # the single-graph form with captured collectives failed: fall back to two graphs with the
            # exchange issued by the host between them, and check again
            step.close()
            step = main_step = make_step(capture_collective=False)
            dp_check = data_parallel_selfcheck(step, dev, n_valid, world)
            dp_check["fallback"] = "captured collectives failed the self-check; two-graph replay"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(batches, read_loss, step=None):
        step = step or main_step
        for i in range(args.warmup):
            step.run(*batches[i % n_batches], n_valid=n_valid[i % n_batches])
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = _lib.launch_count()
        e0.record()
        # D2H read of every step's loss, one step behind: the copy of step i's loss into pinned memory is enqueued
        # right after step i, and the host reads it once step i + 1 has been enqueued, so the GPU is not left idle
        # while the host waits (what a training loop that logs its loss does)
        host_loss = [torch.zeros(1, dtype=torch.float32, pin_memory=True) for _ in range(2)]
        copied = [torch.cuda.Event(), torch.cuda.Event()]
        losses = []
        for i in range(args.steps):
            loss = step.run(*batches[i % n_batches], n_valid=n_valid[i % n_batches])
            if read_loss:
                host_loss[i & 1].copy_(loss.reshape(1), non_blocking=True)
                copied[i & 1].record()
                if i > 0:
                    copied[(i - 1) & 1].synchronize()
                    losses.append(float(host_loss[(i - 1) & 1]))
        if read_loss:
            copied[(args.steps - 1) & 1].synchronize()
            losses.append(float(host_loss[(args.steps - 1) & 1]))
            assert len(losses) == args.steps
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        n = _lib.launch_count() - l0
        if step.graphs is not None:   # replayed launches are not host calls: count what the graph holds
            n += step.launches_per_step * args.steps
        return ms, n

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    ms_dev, launches = timed(dev, read_loss=False)
    clocks = sampler.stop() if sampler else None
    ms_e2e, _ = timed(host, read_loss=True)
    final_loss = float(step.loss)

    # ---- secondary figures (same K steps each, outside the headline's timed region) -------------------
    # every row through the vocab projection (no row compaction: what the headline would be without -100 targets)
    ms_full = None
    if step.compact_head and step.graphs is not None:
        step.compact_head = False
        ms_full, _ = timed(dev, read_loss=False)
        step.compact_head = True
    # the drop-in API itself: Trainer.train_batch, eager (no graph), host batches
    k_eager = min(args.steps, 5)
    for i in range(2):
        tr.train_batch(*host[i % n_batches])
    barrier()
    ee0, ee1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ee0.record()
    for i in range(k_eager):
        tr.train_batch(*host[i % n_batches])
    ee1.record()
    barrier()
    ms_eager = ee0.elapsed_time(ee1)
    if world > 1:
        t = torch.tensor([ms_eager], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_eager = float(t)
    # what the gradient exchange costs a step: the same step with the exchange left out (ranks then diverge, so
    # this runs last; nothing after it uses the weights)
    comm_exposed_ms = None
    if world > 1 and step.graphs is not None:
        local = TrainStep(tr, (MICRO_BATCH, SEQ_LEN), use_graph=True)
        local.ddp = None
        local.compact_head = step.compact_head
        local.capture(warmup=1)
        for n in n_valid:
            local.prepare(n)
        ms_local, _ = timed(dev, read_loss=False, step=local)
        comm_exposed_ms = (ms_dev - ms_local) / args.steps
        del local

    tokens_step = MICRO_BATCH * SEQ_LEN * world
    value = tokens_step * args.steps / (ms_dev / 1e3)
    e2e = tokens_step * args.steps / (ms_e2e / 1e3)

    # ---- roofline of the dominant kernel: instrumented eager pass (not part of the timed runs)
    roof = None
    if rank == 0:
        peaks = load_peaks()
        prof = kernels.GemmProfile()
        kernels.set_gemm_profile(prof)
        eager = TrainStep(tr, (MICRO_BATCH, SEQ_LEN), use_graph=False)
        eager.ddp = None                          # rank-0-only pass: no collective may be issued here
        if getattr(tr.model, "detach", None):
            tr.model.detach()                     # ... nor by the wrapper's gradient-ready hooks
        eager.compact_head = step.compact_head
        eager.x.copy_(dev[0][0]); eager.y.copy_(dev[0][1])
        eager._graphs_for(n_valid[0])             # same vocab-projection capacity as the timed steps
        eager._refresh_hyper(peek=True)
        eager._eager_step()                       # warm
        prof.reset()
        eager._eager_step()
        torch.cuda.synchronize()
        kernels.set_gemm_profile(None)
        # the GEMM whose epilogue also runs the SwiGLU backward (3x the epilogue bytes) is reported on its own
        # line below: its time is not pure tensor work
        by_shape = prof.by_shape()
        fused = {k: v for k, v in by_shape.items() if isinstance(k[0], str) and k[0].startswith("swiglu")}
        # the rank-64 LoRA products run in other kernels (gemm_grouped / gemm_streamk) and are HBM-bound streams:
        # they are rows of other_kernels (bytes / time against the copy bandwidth), not part of the tensor roofline
        skinny = {k: v for k, v in by_shape.items() if isinstance(k[0], str) and k[0] == "grouped"}
        plain = {k: v for k, v in by_shape.items() if k not in fused and k not in skinny}
        flops = sum(v[2] for v in plain.values())
        ms_gemm = sum(v[1] for v in plain.values())
        n_gemm = sum(v[0] for v in plain.values())
        fused_flops = sum(v[2] for v in fused.values())
        fused_ms = sum(v[1] for v in fused.values())
        fused_n = sum(v[0] for v in fused.values())
        skinny_flops = sum(v[2] for v in skinny.values())
        fpt = flops_per_token(cfg, LORA_RANK, SEQ_LEN)
        roof = {"bound": "tensor", "kernel": "gemm_tcgen05_2cta_kernel (every launch of one step: projections, LoRA K extensions, residual / RoPE / SwiGLU-forward epilogues; the SwiGLU-backward-epilogue launches are listed under other_kernels)",
                "achieved": flops / (ms_gemm * 1e-3) / 1e12, "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                "frac": flops / (ms_gemm * 1e-3) / 1e12 / peaks["tf_sustained"], "traffic": None,
                "peak_source": f"{peaks['src']} (sustained; burst {peaks['tf_burst']})",
                "gemm_launches_per_step": n_gemm, "gemm_ms_per_step": ms_gemm,
                "gemm_share_of_step": (ms_gemm + fused_ms) / (ms_dev / args.steps),
                "rank_r_products": {"launches_per_step": sum(v[0] for v in skinny.values()),
                                    "gflop_per_step": sum(v[2] for v in skinny.values()) / 1e9,
                                    "note": "gemm_grouped / gemm_streamk launches (t = xA, u = dY B^T, dA, dB): HBM-bound, see other_kernels"},
                "step_frac_of_burst_peak": value / world * fpt / 1e12 / peaks["tf_burst"],
                "algorithmic_gflops_per_token": fpt / 1e9,
                # FLOPs the kernels really executed (LoRA K-padding included, ignored-target rows of the
                # vocab projection excluded when it is row-compacted) + causal attention
                "executed_gflops_per_token": ((flops + fused_flops + skinny_flops) / (MICRO_BATCH * SEQ_LEN) + attn_flops_per_token(cfg, SEQ_LEN)) / 1e9,
                "step_frac_of_burst_peak_executed": value / world * ((flops + fused_flops + skinny_flops) / (MICRO_BATCH * SEQ_LEN) + attn_flops_per_token(cfg, SEQ_LEN)) / 1e12 / peaks["tf_burst"]}
        roof["other_kernels"] = kernel_rooflines(device, peaks)
        if fused_n:
            tf = fused_flops / (fused_ms * 1e-3) / 1e12
            roof["other_kernels"].insert(0, {
                "kernel": "dh GEMM with SwiGLU backward in the epilogue [8192 x 8192 x (2048+64)], per launch in the step",
                "bound": "tensor + hbm epilogue (reads a|b, writes da|db: 8 B/elt)", "us": round(fused_ms / fused_n * 1e3, 1),
                "achieved": round(tf, 1), "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                "frac": round(tf / peaks["tf_sustained"], 3), "launches_per_step": fused_n})
            roof["gemm_ms_per_step_incl_fused_swiglu"] = ms_gemm + fused_ms
        tr_info = gemm_traffic_from_profile()
        if tr_info is not None:       # bytes per launch of the CTA-pair GEMM, from the committed ncu capture
            roof["traffic"] = tr_info["dram_bytes_per_launch"]
            roof["traffic_detail"] = tr_info
    if world > 1:
        dist.barrier()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = reference_cpu_steps(1, 2)       # ~20-30 s of CPU work: 1 warm-up + 2 timed reference steps
        cpu = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": workload_config(cfg, world),
                "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": 2 * MICRO_BATCH * SEQ_LEN * 8 + 28,
                        "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e / args.steps,
                        "loss_read": "every step, async copy to pinned memory, read by the host one step behind"},
                "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "cpu_baseline": cpu,
                "cuda_graph": step.graphs is not None, "final_loss": final_loss,
                "value_full_head": (tokens_step * args.steps / (ms_full / 1e3)) if ms_full else None,
                "eager_trainer_tokens_per_s": tokens_step * k_eager / (ms_eager / 1e3),
                "data_parallel": None if world == 1 else dict(
                    dp_check, comm_exposed_ms=comm_exposed_ms,
                    exchange=("bucketed NCCL all-reduce (AVG) of the flat fp32 gradient buffer, captured in the step's "
                              "graph and overlapped with backward" if step.capture_collective else
                              "two graphs per step, NCCL all-reduce (AVG) of the flat fp32 gradient buffer issued between them")),
                "ignored_target_fraction": ignored,
                "vocab_projection": ("all rows" if args.full_head else
                                     "non-ignored target rows only (exact: ignored rows reach neither the loss nor any "
                                     f"gradient); capacities {sorted(set(step.bucket_graphs) | {0})} rows, 0 = all")}
        if args.layers is not None:
            line["debug_layers"] = args.layers
        print(json.dumps(line), flush=True)
    if world > 1:
        step.close()          # graphs that captured collectives must go before the communicator
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
