"""One eager training step of the bench workload inside a cudaProfilerStart/Stop range, for
`ncu --profile-from-start off` (launch list / full capture). Not a benchmark."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=2)
    ap.add_argument("--config", default="llama32-1b-lora", choices=["llama32-1b-lora", "gpt2-774m-lora", "gpt2-1.5b", "llama3-8b"],
                    help="llama32-1b-lora = the bench workload; the others are tools/run_config.py's")
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--shapes", default=None, help="write per-GEMM-shape timing JSON here")
    args = ap.parse_args()
    from building_llm_from_scratch_b200 import kernels
    from building_llm_from_scratch_b200.datautils.synthetic import alpaca_batch
    from building_llm_from_scratch_b200.engine import TrainStep
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    if args.config == "llama32-1b-lora":
        cfg = bench.llama32_1b_cfg(bench.SEQ_LEN, n_layers=args.layers)
        tr = bench.build_trainer(cfg, dev, 1)
        x, y = alpaca_batch(torch.Generator().manual_seed(1234), bench.MICRO_BATCH, bench.SEQ_LEN, cfg["vocab_size"])
        step = TrainStep(tr, (bench.MICRO_BATCH, bench.SEQ_LEN), use_graph=False)
        step.x.copy_(x.to(dev)); step.y.copy_(y.to(dev))
        step._graphs_for(int((y != -100).sum()))     # row-compacted vocab projection, like bench.py
    else:
        sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
        from run_config import build_config
        from building_llm_from_scratch_b200.datautils.synthetic import gutenberg_batch
        from building_llm_from_scratch_b200.optim import B200AdamW
        from building_llm_from_scratch_b200.train import Trainer
        model, cfg, T, _ = build_config(args.config, dev, args.layers)
        opt = B200AdamW(model.parameters(), lr=5e-4, weight_decay=0.1)
        tr = Trainer(model, opt, cfg, [], None, "/tmp", device=dev)
        tr.set_schedule(1000)
        x, y = gutenberg_batch(torch.Generator().manual_seed(7), args.batch, T, cfg["vocab_size"])
        step = TrainStep(tr, (args.batch, T), use_graph=False)
        step.compact_head = False
        step.x.copy_(x.to(dev)); step.y.copy_(y.to(dev))
    step._refresh_hyper(peek=True)
    for _ in range(2):
        step._eager_step()
    torch.cuda.synchronize()
    if args.shapes:
        prof = kernels.GemmProfile()
        kernels.set_gemm_profile(prof)
        step._eager_step()
        torch.cuda.synchronize()
        kernels.set_gemm_profile(None)
        rows = [dict(shape=list(k), launches=v[0], ms=v[1], tflops=v[2] / (v[1] * 1e-3) / 1e12 if v[1] > 0 else 0)
                for k, v in sorted(prof.by_shape().items(), key=lambda kv: -kv[1][1])]
        json.dump(rows, open(args.shapes, "w"), indent=1)
    from building_llm_from_scratch_b200 import _lib
    before = _lib.gemm_path_counts()
    torch.cuda.profiler.start()
    step._eager_step()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    after = _lib.gemm_path_counts()
    print("profiled one step, loss", float(step.loss))
    print("GEMM launches by scheduling path:", {k: after[k] - before[k] for k in after if after[k] != before[k]})


if __name__ == "__main__":
    main()
