"""AdamW (bf16 params, fp32 grads) and the fp32 gradient norm at the bench's 53.4 M LoRA parameters and at 400 M, cold L2
(graph replays, memset before each call). Not a bench value."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from building_llm_from_scratch_b200 import kernels as K
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn):
    fn(); torch.cuda.synchronize()
    ga, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
    with torch.cuda.graph(ga):
        for _ in range(5):
            flush.zero_()
    with torch.cuda.graph(gb):
        for _ in range(5):
            flush.zero_(); fn()
    def run(g):
        ts = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        return sorted(ts)[2]
    return (run(gb) - run(ga)) / 5


for n in (53_428_224, 400_000_000):
    p = (torch.randn(n, device=dev) * 0.02).to(torch.bfloat16)
    g = torch.randn(n, device=dev) * 1e-3
    m = torch.zeros(n, device=dev, dtype=torch.bfloat16)
    v = torch.zeros(n, device=dev, dtype=torch.bfloat16)
    ms = timeit(lambda: K.adamw_(p, g, m, v, 1, 1e-4))
    print(f"adamw bf16 n={n/1e6:.1f} M: {ms*1e3:8.1f} us  {16*n/ms/1e6:7.0f} GB/s at 16 B/param")
    ms = timeit(lambda: K.grad_norm(g))
    print(f"grad-norm fp32 n={n/1e6:.1f} M: {ms*1e3:8.1f} us  {4*n/ms/1e6:7.0f} GB/s at 4 B/param")
    del p, g, m, v
