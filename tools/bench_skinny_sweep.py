"""K sweep of one skinny product (M=8192, N=64) on the tile-scheduled grouped kernel: separates the fixed cost of a launch
from the per-byte cost (64 CTAs, one 128-row tile each). Graph replays, warm (operand L2-resident when it fits) and cold."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from building_llm_from_scratch_b200 import kernels as K

dev = torch.device("cuda", 0)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)


def bf(*shape):
    return (torch.randn(*shape, device=dev) * 0.05).to(torch.bfloat16)


def graph_us(fn, n, cold):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(n):
            if cold:
                flush.zero_()
            fn()
    gf = None
    if cold:
        gf = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gf):
            for _ in range(n):
                flush.zero_()
    def t(gr):
        gr.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1)
    return (t(g) - (t(gf) if cold else 0.0)) * 1e3 / n


for sk in (False, True):
    K.STREAMK = sk
    for M in (8192,):
        for Kd in (512, 1024, 2048, 4096, 8192):
            x = bf(M, Kd); A = bf(Kd, 64); out = torch.empty(M, 64, device=dev, dtype=torch.bfloat16)
            fn = lambda: K.gemm_grouped([dict(a=x, b=A, out=out, b_mn=True)])
            w = graph_us(fn, 20, False); c = graph_us(fn, 20, True)
            print(f"streamk={int(sk)} M={M} K={Kd}: warm {w:6.1f} us  cold {c:6.1f} us   ({x.numel()*2/1e6:.0f} MB)", flush=True)
K.STREAMK = True
