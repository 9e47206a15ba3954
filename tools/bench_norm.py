"""RMSNorm backward with the residual-stream gradient added (the form the Llama block's backward runs), cold L2. Not a bench value."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from building_llm_from_scratch_b200 import kernels as K
dev = torch.device("cuda", 0)
M, d = 8192, 2048
bf = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16)
x, dy, dres, w = bf(M, d), bf(M, d), bf(M, d), torch.ones(d, device=dev)
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
        for _ in range(7):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        return sorted(ts)[3]
    return (run(gb) - run(ga)) / 5


ms = timeit(lambda: K.rmsnorm_bwd(x, w, dy, dres=dres))
print(f"rmsnorm_bwd + dres [8192x2048] (8 B/elt): {ms*1e3:6.1f} us  {8*M*d/ms/1e6:6.0f} GB/s")
ms = timeit(lambda: K.rmsnorm_bwd(x, w, dy))
print(f"rmsnorm_bwd [8192x2048] (6 B/elt): {ms*1e3:6.1f} us  {6*M*d/ms/1e6:6.0f} GB/s")
ms = timeit(lambda: K.rmsnorm_fwd(x, w, res=dres))
print(f"rmsnorm_fwd + residual [8192x2048] (8 B/elt): {ms*1e3:6.1f} us  {8*M*d/ms/1e6:6.0f} GB/s")
for N in (1600, 6400):
    dyb = bf(4096, N)
    acc = torch.zeros(N, device=dev)
    ms = timeit(lambda: K.colsum(dyb, out=acc))
    print(f"colsum (bias gradient) [4096x{N}] (2 B/elt): {ms*1e3:6.1f} us  {2*4096*N/ms/1e6:6.0f} GB/s")
xl, dyl, wl, bl = bf(4096, 1600), bf(4096, 1600), bf(1600), bf(1600)
_, mean, rstd = K.layernorm_fwd(xl, wl, bl)
ms = timeit(lambda: K.layernorm_bwd(xl, wl, dyl, mean, rstd, need_dw=True))
print(f"layernorm_bwd with dw, db [4096x1600] (6 B/elt): {ms*1e3:6.1f} us  {6*4096*1600/ms/1e6:6.0f} GB/s")
ms = timeit(lambda: K.layernorm_bwd(xl, wl, dyl, mean, rstd, need_dw=False))
print(f"layernorm_bwd dx only [4096x1600] (6 B/elt): {ms*1e3:6.1f} us  {6*4096*1600/ms/1e6:6.0f} GB/s")
ms = timeit(lambda: K.layernorm_fwd(xl, wl, bl))
print(f"layernorm_fwd [4096x1600] (4 B/elt): {ms*1e3:6.1f} us  {4*4096*1600/ms/1e6:6.0f} GB/s")
