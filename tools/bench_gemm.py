"""Micro-benchmark of the GEMM at the bench workload's shapes (not a bench value)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from building_llm_from_scratch_b200 import kernels as K

flush = torch.empty(256 * 1024 * 1024, device="cuda", dtype=torch.uint8)


def timeit(fn, n=8):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


def rnd(*s):
    return (torch.randn(*s, device="cuda") * 0.05).to(torch.bfloat16)


M = 8192
tag = "1cta" if os.environ.get("B200_GEMM_1CTA") == "1" else "2cta"
for (N, Kd, b_mn, r) in [(2048, 2048, 0, 64), (8192, 2048, 0, 64), (2048, 8192, 0, 64), (2048, 8192, 1, 64),
                         (8192, 2048, 1, 64), (512, 2048, 0, 64), (128256, 2048, 0, 64), (2048, 128256, 1, 64),
                         (3072, 2048, 0, 0), (16384, 2048, 0, 0)]:
    a = rnd(M, Kd)
    b = rnd(Kd, N) if b_mn else rnd(N, Kd)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    kw = {}
    if r:
        kw = dict(a2=rnd(M, r), b2=rnd(N, r))
    ms = timeit(lambda: K.gemm_ex(a, b, b1_mn=bool(b_mn), out=out, **kw))
    fl = 2.0 * M * N * (Kd + r)
    print(f"{tag} M={M} N={N} K={Kd}+{r} b_mn={b_mn}: {ms*1e3:8.1f} us  {fl/ms/1e9:7.1f} TF/s")

# the two input-gradient GEMMs with the most DRAM traffic per launch: fc1|fc2 (K = 16384) and the vocab projection
for (Mx, N, Kd, r) in [(8192, 2048, 16384, 128), (3584, 2048, 128256, 64), (3072, 2048, 128256, 64)]:
    a, b = rnd(Mx, Kd), rnd(Kd, N)
    out = torch.empty(Mx, N, device="cuda", dtype=torch.bfloat16)
    a2, b2 = rnd(Mx, r), rnd(N, r)
    ms = timeit(lambda: K.gemm_ex(a, b, b1_mn=True, out=out, a2=a2, b2=b2))
    print(f"{tag} dX GEMM M={Mx} N={N} K={Kd}+{r} b_mn=1: {ms*1e3:8.1f} us  {2.0*Mx*N*(Kd+r)/ms/1e9:7.1f} TF/s")
    del a, b

# short-K GEMMs with a residual epilogue (the dX chain of the k / v projections) and o-proj
for (N, Kd) in [(2048, 512), (2048, 2048)]:
    a, b = rnd(M, Kd), rnd(Kd, N)
    res = rnd(M, N)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    ms = timeit(lambda: K.gemm_ex(a, b, b1_mn=True, out=out, residual=res, a2=rnd(M, 64), b2=rnd(N, 64)))
    print(f"{tag} residual epilogue M={M} N={N} K={Kd}+64: {ms*1e3:8.1f} us  {2.0*M*N*(Kd+64)/ms/1e9:7.1f} TF/s")
