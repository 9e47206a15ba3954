"""GPT2-1.5B's first MLP GEMM with bias + GELU in the epilogue ([4096 x 6400 x 1600], two outputs) against the plain GEMM
of the same shape. `once` = one launch of each for ncu. Not a bench value."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from building_llm_from_scratch_b200 import kernels as K

dev = torch.device("cuda", 0)
M, N, Kd = 4096, 6400, 1600
g = torch.Generator(device="cuda").manual_seed(0)
x = (torch.randn(M, Kd, generator=g, device=dev) * 0.5).to(torch.bfloat16)
w = (torch.randn(N, Kd, generator=g, device=dev) * 0.05).to(torch.bfloat16)
b = (torch.randn(N, generator=g, device=dev) * 0.1).to(torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


if "once" in sys.argv:
    K.gemm_ex(x, w, bias=b)
    K.gemm_gelu_fwd(x, w, b)
    torch.cuda.synchronize()
    sys.exit(0)
fl = 2.0 * M * N * Kd
for name, fn in (("plain + bias", lambda: K.gemm_ex(x, w, bias=b)), ("bias + GELU, z and h stored", lambda: K.gemm_gelu_fwd(x, w, b))):
    ms = timeit(fn)
    print(f"[{M} x {N} x {Kd}] {name}: {ms*1e3:7.1f} us  {fl/ms/1e9:7.1f} TF/s")
