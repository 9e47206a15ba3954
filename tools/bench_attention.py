"""Micro-benchmark of the attention kernels at the bench shape (not a bench value)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from building_llm_from_scratch_b200 import kernels as K

# default: the bench shape (Llama-3.2-1B); `hd128` = Llama-3-8B's attention at B1 x T4096
if "hd128" in sys.argv:
    B, T, h, hkv, hd = 1, 4096, 32, 8, 128
else:
    B, T, h, hkv, hd = 4, 2048, 32, 8, 64
g = torch.Generator(device="cuda").manual_seed(0)
qkv = (torch.randn(B * T, (h + 2 * hkv) * hd, generator=g, device="cuda")).to(torch.bfloat16)
q, k, v = qkv[:, :h * hd], qkv[:, h * hd:(h + hkv) * hd], qkv[:, (h + hkv) * hd:]
do = torch.randn(B * T, h * hd, generator=g, device="cuda").to(torch.bfloat16)
flush = torch.empty(256 * 1024 * 1024, device="cuda", dtype=torch.uint8)


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


if len(sys.argv) > 1 and sys.argv[1] == "once":     # for ncu: one forward + one backward
    o, lse = K.attention_fwd(q, k, v, B, T, h, hkv, hd)
    K.attention_bwd(q, k, v, o, do, lse, B, T, h, hkv, hd)
    torch.cuda.synchronize()
    sys.exit(0)
fwd_flops = 2 * 2 * B * h * T * T * hd / 2
o, lse = K.attention_fwd(q, k, v, B, T, h, hkv, hd)
t_f = timeit(lambda: K.attention_fwd(q, k, v, B, T, h, hkv, hd))
t_b = timeit(lambda: K.attention_bwd(q, k, v, o, do, lse, B, T, h, hkv, hd))
print(f"B{B} T{T} h{h} kv{hkv} hd{hd} legacy={os.environ.get('B200_ATTN_LEGACY','0')} fwd {t_f*1e3:.1f} us ({fwd_flops/t_f/1e9:.0f} TF/s)  "
      f"bwd {t_b*1e3:.1f} us ({2.5*fwd_flops/t_b/1e9:.0f} TF/s algorithmic)")
