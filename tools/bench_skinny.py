"""Micro-benchmark of the skinny (N <= 64) grouped GEMM at the bench workload's shapes, both schedulers
(stream-K vs tile-scheduled). HBM-bound: GB/s = operand bytes read once / time. Not a bench value."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from building_llm_from_scratch_b200 import kernels as K

dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def bf(*shape):
    return (torch.randn(*shape, device=dev) * 0.05).to(torch.bfloat16)


def out_bf(m, n):
    return torch.empty(m, n, device=dev, dtype=torch.bfloat16)


ONCE = "once" in sys.argv          # for ncu: one launch per group, default scheduler only


def bench_group(label, probs, nbytes):
    K.gemm_grouped(probs)
    if ONCE:
        torch.cuda.synchronize()
        return
    torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        K.gemm_grouped(probs)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    print(f"streamk={int(K.STREAMK)} {label}: {ms*1e3:8.1f} us  {nbytes/ms/1e6:8.1f} GB/s", flush=True)


def main():
    x = bf(8192, 2048)
    t_qkv = [dict(a=x, b=bf(2048, 64), out=out_bf(8192, 64), b_mn=True) for _ in range(3)]
    t_o = [dict(a=x, b=bf(2048, 64), out=out_bf(8192, 64), b_mn=True)]
    h = bf(8192, 8192)
    t_fc3 = [dict(a=h, b=bf(8192, 64), out=out_bf(8192, 64), b_mn=True)]
    u_fc12 = [dict(a=h, b=bf(64, 8192), out=out_bf(8192, 64)), dict(a=bf(8192, 8192), b=bf(64, 8192), out=out_bf(8192, 64))]
    dl = bf(4096, 128256)
    u_head = [dict(a=dl, b=bf(64, 128256), out=out_bf(4096, 64))]
    wg, wg_bytes = [], 0
    for d_in, d_out in [(2048, 2048), (2048, 512), (2048, 512), (2048, 2048), (2048, 8192), (2048, 8192), (8192, 2048)]:
        xi, u, dy, t = bf(8192, d_in), bf(8192, 64), bf(8192, d_out), bf(8192, 64)
        wg.append(dict(a=xi, b=u, out=torch.zeros(d_in, 64, device=dev), a_mn=True, b_mn=True, accumulate=True))
        wg.append(dict(a=dy, b=t, out=torch.zeros(64, d_out, device=dev), a_mn=True, b_mn=True, accumulate=True, trans_out=True))
        wg_bytes += (xi.numel() + dy.numel()) * 2
    for sk in ((True,) if ONCE else (True, False)):
        K.STREAMK = sk
        bench_group("t = x A for q,k,v (3 x 8192x64x2048, shared x)", t_qkv, x.numel() * 2)
        bench_group("t = x A for o (8192x64x2048)", t_o, x.numel() * 2)
        bench_group("t = h A for fc3 (8192x64x8192)", t_fc3, h.numel() * 2)
        bench_group("u = dY B^T for fc1, fc2 (2 x 8192x64x8192)", u_fc12, 2 * h.numel() * 2)
        bench_group("u = dlogits B^T (4096x64x128256)", u_head, dl.numel() * 2)
        bench_group("dA, dB of one Llama block (14 problems)", wg, wg_bytes)
    K.STREAMK = True


if __name__ == "__main__":
    main()
