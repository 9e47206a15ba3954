"""What a rank-64 product costs INSIDE a step: CUDA-graph replays of [producer -> t = x A -> large GEMM that consumes
t as its K extension] over 8 sets of buffers, with and without the skinny launch (the difference is its cost with the
producer's output as warm in L2 as it is in the real step). Not a bench value."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from building_llm_from_scratch_b200 import kernels as K

dev = torch.device("cuda", 0)
L = 8


def bf(*shape):
    return (torch.randn(*shape, device=dev) * 0.05).to(torch.bfloat16)


def graph_ms(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    for _ in range(2):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def case(label, d_in, d_out, n_t, producer):
    """producer(l) -> x_l [8192, d_in]; t = x A (n_t problems sharing x); y = x W^T + t B^T."""
    W = bf(d_out, d_in)
    A = [bf(d_in, 64) for _ in range(n_t)]          # stored [K, r]: b_mn
    Bl = bf(d_out, 64)
    t = [[torch.zeros(8192, 64, device=dev, dtype=torch.bfloat16) for _ in range(n_t)] for _ in range(L)]
    y = [torch.empty(8192, d_out, device=dev, dtype=torch.bfloat16) for _ in range(L)]

    def run(with_t):
        for l in range(L):
            x = producer(l)
            if with_t:
                K.gemm_grouped([dict(a=x, b=A[i], out=t[l][i], b_mn=True) for i in range(n_t)])
            K.gemm_ex(x, W, a2=t[l][0], b2=Bl, out=y[l])

    a = graph_ms(lambda: run(True))
    b = graph_ms(lambda: run(False))
    print(f"{label}: with t {a*1e3/L:7.1f} us/layer, without {b*1e3/L:7.1f} us/layer -> skinny launch costs {(a-b)*1e3/L:6.1f} us", flush=True)


def main():
    wn = torch.ones(2048, device=dev, dtype=torch.float32)
    h = [bf(8192, 2048) for _ in range(L)]
    case("t = norm(h) A, 1 problem, K=2048 (out_proj-like), main 8192x2048x2048", 2048, 2048, 1,
         lambda l: K.rmsnorm_fwd(h[l], wn))
    case("t = norm(h) A, 3 problems sharing x, K=2048 (q|k|v-like), main 8192x2048x2048", 2048, 2048, 3,
         lambda l: K.rmsnorm_fwd(h[l], wn))
    a = [bf(8192, 8192) for _ in range(L)]
    b = [bf(8192, 8192) for _ in range(L)]
    case("t = swiglu(a,b) A, K=8192 (fc3), main 8192x2048x8192", 8192, 2048, 1, lambda l: K.swiglu_fwd(a[l], b[l]))


if __name__ == "__main__":
    main()
