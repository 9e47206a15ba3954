"""Cross-entropy forward + backward in place at the bench shape, cold L2 (graph replays, memset before each call). Not a bench value."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from building_llm_from_scratch_b200 import kernels as K
dev = torch.device("cuda", 0)
rows, V = 3584, 128256
logits = (torch.randn(rows, V, device=dev) * 2).to(torch.bfloat16)
tg = torch.randint(0, V, (rows,), device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
K.cross_entropy_(logits, tg)
torch.cuda.synchronize()
ga, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
with torch.cuda.graph(ga):
    for _ in range(5):
        flush.zero_()
with torch.cuda.graph(gb):
    for _ in range(5):
        flush.zero_()
        K.cross_entropy_(logits, tg)
def run(g):
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[2]
ms = (run(gb) - run(ga)) / 5
print(f"rows/SM={os.environ.get('B200_CE_ROWS_PER_SM','4')}: {ms*1e3:.1f} us  {6*rows*V/ms/1e6:.0f} GB/s at 6 B/logit")
