import sys, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
torch.cuda.set_device(0)
for r in bench.kernel_rooflines(torch.device("cuda", 0), bench.load_peaks()):
    print(r)
