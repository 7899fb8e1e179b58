import sys, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
print(bench.k1_roofline(torch, torch.device("cuda", 0), 1 << 20, iters=3, warmup=2))
