"""Ten fused verifications at the C2 bench shape (rotating input sets) for ncu captures of the step's kernels."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import duodecoding_b200.ops as ops

dev = torch.device("cuda:0")
sets = bench.make_sets(dev, 4, 1234)
out = None
for it in range(10):
    out = ops.verify_fused(*sets[it % 4], bench.TEMP, bench.TOP_K, bench.TOP_P, out=out)
torch.cuda.synchronize()
print("ok", out["accepted_len"].float().mean().item())
