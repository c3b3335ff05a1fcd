"""Small driver for ncu: a few fused-verify launches at the C2 bench shape (no graphs)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import duodecoding_b200.ops as ops

dev = torch.device("cuda:0")
sets = bench.make_sets(dev, 2, 1234)
out = None
for i in range(6):
    out = ops.verify_fused(*sets[i % 2], bench.TEMP, bench.TOP_K, bench.TOP_P, out=out)
torch.cuda.synchronize()
print("accepted", out["accepted_len"].float().mean().item())
