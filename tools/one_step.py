"""Minimal driver for ncu captures: a few fused-verify steps at the bench shape (C2)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import duodecoding_b200.ops as ops

dev = torch.device("cuda:0")
sets = bench.make_sets(dev, 2, 1234)
top_p = float(sys.argv[1]) if len(sys.argv) > 1 else bench.TOP_P
out = None
for it in range(6):
    out = ops.verify_fused(*sets[it % 2], bench.TEMP, bench.TOP_K, top_p, out=out)
torch.cuda.synchronize()
print("accepted", out["accepted_len"].float().mean().item())
