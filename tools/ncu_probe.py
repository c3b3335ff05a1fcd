"""Small driver for ncu: a few fused-verify launches at the C2 bench shape (no graphs).  argv[1]: top_p (default bench)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import duodecoding_b200.ops as ops

top_p = float(sys.argv[1]) if len(sys.argv) > 1 else bench.TOP_P
dev = torch.device("cuda:0")
sets = bench.make_sets(dev, 2, 1234)
out = None
for i in range(6):
    out = ops.verify_fused(*sets[i % 2], bench.TEMP, bench.TOP_K, top_p, out=out)
torch.cuda.synchronize()
print("accepted", out["accepted_len"].float().mean().item())
