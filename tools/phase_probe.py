"""Tuning probe: per-phase timing (both kernels / scan only / sample only) of the fused verify at C2, graph-replayed."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import time
import torch

import bench
import duodecoding_b200.ops as ops

dev = torch.device("cuda:0")
sets = bench.make_sets(dev, 4, 1234)
out = None
warps = [(1.0, 0, 0.9)] if len(sys.argv) < 2 else [(1.0, 0, float(a)) for a in sys.argv[1:]]
for warp in warps:
    for mask in (3, 1, 2):
        ops.set_phase_mask(mask)
        for i in range(5):
            out = ops.verify_fused(*sets[i % 4], *warp, out=out)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for i in range(100):
                ops.verify_fused(*sets[i % 4], *warp, out=out)
        t_end = time.perf_counter() + 0.3
        while time.perf_counter() < t_end:
            gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 100
        print(f"warp={warp} phases={mask}: {us:.1f} us/call", flush=True)
    ops.set_phase_mask(3)
