"""Tuning probe: row-capture kernel statistics and timing of the fused verify at the bench shape.
usage: stats_probe.py [path]   (6 row capture, 5 row pairs, 0 general)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import duodecoding_b200.ops as ops

dev = torch.device("cuda:0")
if len(sys.argv) > 1:
    ops.set_stream_path(int(sys.argv[1]))
sets = bench.make_sets(dev, 4, 1234)
stats = torch.zeros(32, dtype=torch.int32, device=dev)
ops.set_stats_buffer(stats)
for it in range(4):
    stats.zero_()
    out = ops.verify_fused(*sets[it % 4], bench.TEMP, bench.TOP_K, bench.TOP_P)
    torch.cuda.synchronize()
    s = stats.tolist()
    print(f"launch {it}: fast-select rows {s[16]}, handed to the general search {s[17]}, general searches {s[19]}, "
          f"acc {out['accepted_len'].float().mean().item():.2f}", flush=True)
ops.set_stats_buffer(None)
for warp in [(1.0, 0, 0.0), (1.0, 0, 0.9)]:
    for mask in (3, 1, 2):
        ops.set_phase_mask(mask)
        for i in range(5):
            out = ops.verify_fused(*sets[i % 4], *warp, out=out)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for i in range(40):
                ops.verify_fused(*sets[i % 4], *warp, out=out)
        gr.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 40
        print(f"warp={warp} phases={mask}: {us:.1f} us/call  {2*64*5*32000*4/us/1e3:.0f} GB/s", flush=True)
    ops.set_phase_mask(3)
