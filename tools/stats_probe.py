"""Tuning probe: nucleus-search pass statistics and timing of the fused verify at the bench shape."""
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
for it in range(8):
    stats.zero_()
    out = ops.verify_fused(*sets[it % 4], bench.TEMP, bench.TOP_K, bench.TOP_P)
    torch.cuda.synchronize()
    print("  list entries: sum", int(stats[21]), "max", int(stats[22]), " us per block (148 blocks): P1 %.1f close %.1f levels %.1f select %.1f  [fine %.1f close %.1f take %.1f reduce %.1f]" % tuple(float(x) * 16 / 148 / 1965 for x in stats[24:32].tolist()))
    print("launch", it, "rows by extra L2 passes", stats[:16].tolist(), "rows", int(stats[16]), "cands", int(stats[18]))
print("passes histogram", stats[:16].tolist(), "fast", int(stats[16]), "bisect", int(stats[17]), "cands", int(stats[18]))
ops.set_stats_buffer(None)
for warp in [(1.0, 0, 0.0), (1.0, 0, 0.9)]:
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
    print(f"warp={warp}: {us:.1f} us/call  {2*64*5*32000*4/us/1e3:.0f} GB/s acc={out['accepted_len'].float().mean().item():.2f}")
