"""Tuning probe: fused-verify step time and logit bandwidth as the batch grows (C2 shape otherwise)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import duodecoding_b200.ops as ops

dev = torch.device("cuda:0")
for B in (16, 64, 128, 256, 512):
    bench.B = B
    nsets = 4 if B <= 128 else 2
    sets = bench.make_sets(dev, nsets, 1234)
    for warp in [(1.0, 0, 0.0), (1.0, 0, 0.9)]:
        out = None
        for i in range(6):
            out = ops.verify_fused(*sets[i % nsets], *warp, out=out)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for i in range(20):
                ops.verify_fused(*sets[i % nsets], *warp, out=out)
        gr.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 20
        gbs = 2 * B * 5 * bench.V * 4 / us / 1e3
        print(f"B={B} top_p={warp[2]}: {us:.1f} us/step  {gbs:.0f} GB/s ({gbs/6520.5*100:.0f}% of peak)  {B*5/us:.2f} M tok/s", flush=True)
    del sets, out, gr
    torch.cuda.empty_cache()
