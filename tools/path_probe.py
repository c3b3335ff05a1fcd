"""Tuning probe: time the fused verify per device path (1 split kernels, 2 persistent row-stream) at the bench shape."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import duodecoding_b200.ops as ops
from duodecoding_b200 import _lib

dev = torch.device("cuda:0")
sets = bench.make_sets(dev, 4, 1234)
paths = [int(a) for a in sys.argv[1:]] or [1, 2]
for path in paths:
    _lib.lib.duodec_set_stream_path(path)
    for warp in [(1.0, 0, 0.0), (1.0, 0, 0.9)]:
        out = None
        for i in range(8):
            out = ops.verify_fused(*sets[i % 4], *warp, out=out)
        torch.cuda.synchronize()
        if path == 2:
            stats = torch.zeros(32, dtype=torch.int32, device=dev)
            ops.set_stats_buffer(stats)
            ops.verify_fused(*sets[0], *warp, out=out)
            torch.cuda.synchronize()
            ops.set_stats_buffer(None)
            print("  path2 stats passes", stats[:6].tolist(), "cycles/16: pass, reduce, resolve, publish, close", stats[20:25].tolist())
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for i in range(40):
                ops.verify_fused(*sets[i % 4], *warp, out=out)
        gr.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 40
        print(f"path={path} warp={warp}: {us:.1f} us/call  {2*64*5*32000*4/us/1e3:.0f} GB/s acc={out['accepted_len'].float().mean().item():.2f}", flush=True)
_lib.lib.duodec_set_stream_path(1)
