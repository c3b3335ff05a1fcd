"""Tuning probe (build with -DDUODEC_TIMELINE via tools/build_variant.sh, run with DUODEC_LIB=...): global-timer
stamps of one fused verify at the bench shape, us after the scan kernel's first block."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import duodecoding_b200.ops as ops

dev = torch.device("cuda:0")
sets = bench.make_sets(dev, 4, 1234)
stats = torch.zeros(32, dtype=torch.int32, device=dev)
out = None
for it in range(6):
    out = ops.verify_fused(*sets[it % 4], bench.TEMP, bench.TOP_K, bench.TOP_P, out=out)
torch.cuda.synchronize()
ops.set_stats_buffer(stats)
names = {9: "last pass-1 end", 10: "last pass-2 end", 25: "  select: level histogram", 26: "  select: scan", 27: "  select: candidates collected", 28: "  select: ordered", 11: "last row closed", 13: "last sample block past the wait",
         14: "last n known (row copies issued)", 24: "last row slices landed", 15: "last slice reduced",
         23: "last sample block end"}
for it in range(4):
    stats.zero_()
    out = ops.verify_fused(*sets[it % 4], bench.TEMP, bench.TOP_K, bench.TOP_P, out=out)
    torch.cuda.synchronize()
    s = stats.tolist()
    t0 = 0x7fffffff - s[8]
    print("launch", it, "| first sample block resident at %.1f us" % ((0x7fffffff - s[12] - t0) / 1e3))
    for k, nm in names.items():
        print("   %-34s %.1f us" % (nm, (s[k] - t0) / 1e3))
ops.set_stats_buffer(None)
