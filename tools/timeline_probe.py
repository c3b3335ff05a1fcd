"""Tuning probe (build with -DDUODEC_TIMELINE): global-timer stamps of one fused verify, us after the scan kernel's first block."""
import os, sys
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
for it in range(4):
    stats.zero_()
    out = ops.verify_fused(*sets[it % 4], bench.TEMP, bench.TOP_K, bench.TOP_P, out=out)
    torch.cuda.synchronize()
    s = stats.tolist()
    t0 = 0x7fffffff - s[8]
    names = {9: "last stream end", 25: "  select: value range", 26: "  select: level-0 histogram", 27: "  select: last scan", 28: "  select: last level>0 loop", 29: "  select: candidates collected", 10: "last select end", 11: "last row closed", 13: "last sample block past the wait",
             14: "last n known", 24: "compact: list entries issued", 15: "compact: gathers consumed", 23: "last sample block end"}
    print("launch pair", it, "| first sample block resident at %.1f us" % ((0x7fffffff - s[12] - t0) / 1e3))
    for k, nm in names.items():
        print("   %-34s %.1f us" % (nm, (s[k] - t0) / 1e3))
    print("   sampler cycles from the grid dependency (max over blocks): n known %d, bit maps %d, list values %d, "
          "gathers done %d, block end %d, finalised %d" % tuple(s[1:7]))
ops.set_stats_buffer(None)
