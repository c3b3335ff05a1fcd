"""cProfile of the reference-named call (decoding_ops.resolve_stage_verification, fused route) at the C2 shape:
where the host time of a step goes."""
import cProfile
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from duodecoding_b200 import decoding_ops, model_gpu

dev = torch.device("cuda:0")
wl = bench.Workload("c2", 64, 4, 32000, torch.float32, 1.0, 0, 0.9, 1.0, "")
dl, tl, toks, _, _ = bench.make_sets(dev, 1, 1234, wl)[0]
B, G, V, P = wl.B, wl.gamma, wl.V, 8
dh = torch.cat((dl[:, :1].expand(B, P - 1, V), dl[:, :G]), 1).contiguous()
th = torch.cat((tl[:, :1].expand(B, P - 1, V), tl[:, :G + 1]), 1).contiguous()
x = torch.cat((torch.zeros(B, P, dtype=torch.int64, device=dev), toks), 1)
dc = model_gpu.KVCacheModel(bench._NoModel(V, dev), wl.T, wl.top_k, wl.top_p)
tc = model_gpu.KVCacheModel(bench._NoModel(V, dev), wl.T, wl.top_k, wl.top_p)
dc._store_rows(dh, 0)
tc._store_rows(th, 0)
dc._past_key_values, tc._past_key_values = bench._Len(P - 1 + G), bench._Len(P + G)


def step():
    dc._current_seq_len, dc._past_key_values.n = P - 1 + G, P - 1 + G
    tc._current_seq_len, tc._past_key_values.n = P + G, P + G
    return decoding_ops.resolve_stage_verification(dc, tc, x, P, G, output_device=dev)


for _ in range(20):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(300):
    step()
torch.cuda.synchronize()
print("us per step", (time.perf_counter() - t0) / 300 * 1e6)
pr = cProfile.Profile()
pr.enable()
for _ in range(300):
    step()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(22)
