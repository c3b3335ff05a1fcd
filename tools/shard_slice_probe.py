"""Tuning probe: the sharded step's kernels on ONE GPU for the slice widths a rank holds at N = 1, 2, 4, 8
(no exchange: world = 1), graph-replayed - the compute floor of the vocabulary-sharded config 5."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from duodecoding_b200.sharded import ShardedVerifier

dev = torch.device("cuda:0")
B, G = 256, 6
for n in (1, 2, 4, 8):
    Vs = 151936 // n
    sv = ShardedVerifier(Vs, rank=0, world=1)
    g = torch.Generator(device=dev).manual_seed(1)
    sets = []
    for _ in range(2):
        tl = (torch.randn(B, G + 1, Vs, device=dev, generator=g) * 3).to(torch.bfloat16)
        dl = (tl[:, :G].float() + torch.randn(B, G, Vs, device=dev, generator=g)).to(torch.bfloat16)
        toks = torch.randint(0, Vs, (B, G), device=dev, generator=g)
        r = torch.rand(B, G, device=dev, generator=g)
        noise = torch.empty(B, Vs, device=dev).exponential_(1.0, generator=g)
        sets.append((dl, tl, toks, r, noise))
    for i in range(6):
        sv.verify(*sets[i % 2], 1.0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(50):
        sv.verify(*sets[i % 2], 1.0)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 50
    print(f"slice of N={n}: Vs={Vs}: {us:.1f} us/step  ({2 * B * (G + 1) * Vs * 2 / us / 1e3:.0f} GB/s of logits)", flush=True)
    del sets, sv
    torch.cuda.empty_cache()
