"""Microseconds of every kernel of the sharded top-p flow at the C5 shape on ONE GPU (slice = V / WORLD)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from duodecoding_b200.sharded import _CudaBackend

WORLD = int(os.environ.get("WORLD", "1"))
B, G, V = 256, 6, 151936
Vs = (V // WORLD + 7) // 8 * 8
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1)
tl = (torch.randn(B, G + 1, Vs, device=dev, generator=g) * 3).to(torch.bfloat16)
dl = (tl[:, :G].float() + torch.randn(B, G, Vs, device=dev, generator=g)).to(torch.bfloat16)
toks = torch.randint(0, Vs, (B, G), device=dev, generator=g)
r = torch.rand(B, G, device=dev, generator=g)
noise = torch.empty(B, Vs, device=dev).exponential_(1.0, generator=g)
be = _CudaBackend()
T, P = 1.0, 0.9


def timed(name, fn, reps=5):
    out = fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:28s} {e0.elapsed_time(e1) * 1e3 / reps:9.1f} us", flush=True)
    return out


stats = timed("stats", lambda: be.row_stats(dl, tl, toks, 0, T, T))
gstats = stats.unsqueeze(0)
levels = be.nucleus_levels(tl, T, T)
sel, quota = None, None
for level in range(levels):
    sel, hist = timed(f"level {level}", lambda: be.nucleus_level(dl, tl, G, 0, T, T, P, P, gstats, sel, level))
    saved = sel.clone()
    sel, quota = timed(f"reduce {level}", lambda: be.nucleus_reduce(hist.unsqueeze(0), 0, B, G, P, P, saved.clone(), level, levels, quota), 1)
cand = timed("tie", lambda: be.nucleus_tie(dl, tl, G, 0, T, T, sel, quota, levels))
print("rows with an index cut:", int((quota > 0).sum()), "of", quota.numel())
cuts = timed("close", lambda: be.nucleus_close(sel, gstats, cand.unsqueeze(0), B, G, levels))
out, plan = timed("accept", lambda: be.accept(gstats, B, G, r, tokens=toks, cuts=cuts))
part = timed("sample (filtered)", lambda: be.sample_partial(dl, tl, plan, noise, 0, T, T, G, filtered=True))
timed("finish", lambda: be.finish(part.unsqueeze(0), B))

# the whole step (one CUDA graph) on this slice as a world-1 problem: the compute floor of one rank
from duodecoding_b200.sharded import ShardedVerifier

sv = ShardedVerifier(Vs, rank=0, world=1)
for name, kw in (("step, no filter", {}), ("step, top_p=0.9", {"top_p": P})):
    for _ in range(3):
        sv.verify(dl, tl, toks, r, noise, T, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        sv.verify(dl, tl, toks, r, noise, T, **kw)
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:28s} {e0.elapsed_time(e1) * 1e3 / 20:9.1f} us (graph replay)", flush=True)
