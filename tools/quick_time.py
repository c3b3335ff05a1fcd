"""Development timing probe (not the bench): fused verify at the C2 shape for several warps / team sizes."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import duodecoding_b200.ops as ops

dev = torch.device("cuda:0")
B, gamma, V = 64, 4, 32000
NSETS = 4
g = torch.Generator(device=dev).manual_seed(1234)
sets = []
for s in range(NSETS):
    tl = torch.randn(B, gamma + 1, V, device=dev, generator=g) * 3
    dl = tl + 0.7 * torch.randn(B, gamma + 1, V, device=dev, generator=g)
    toks = torch.randint(0, V, (B, gamma), device=dev, generator=g)
    r = torch.rand(B, gamma, device=dev, generator=g)
    noise = torch.empty(B, V, device=dev).exponential_(1.0, generator=g)
    sets.append((dl, tl, toks, r, noise))
out = None
for warp in [(1.0, 0, 0.0), (1.0, 0, 0.9), (1.0, 50, 0.0), (0.0, 0, 0.0)]:
    for team in [0, 1, 2, 4, 8]:
        ops.set_team_size(team)
        try:
            for i in range(5):
                out = ops.verify_fused(*sets[i % NSETS], *warp)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n = 40
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                for i in range(NSETS):
                    ops.verify_fused(*sets[i % NSETS], *warp, out=out)
            gr.replay(); torch.cuda.synchronize()
            e0.record()
            for i in range(n // NSETS):
                gr.replay()
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / n
            byt = 2 * B * (gamma + 1) * V * 4
            print(f"warp={warp} team={team} auto->{ops.team_size(V,2)}: {us:8.1f} us/call  {byt/us/1e3:8.1f} GB/s  acc={out['accepted_len'].float().mean().item():.2f}", flush=True)
        except Exception as ex:
            print(f"warp={warp} team={team}: ERR {ex}", flush=True)
ops.set_team_size(0)
x = sets[0][1].reshape(-1, V)
for warp in [(1.0, 0, 0.0), (1.0, 0, 0.9)]:
    for i in range(3):
        y = ops.norm_logits(x, *warp)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(20):
        ops.norm_logits(sets[i % NSETS][1].reshape(-1, V), *warp, out=y)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 20
    print(f"norm_logits warp={warp}: {us:.1f} us  {2*x.numel()*4/us/1e3:.1f} GB/s")
