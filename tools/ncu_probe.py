"""Small driver for ncu: a few fused-verify launches at the C2 shape (no graphs)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import duodecoding_b200.ops as ops

dev = torch.device("cuda:0")
B, gamma, V = 64, 4, 32000
team = int(sys.argv[1]) if len(sys.argv) > 1 else 0
g = torch.Generator(device=dev).manual_seed(1234)
tl = torch.randn(B, gamma + 1, V, device=dev, generator=g) * 3
dl = tl + 0.7 * torch.randn(B, gamma + 1, V, device=dev, generator=g)
q = ops.norm_logits(dl, 1.0, 0, 0.9)
toks = torch.stack([ops.sample_rows(q[:, j])[:, 0] for j in range(gamma)], 1)
r = torch.rand(B, gamma, device=dev, generator=g)
noise = torch.empty(B, V, device=dev).exponential_(1.0, generator=g)
ops.set_team_size(team)
for warp in [(1.0, 0, 0.0), (1.0, 0, 0.9)]:
    for i in range(3):
        out = ops.verify_fused(dl, tl, toks, r, noise, *warp)
    torch.cuda.synchronize()
    print(warp, "accepted", out["accepted_len"].float().mean().item())
