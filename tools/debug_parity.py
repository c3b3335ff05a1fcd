import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, oracle
import duodecoding_b200.ops as ops
dev = torch.device("cuda:0")

def make_case(seed, B, gamma, V, sigma, T, k, p):
    g = torch.Generator().manual_seed(seed)
    tl = torch.randn(B, gamma + 1, V, generator=g) * 3
    dl = tl[:, :gamma] + sigma * torch.randn(B, gamma, V, generator=g)
    q = oracle.norm_logits(dl, T, k, p)
    toks = torch.stack([oracle.sample(q[:, j], torch.empty(B, V).exponential_(1, generator=g))[:, 0] for j in range(gamma)], 1)
    r = torch.rand(B, gamma, generator=g)
    noise = torch.empty(B, V).exponential_(1, generator=g)
    return dl, tl, toks, r, noise

print("== verify V=32000 top_p")
dl, tl, toks, r, noise = make_case(22, 8, 4, 32000, 1.0, 1.0, 0, 0.9)
want = oracle.fused_verify(dl, tl, toks, r, noise, 1.0, 0, 0.9)
for team in (0, 2, 4, 8):
    ops.set_team_size(team)
    got = ops.verify_fused(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev), 1.0, 0, 0.9)
    for k in want:
        a, b = got[k].cpu(), want[k]
        if a.dtype.is_floating_point:
            rel = ((a - b).abs() / b.abs().clamp_min(1e-30)).max().item()
            print(team, k, "max rel", rel)
        else:
            print(team, k, "mismatch", int((a != b).sum()), a.flatten()[:8].tolist(), b.flatten()[:8].tolist())
ops.set_team_size(0)
print("== norm_logits")
g = torch.Generator().manual_seed(12)
for V in (32000, 151936):
    x = torch.randn(3, V, generator=g) * 3
    for T, k, p in [(1.0, 0, 0.0), (1.0, 0, 0.9), (0.8, 1024, 0.95)]:
        want = oracle.norm_logits(x, T, k, p)
        got = ops.norm_logits(x.to(dev), T, k, p).cpu()
        sup = int(((got == 0) != (want == 0)).sum())
        nz = (want != 0) & (got != 0)
        rel = ((got - want).abs()[nz] / want[nz]).max().item()
        print(V, T, k, p, "team", ops.team_size(V, 1), "support mismatches", sup, "kept", int((want != 0).sum()), int((got != 0).sum()), "max rel", rel,
              "rowsum", got.sum(-1).tolist())
print("== topk")
g = torch.Generator().manual_seed(33)
p = torch.softmax((torch.randn(5, 32000, generator=g) * 3).bfloat16().float(), -1)
for team in (1, 2, 8):
    ops.set_team_size(team)
    for k in (100, 1024):
        want = oracle.rebuild_topk_probs(p, k)
        got = ops.topk_rebuild(p.to(dev), k, 0).cpu()
        want1 = torch.stack([oracle.apply_top_k_compression(row, k) for row in p])
        got1 = ops.topk_rebuild(p.to(dev), k, 1).cpu()
        print("team", team, "k", k, "mode0 maxabs", (got - want).abs().max().item(), "mode1 mismatches", int((got1 != want1).sum()),
              "nnz", int((got1 != 0).sum(-1)[0]), int((want1 != 0).sum(-1)[0]))
ops.set_team_size(0)
pr = torch.softmax(torch.randn(4, 1000, generator=g) * 3, -1)
for k in (1, 7, 64, 999):
    want = oracle.rebuild_topk_probs(pr, k); got = ops.topk_rebuild(pr.to(dev), k, 0).cpu()
    print("k", k, "maxrel", ((got - want).abs() / want.clamp_min(1e-30)).max().item())
