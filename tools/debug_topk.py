import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, oracle
import duodecoding_b200.ops as ops
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(33)
p = torch.softmax((torch.randn(5, 32000, generator=g) * 3).bfloat16().float(), -1)
for team in (2,):
    ops.set_team_size(team)
    for k in (100, 1024):
        want1 = torch.stack([oracle.apply_top_k_compression(row, k) for row in p])
        got1 = ops.topk_rebuild(p.to(dev), k, 1).cpu()
        for row in range(5):
            diff = (got1[row] != want1[row]).nonzero().flatten()
            kth = torch.sort(p[row], descending=True)[0][k - 1].item()
            ties = (p[row] == kth).nonzero().flatten()
            greater = int((p[row] > kth).sum())
            print("team", team, "k", k, "row", row, "kth", kth, "greater", greater, "ties", ties.tolist()[:12], "need", k - greater)
            for i in diff.tolist():
                print("    idx", i, "p", p[row, i].item(), "got", got1[row, i].item(), "want", want1[row, i].item())
