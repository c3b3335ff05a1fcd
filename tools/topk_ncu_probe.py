"""One transfer_top_k rebuild at the C4 stage-2 shape ([8, 9, 32000] bf16, k = 1024) for an ncu capture."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import duodecoding_b200.ops as ops

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(5)
p = torch.softmax(torch.randn(8, 9, 32000, device=dev, generator=g) * 3, -1).to(torch.bfloat16)
for _ in range(6):
    out = ops.topk_rebuild(p, 1024, 0)
torch.cuda.synchronize()
print("ok", float(out.float().sum()))
