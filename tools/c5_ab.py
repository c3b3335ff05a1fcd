"""Tuning probe (torchrun, one rank per GPU): the vocabulary-sharded C5 step with the three exchange variants in ONE
run - NCCL all-gathers in the graph, fused peer stores per block (mode 0), fused slab copy by the last block (mode 1)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from duodecoding_b200.sharded import ShardedVerifier

world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
B, G, V = 256, 6, 151936
variants = [("nccl", False, 0), ("fused-per-block", True, 0), ("fused-slab", True, 1)]
svs = []
for name, fused, mode in variants:
    sv = ShardedVerifier(V, rank=rank, world=world, fused_exchange=fused)
    sv.xchg_mode = mode
    svs.append(sv)
Vs = svs[0].v_end - svs[0].v_begin
g = torch.Generator(device=dev).manual_seed(100 + rank)
gt = torch.Generator(device=dev).manual_seed(7)
sets = []
for _ in range(2):
    tl = (torch.randn(B, G + 1, Vs, device=dev, generator=g) * 3).to(torch.bfloat16)
    dl = (tl[:, :G].float() + torch.randn(B, G, Vs, device=dev, generator=g)).to(torch.bfloat16)
    toks = torch.randint(0, V, (B, G), device=dev, generator=gt)
    r = torch.rand(B, G, device=dev, generator=gt)
    noise = torch.empty(B, Vs, device=dev).exponential_(1.0, generator=g)
    sets.append((dl, tl, toks, r, noise))
for rep in range(3):
    for (name, _, _), sv in zip(variants, svs):
        for i in range(6):
            sv.verify(*sets[i % 2], 1.0)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(100):
            sv.verify(*sets[i % 2], 1.0)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) * 10.0], device=dev)  # us per step
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"N={world} rep {rep} {name}: {float(t):.1f} us/step", flush=True)
dist.barrier()
dist.destroy_process_group()
