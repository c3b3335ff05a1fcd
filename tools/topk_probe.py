"""Times duodec_topk_rebuild (row-select kernels vs the forced team kernel) at the C4 shapes; launches are captured
in a CUDA graph so that the host side of the op does not hide the kernel."""
import torch
import duodecoding_b200.ops as ops

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(5)
for dt in (torch.bfloat16, torch.float32):
    for V in (32000, 32004 if dt == torch.float32 else 32008):
        for rows in (8, 72, 148, 592):
            p = torch.softmax(torch.randn(rows, V, device=dev, generator=g) * 3, -1).to(dt)
            if V != 32000:
                p = p[:, 1:]  # unaligned rows: the shared-memory kernel
            for team in (0, 2):
                ops.set_team_size(team)
                for _ in range(3):
                    ops.topk_rebuild(p, 1024, 0)
                torch.cuda.synchronize()
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr):
                    for _ in range(20):
                        o = ops.topk_rebuild(p, 1024, 0)
                gr.replay()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                gr.replay()
                e1.record()
                torch.cuda.synchronize()
                us = e0.elapsed_time(e1) * 1e3 / 20
                gb = 2 * p.numel() * p.element_size() / us / 1e3
                name = ("regs" if V == 32000 else "smem") if team == 0 else "team2"
                print(f"{dt} V={p.shape[1]} rows={rows} {name}: {us:.1f} us  {gb:.0f} GB/s", flush=True)
            ops.set_team_size(0)
