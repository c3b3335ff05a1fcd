import torch, time
dev = torch.device("cuda:0")
for mb in (8, 36.8, 73.7, 256):
    n = int(mb * 1e6 / 4)
    h = torch.empty(n, dtype=torch.float32).pin_memory()
    d = torch.empty(n, dtype=torch.float32, device=dev)
    for _ in range(3): d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): d.copy_(h, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print(f"H2D pinned {mb} MB: {n*4*10/e0.elapsed_time(e1)/1e6:.1f} GB/s")
    e0.record()
    for _ in range(10): h.copy_(d, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print(f"D2H pinned {mb} MB: {n*4*10/e0.elapsed_time(e1)/1e6:.1f} GB/s")
import subprocess
print(subprocess.run("nvidia-smi -q | grep -A6 'GPU Link Info'; nvidia-smi topo -m | head -5; lscpu | grep -E 'NUMA|Socket|Model name' | head", shell=True, capture_output=True, text=True).stdout)
