"""Raw pinned host -> device copy rate on this box (what bounds the e2e leg of bench.py: 25.6 MB of candidates per step)."""
import torch
for mb in (6.4, 25.6, 256.0):
    n = int(mb * 1e6 / 8)
    h = torch.empty(n, dtype=torch.float64).pin_memory()
    d = torch.empty(n, dtype=torch.float64, device="cuda")
    for _ in range(3):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        d.copy_(h, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"{mb} MB: {ms:.3f} ms = {mb / ms:.1f} GB/s")
