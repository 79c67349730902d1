"""Pinned host->device copy bandwidth by transfer size (context for the end-to-end number, which is PCIe-bound)."""
import json, torch
dev = torch.device("cuda", 0)
out = {}
for mb in (8, 24, 64, 256, 1024):
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device=dev)
    for _ in range(3):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = max(4, 2048 // mb)
    e0.record()
    for _ in range(reps):
        d.copy_(h, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    out[f"{mb}MiB"] = round(n * reps / e0.elapsed_time(e1) / 1e6, 2)
print(json.dumps({"h2d_pinned_GBps": out}))
