"""HBM bandwidth by access mix (torch ops as the probe, not product code): write-only (fill_), copy (1 read : 1 write), read-only (sum)."""
import torch

def t(fn, iters=20):
    fn(); fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3

n = 1 << 30                      # 1 GiB buffers (>> 126 MB L2)
a = torch.empty(n, dtype=torch.uint8, device="cuda").view(torch.float32)
b = torch.empty(n, dtype=torch.uint8, device="cuda").view(torch.float32)
print(f"write-only fill_ : {n / t(lambda: a.fill_(1.0)) / 1e9:8.1f} GB/s")
print(f"write-only zero_ : {n / t(lambda: a.zero_()) / 1e9:8.1f} GB/s")
print(f"copy (r+w bytes) : {2 * n / t(lambda: b.copy_(a)) / 1e9:8.1f} GB/s")
print(f"read-only sum    : {n / t(lambda: a.sum()) / 1e9:8.1f} GB/s")
ab = a.view(torch.bfloat16)
print(f"read-only bf16 max: {n / t(lambda: ab.max()) / 1e9:8.1f} GB/s")
