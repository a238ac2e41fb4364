"""Pure-write / pure-read / copy HBM bandwidth probes (context for the rooflines; torch ops only)."""
import torch
x = torch.empty(1 << 30, dtype=torch.float32, device="cuda")  # 4 GiB
y = torch.empty(1 << 30, dtype=torch.float32, device="cuda")
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e-3
print("fill  (write only) GB/s", 4.294967296 / t(lambda: x.fill_(1.0)))
print("sum   (read only)  GB/s", 4.294967296 / t(lambda: x.sum()))
print("copy  (r+w)        GB/s", 2 * 4.294967296 / t(lambda: y.copy_(x)))
