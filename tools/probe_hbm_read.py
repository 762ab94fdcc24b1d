import torch, time
x = torch.empty(4 << 30, dtype=torch.uint8, device="cuda").view(torch.float32)
y = torch.empty_like(x)
def t(f, n=10):
    f(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); f(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best
ms = t(lambda: y.copy_(x)); print("copy GB/s", 2 * x.numel() * 4 / ms / 1e6)
ms = t(lambda: x.sum()); print("read-only sum GB/s", x.numel() * 4 / ms / 1e6)
ms = t(lambda: torch.max(x)); print("read-only max GB/s", x.numel() * 4 / ms / 1e6)
ms = t(lambda: y.fill_(1.0)); print("write-only GB/s", x.numel() * 4 / ms / 1e6)
