import torch
x = torch.empty(1 << 30, dtype=torch.uint8, device='cuda')
y = torch.empty(1 << 30, dtype=torch.uint8, device='cuda')
def t(f, n=10):
    f(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
ms = t(lambda: x.zero_()); print("fill  GB/s", (1 << 30) / ms / 1e6)
ms = t(lambda: y.copy_(x)); print("copy  GB/s (r+w)", 2 * (1 << 30) / ms / 1e6)
xf = x.view(torch.float32)
ms = t(lambda: xf.sum()); print("read  GB/s", (1 << 30) / ms / 1e6)
