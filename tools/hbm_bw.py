import torch
x = torch.empty(1 << 30, dtype=torch.float32, device="cuda")  # 4 GB
y = torch.empty_like(x)
def t(f, n=10):
    for _ in range(2): f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
ms = t(lambda: x.fill_(1.0)); print("fill  4 GB: %.3f ms = %.0f GB/s write" % (ms, 4.295 / ms * 1e3))
ms = t(lambda: x.zero_()); print("memset 4 GB: %.3f ms = %.0f GB/s write" % (ms, 4.295 / ms * 1e3))
ms = t(lambda: x.sum()); print("sum   4 GB: %.3f ms = %.0f GB/s read" % (ms, 4.295 / ms * 1e3))
ms = t(lambda: y.copy_(x)); print("copy  4 GB: %.3f ms = %.0f GB/s read+write" % (ms, 2 * 4.295 / ms * 1e3))
