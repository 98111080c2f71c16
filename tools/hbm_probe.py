"""HBM ceilings by direction on the box (torch fill / copy / reduce over 1 GiB): the write-only ceiling is well below
the copy figure in MEASURED_PEAKS.json, which is what bounds the stem and the ConvTranspose scatter."""
import torch, time
x = torch.empty(1<<30, dtype=torch.uint8, device="cuda")
y = torch.empty(1<<30, dtype=torch.uint8, device="cuda")
def t(f, n=20):
    f(); torch.cuda.synchronize()
    a,b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b)/n
ms = t(lambda: x.zero_()); print("fill 1 GiB: %.3f ms -> %.2f TB/s write" % (ms, (1<<30)/ms/1e9))
ms = t(lambda: y.copy_(x)); print("copy 1 GiB: %.3f ms -> %.2f TB/s r+w" % (ms, 2*(1<<30)/ms/1e9))
ms = t(lambda: x.view(torch.float32).sum()); print("read 1 GiB: %.3f ms -> %.2f TB/s read" % (ms, (1<<30)/ms/1e9))
