"""Pure-store bandwidth of the box (context for the several-link kernels, whose traffic is > 97 % writes):
torch fill / zero_ / copy over a buffer far larger than L2, timed with CUDA events."""
import torch

def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e-3

GB = 4
x = torch.empty(GB * (1 << 30) // 8, dtype=torch.float64, device="cuda")
y = torch.empty_like(x)
nb = x.numel() * 8
print("fill_(1.0)  %.0f GB/s written" % (nb / t(lambda: x.fill_(1.0)) / 1e9))
print("zero_()     %.0f GB/s written" % (nb / t(lambda: x.zero_()) / 1e9))
print("copy_       %.0f GB/s read+written" % (2 * nb / t(lambda: y.copy_(x)) / 1e9))
print("sum         %.0f GB/s read" % (nb / t(lambda: x.sum()) / 1e9))
