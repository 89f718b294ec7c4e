"""HBM bandwidth by access mix on this GPU: pure write (fill), pure read (sum), copy (read + write), and a 1:2
read:write mix -- the mix of the forward contractions' output-dominated traffic.  torch kernels, CUDA events."""
import torch

dev = torch.device("cuda", 0)
n = 1 << 29                                     # 2 GiB of fp32 per buffer (>> 126 MB L2)
a = torch.empty(n, device=dev)
b = torch.empty(n, device=dev)
half = torch.empty(n // 2, device=dev)


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e-3


gb = n * 4 / 1e9
print(f"fill  (write only)        {gb / timed(lambda: a.fill_(1.0)):8.0f} GB/s")
print(f"sum   (read only)         {gb / timed(lambda: a.sum()):8.0f} GB/s")
print(f"copy  (1 read : 1 write)  {2 * gb / timed(lambda: b.copy_(a)):8.0f} GB/s")
print(f"add   (2 reads : 1 write) {3 * gb / timed(lambda: torch.add(a, b, out=a)):8.0f} GB/s")
# 1 read : 2 writes -- repeat_interleave-free form: write two outputs from one half-size input
av, bv = a[: n // 2], b[: n // 2]
def one_to_two():
    torch.mul(half, 2.0, out=av)
    # second write of the same input would re-read it; emulate with a fused form instead:
print(f"mul   (1 read : 1 write, half size) {gb / timed(lambda: torch.mul(half, 2.0, out=av)):8.0f} GB/s")
