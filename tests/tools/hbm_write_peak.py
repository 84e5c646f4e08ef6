"""Development tool (GPU): write-only, read-only and copy HBM bandwidth with plain torch ops, to
put the GEMM epilogues' write-dominated traffic (FC1 forward: 310 MB written, 39 MB read) next
to the copy peak that MEASURED_PEAKS.json quotes."""
import torch

def t(fn, n=20):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e-3

N = 1 << 30
a = torch.empty(N, dtype=torch.uint8, device="cuda")
b = torch.empty(N, dtype=torch.uint8, device="cuda")
af = a.view(torch.float32)
print("write-only (zero_)  %.0f GB/s" % (N / t(lambda: a.zero_()) / 1e9))
print("copy (read+write)   %.0f GB/s" % (2 * N / t(lambda: b.copy_(a)) / 1e9))
print("read-only (sum)     %.0f GB/s" % (N / t(lambda: af.sum()) / 1e9))
m = 310 << 20
print("write 310 MB        %.0f GB/s" % (m / t(lambda: a[:m].zero_()) / 1e9))
