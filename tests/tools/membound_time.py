"""Development tool (GPU): the HBM-bound kernels of the block, timed alone with CUDA events over a
rotating set of buffers larger than L2 (cold) and on one buffer (warm), against the bytes they move.

    python tests/tools/membound_time.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "self-attention-experiments-vision_b200"))
from sav import ops  # noqa: E402

M, C = 50432, 384


def timeit(fns, reps=6):
    """Replays the calls from a CUDA graph: eager launches through ctypes cost 20-30 us of host time
    each, more than some of these kernels run."""
    for f in fns:
        f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for f in fns:
                f()
    g.replay()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * len(fns)) * 1e3


def report(name, us_cold, us_warm, nbytes):
    print(f"{name:22s} cold {us_cold:7.1f} us {nbytes / us_cold / 1e3:6.0f} GB/s   warm {us_warm:7.1f} us "
          f"{nbytes / us_warm / 1e3:6.0f} GB/s   ({nbytes / 1e6:.0f} MB)", flush=True)


def main():
    dev = "cuda"
    for N in (384, 1152, 1536):
        nbuf = max(2, int(400e6 // (M * N * 2)) + 1)
        xs = [torch.randn(M, N, device=dev).bfloat16() for _ in range(nbuf)]
        out = torch.zeros(N, device=dev)
        cold = timeit([(lambda x=x: ops.colsum(x, out=out, accumulate=True)) for x in xs])
        warm = timeit([lambda: ops.colsum(xs[0], out=out, accumulate=True)] * 4)
        report(f"colsum N={N}", cold, warm, M * N * 2)
        del xs
    nbuf = 4
    x = [torch.randn(M, C, device=dev) for _ in range(nbuf)]
    dy = [torch.randn(M, C, device=dev).bfloat16() for _ in range(nbuf)]
    g = [torch.randn(M, C, device=dev) for _ in range(nbuf)]
    dx = [torch.empty(M, C, device=dev) for _ in range(nbuf)]
    dx16 = [torch.empty(M, C, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
    scale, bias = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    dscale, dbias = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
    y, mean, rstd = ops.layernorm_fwd(x[0], scale, bias)
    cold = timeit([(lambda i=i: ops.layernorm_fwd(x[i], scale, bias)) for i in range(nbuf)])
    warm = timeit([lambda: ops.layernorm_fwd(x[0], scale, bias)] * 4)
    report("ln_fwd", cold, warm, M * C * 6)
    cold = timeit([(lambda i=i: ops.layernorm_bwd(dy[i], x[i], scale, mean, rstd, dscale, dbias, dx_add=g[i],
                                                  dx=dx[i], dx16=dx16[i])) for i in range(nbuf)])
    warm = timeit([lambda: ops.layernorm_bwd(dy[0], x[0], scale, mean, rstd, dscale, dbias, dx_add=g[0],
                                             dx=dx[0], dx16=dx16[0])] * 4)
    report("ln_bwd (+stream grad)", cold, warm, M * C * 16)
    dsum = torch.zeros(C, device=dev)
    cold = timeit([(lambda i=i: ops.layernorm_bwd(dy[i], x[i], scale, mean, rstd, dscale, dbias, dx_add=g[i],
                                                  dx=dx[i], dx16=dx16[i], dx16_colsum=dsum)) for i in range(nbuf)])
    warm = timeit([lambda: ops.layernorm_bwd(dy[0], x[0], scale, mean, rstd, dscale, dbias, dx_add=g[0],
                                             dx=dx[0], dx16=dx16[0], dx16_colsum=dsum)] * 4)
    report("ln_bwd + bias colsum", cold, warm, M * C * 16)


if __name__ == "__main__":
    main()
