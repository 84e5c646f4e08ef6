"""Development tool (GPU): the talking-heads row kernels (sav_thsoftmax_fwd / _bwd) at CaiT-S24's shape
(B = 128 images, 8 heads x 48, 197 tokens) alone: CUDA-event times, or the launches ncu attaches to.

    python tests/tools/th_once.py [--batch 128] [--heads 8] [--reps 10]
"""
import argparse
import json
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "self-attention-experiments-vision_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--heads", type=int, default=8)
    ap.add_argument("--dim", type=int, default=48)
    ap.add_argument("--tokens", type=int, default=197)
    ap.add_argument("--reps", type=int, default=10)
    a = ap.parse_args()
    from sav import ops
    B, H, D, N = a.batch, a.heads, a.dim, a.tokens
    g = torch.Generator(device="cuda").manual_seed(0)
    q, k, v, d_o = (torch.randn((B, N, H, D), device="cuda", generator=g).bfloat16() for _ in range(4))
    t1 = torch.eye(H, device="cuda") + 0.1 * torch.randn((H, H), device="cuda", generator=g)
    t2 = torch.eye(H, device="cuda") + 0.1 * torch.randn((H, H), device="cuda", generator=g)
    d1, d2 = torch.zeros_like(t1), torch.zeros_like(t2)
    scale = 1.0 / math.sqrt(D)
    dq, dk, dv = (torch.empty_like(q) for _ in range(3))

    def fwd():
        return ops.attn_mat_fwd(q, k, v, scale, t1, t2)

    o, S, P = fwd()

    def bwd():
        ops.attn_mat_bwd(q, k, v, S, P, d_o, dq, dk, dv, scale, t1, t2, d1, d2)

    def timed(fn):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(a.reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / a.reps * 1e3
    print("RESULT " + json.dumps({"shape": [B, H, N, D], "attn_mat_fwd_us (2 bgemm + row kernel)": timed(fwd),
                                  "attn_mat_bwd_us (4 bgemm + row kernel)": timed(bwd)}))


if __name__ == "__main__":
    main()
