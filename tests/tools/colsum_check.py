"""Development tool (GPU): two-pass column sum against torch for awkward shapes."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "self-attention-experiments-vision_b200"))
from sav import ops
for M, N in [(1, 8), (7, 24), (256, 1000), (50432, 384), (1576, 3072), (333, 2056), (12608, 768), (50432, 1152)]:
    x = torch.randn(M, N, device="cuda").bfloat16()
    ref = x.float().sum(0)
    out = ops.colsum(x)
    out2 = ops.colsum(x, out=out.clone(), accumulate=True)
    e1 = (out - ref).abs().max().item() / max(1e-6, ref.abs().max().item())
    e2 = (out2 - 2 * ref).abs().max().item() / max(1e-6, ref.abs().max().item())
    xs = torch.randn(M, N + 16, device="cuda").bfloat16()[:, 8:N + 8]   # strided view, 16 B aligned
    e3 = (ops.colsum(xs) - xs.float().sum(0)).abs().max().item()
    print(f"M={M} N={N} err {e1:.2e} accumulate {e2:.2e} strided {e3:.2e}", "ok" if max(e1, e2) < 1e-5 and e3 < 1e-2 else "FAIL")
