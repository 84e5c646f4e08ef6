"""Development tool (GPU): a few eager LayerNorm backward / forward launches at the bench shape, for
ncu (`-k regex:ln_bwd_pieces`)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "self-attention-experiments-vision_b200"))
from sav import ops  # noqa: E402

M, C = 50432, 384
dev = "cuda"
x = torch.randn(M, C, device=dev)
dy = torch.randn(M, C, device=dev).bfloat16()
g = torch.randn(M, C, device=dev)
sc, bi = torch.ones(C, device=dev), torch.zeros(C, device=dev)
ds, db, dsum = (torch.zeros(C, device=dev) for _ in range(3))
for _ in range(6):
    y, mean, rstd = ops.layernorm_fwd(x, sc, bi)
    ops.layernorm_bwd(dy, x, sc, mean, rstd, ds, db, dx_add=g, dx16_colsum=dsum)
torch.cuda.synchronize()
print("ok")
