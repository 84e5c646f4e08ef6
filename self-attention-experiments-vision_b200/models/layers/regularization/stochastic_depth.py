"""StochasticDepthBlock (models/layers/regularization/stochastic_depth.py:6-28): per-sample
drop-path, x/keep * floor(keep + U); the per-sample factor is a row scale in the producing GEMM."""
import torch

from sav import functional as F
from sav import linen as nn


class StochasticDepthBlock(nn.Module):
    drop_rate: float = 0.
    scale_by_keep: bool = True

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        if not is_training or self.drop_rate == 0.:
            return inputs
        if not self.scale_by_keep:
            raise NotImplementedError("scale_by_keep=False is broken in the reference (x unbound)")
        keep = 1.0 - self.drop_rate
        b = inputs.shape[0]
        u = torch.rand((b,), generator=self.make_rng('stochastic_depth'))
        factor = (torch.floor(keep + u) / keep).to(torch.float32)
        rows = 1
        for s in inputs.shape[1:-1]:
            rows *= s
        dev = inputs.x.v.device
        return F.scale_rows(inputs, factor.to(dev) if dev.type != "meta" else factor, rows)
