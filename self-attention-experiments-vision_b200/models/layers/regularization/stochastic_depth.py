"""StochasticDepthBlock (models/layers/regularization/stochastic_depth.py:6-28): per-sample
drop-path, x/keep * floor(keep + U); the per-sample factor is drawn by a device Philox kernel
(sav/rng.py) and applied as a row scale in the producing GEMM's epilogue."""
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
        rows = 1
        for s in inputs.shape[1:-1]:
            rows *= s
        dev = inputs.x.v.device
        rng = self.make_device_rng('stochastic_depth')   # KeyError without the stream, like Flax
        if dev.type == "meta":
            return F.scale_rows(inputs, torch.empty((b,), dtype=torch.float32, device="meta"), rows)
        # per-sample floor(keep + U) / keep drawn ON THE DEVICE (Philox keyed by seed, call index and
        # the device step counter): graph replays resample, nothing is copied from the host
        factor = rng.next_factors(b, keep, dev)
        return F.scale_rows(inputs, factor, rows)
