"""Device-side random streams.  The only consumer on the hot path is StochasticDepthBlock
(models/layers/regularization/stochastic_depth.py:16-27): its per-sample keep factor comes from a
counter-based Philox4x32-10 generator keyed by (seed; sample, call index, step), evaluated by a
kernel (sav_drop_path_factors).  The step is read from a device counter when one is attached, so a
captured CUDA graph draws fresh masks on every replay without any host work."""
import torch

from . import ops


class DeviceRng:
    """Value of rngs['stochastic_depth']: a seed plus (optionally) the device int64 step counter the
    training step advances (TrainState.step_dev).  A plain int seed means step 0."""

    def __init__(self, seed, step_dev=None, step=0):
        self.seed, self.step_dev, self.step = int(seed), step_dev, int(step)
        self.calls = 0

    def next_factors(self, batch, keep, device):
        i = self.calls
        self.calls += 1
        return ops.drop_path_factors(self.seed, self.step_dev, self.step, i, batch, keep, device=device)


def drop_path_factors(seed, step, n_calls, batch, keep, device="cuda"):
    """The factors the first `n_calls` StochasticDepthBlock calls of one apply() draw, in execution
    order -- for feeding the same masks to a checker."""
    rng = DeviceRng(seed, None, step)
    return [rng.next_factors(batch, keep, torch.device(device)) for _ in range(n_calls)]
