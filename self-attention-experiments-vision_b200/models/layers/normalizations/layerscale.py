"""LayerScaleBlock (models/layers/normalizations/layerscale.py:13-23): per-channel scale, init eps;
applied as a column scale in the producing GEMM's epilogue."""
from sav import functional as F
from sav import linen as nn


class LayerScaleBlock(nn.Module):
    eps: float = 1e-5
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, *unused_args, **unused_kwargs):
        scale = self.param('layerscale', nn.initializers.full(self.eps), (inputs.shape[-1],))
        return F.scale_channels(inputs, scale)
