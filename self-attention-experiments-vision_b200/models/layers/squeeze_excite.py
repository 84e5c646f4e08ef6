"""SqueezeExciteBlock (models/layers/squeeze_excite.py:13-38): global mean over (H, W) -> Dense(hidden)
-> activation -> Dense(C) -> sigmoid -> scale the input per (sample, channel).

The reference calls the result of jnp.mean as a function (squeeze_excite.py:32-33); the
intended x = mean(inputs, (1, 2), keepdims=True) is built.  `gate_only=True` returns the [B, C] sigmoid gate so
that a bottleneck block can fold gate * y + residual + activation into one pass
(functional.gate_residual_act); calling the module applies the gate (squeeze_excite.py:37).
"""
from typing import Callable

from sav import functional as F
from sav import linen as nn


class SqueezeExciteBlock(nn.Module):
    se_ratio: float = None
    hidden_ch: int = None
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    def _hidden(self, in_ch):
        if self.se_ratio is None:
            if self.hidden_ch is None:
                raise ValueError('Must provide one of se_ratio or hidden_ch')        # squeeze_excite.py:24
            return self.hidden_ch
        return max(1, int(in_ch * self.se_ratio))                                     # squeeze_excite.py:28

    def _activate(self, x):
        kind = getattr(self.activation_fn, "kind", None)
        if kind is not None:
            return F.activation(x, kind)
        return self.activation_fn(x)          # gelu: fused into the producing Dense's epilogue

    @nn.compact
    def __call__(self, inputs, gate_only=False):
        in_ch = inputs.shape[-1]
        hidden = self._hidden(in_ch)
        x = F.spatial_mean(inputs)                                                    # squeeze_excite.py:32
        x = nn.Dense(features=hidden, use_bias=True)(x)
        x = self._activate(x)
        x = nn.Dense(features=in_ch, use_bias=True)(x)
        gate = F.activation(x, "sigmoid")
        if gate_only:
            return gate
        return F.gate_residual_act(inputs, gate, None, "none")                        # squeeze_excite.py:37
