"""FFBlock (models/layers/feedforwards/ff.py:8-34): Dense_0 + bias -> tanh-GELU -> Dense_1 + bias.

The GELU (and the pre-activation saved for backward) lives in the first GEMM's epilogue; the bias
and the caller's residual add live in the second GEMM's epilogue.
"""
from typing import Callable

from sav import linen as nn


class FFBlock(nn.Module):
    expand_ratio: float = None
    hidden_ch: int = None
    dropout_rate: float = 0.
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        in_ch = inputs.shape[-1]
        if self.expand_ratio is None:
            if self.hidden_ch is None:
                raise ValueError('Must provide one of expand_ratio or hidden_ch')   # ff.py:18-21
            hidden_ch = self.hidden_ch
        else:
            hidden_ch = max(1, int(self.expand_ratio * in_ch))
        if self.dropout_rate and is_training:
            raise NotImplementedError("FFBlock dropout with rate > 0")
        h = nn.Dense(features=hidden_ch, use_bias=True)(inputs)
        h = self.activation_fn(h)
        return nn.Dense(features=in_ch, use_bias=True)(h)
