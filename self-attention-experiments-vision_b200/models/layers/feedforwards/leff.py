"""LeFFBlock (models/layers/feedforwards/leff.py:8-63), CeiT's locally-enhanced feed-forward: the
patch tokens go Dense -> BatchNorm -> GELU -> (back on the 14x14 grid) Conv kxk -> BatchNorm -> GELU
-> Dense -> BatchNorm -> GELU; the class token bypasses it.  BatchNorm + GELU are one fused kernel
pair; the convolution is im2col + tcgen05 GEMM with the bias in the epilogue.  Returns the patch
tokens only: the caller's `x + concat([cls, tokens])` is sav.functional.leff_residual."""
import math
from typing import Callable

from sav import functional as F
from sav import linen as nn


class LeFFBlock(nn.Module):
    expand_ratio: float = None
    hidden_ch: int = None
    kernel_size: int = 5
    activation_fn: Callable = nn.activation.gelu
    bn_momentum: float = 0.9
    bn_epsilon: float = 1e-5
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        tokens = F.stream_tokens_operand(inputs)                  # leff.py:22: cls, tokens = x[:, 0], x[:, 1:]
        b, l, in_ch = tokens.shape
        if self.expand_ratio is None:
            if self.hidden_ch is None:
                raise ValueError('Must provide one of expand_ratio or hidden_ch')   # leff.py:25-27
            hidden_ch = self.hidden_ch
        else:
            hidden_ch = max(1, int(self.expand_ratio * in_ch))

        def batch_norm(x):
            return nn.BatchNorm(use_running_average=not is_training, momentum=self.bn_momentum,
                                epsilon=self.bn_epsilon)(x, act=self.activation_fn)

        x = nn.Dense(features=hidden_ch, use_bias=True)(tokens)                     # leff.py:41
        x = batch_norm(x)                                                           # :42-43
        side = int(math.sqrt(l))                                                    # :45
        x = F.reshape(x, (b, side, side, hidden_ch))                                # :46
        x = nn.Conv(features=hidden_ch, kernel_size=(self.kernel_size, self.kernel_size),
                    padding='SAME', use_bias=True)(x)                               # :48-53
        x = batch_norm(x)                                                           # :54-55
        x = F.reshape(x, (b, l, hidden_ch))                                         # :57
        x = nn.Dense(features=in_ch, use_bias=True)(x)                              # :59
        return batch_norm(x)                                                        # :60-61
