"""CvT stage block behind the reference surface (models/cvt.py:10-68): tokens -> zero-pad to a
square map -> CvT conv-projection self-attention -> + inputs -> LayerNorm -> FFBlock -> +.

Only the attention + FF block is on the hot path; the strided-conv token embedding
and the Stage / CvT wrappers around it (cvt.py:19-35,71-171) are out of scope this round, so
`create_model('cvt-*')` is not provided.  Note there is NO LayerNorm before the attention
(cvt.py:50-61): the depthwise conv reads the fp32 residual stream directly.
"""
from typing import Callable

from sav import functional as F
from sav import linen as nn

from models.layers import FFBlock
from models.layers.attentions.cvt_attention import CvTSelfAttentionBlock


def zero_pad_and_reshape(inputs):
    """cvt.py:10-16: [B, L, C] -> ([B, S, S, C] view Var, [B, S*S, C] padded Var)."""
    assert inputs.ndim == 3
    padded, S = F.pad_tokens_to_square(inputs)
    B, _, C = padded.shape
    return F.reshape(padded, (B, S, S, C)), padded


class StageBlock(nn.Module):
    num_heads: int
    embed_dim: int
    kernel_size: int = 3
    use_bias: bool = False
    activation_fn: Callable = nn.activation.gelu
    bn_momentum: float = 0.9
    bn_epsilon: float = 1e-5
    expand_ratio: float = 4
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        grid, tokens = zero_pad_and_reshape(inputs)                              # cvt.py:51
        x = CvTSelfAttentionBlock(num_heads=self.num_heads, kernel_size=self.kernel_size,
                                  use_bias=self.use_bias, bn_momentum=self.bn_momentum,
                                  bn_epsilon=self.bn_epsilon)(grid, is_training=is_training)
        x = x + tokens                                                           # cvt.py:61
        y = nn.LayerNorm()(x)                                                    # cvt.py:63
        y = FFBlock(expand_ratio=self.expand_ratio,
                    activation_fn=self.activation_fn)(y, is_training=is_training)
        return x + y                                                             # cvt.py:67
