"""CvT stage block behind the reference surface (models/cvt.py:10-68): tokens -> zero-pad to a
square map -> CvT conv-projection self-attention -> + inputs -> LayerNorm -> FFBlock -> +.

The strided-conv token embedding (cvt.py:19-35) is im2col + the tcgen05 GEMM (csrc/conv2d.cu) and
its LayerNorm writes the fp32 residual stream; Stage / CvT (cvt.py:71-171) compose the blocks, the
last stage prepends the class token.  Note there is NO LayerNorm before the attention
(cvt.py:50-61): the depthwise conv reads the fp32 residual stream directly, and the head reads the
class-token row of the stream without a final LayerNorm (cvt.py:165-170).
"""
from typing import Callable, Tuple

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


class ConvTokenEmbedBlock(nn.Module):
    out_ch: int
    kernel_size: int
    strides: int
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, **unused_kwargs):
        assert len(inputs.shape) == 4                                            # cvt.py:26
        in_ch = inputs.shape[-1]
        scope = self._scope_path()
        k = self.kernel_size
        w = self._param(scope + ("Conv_0", "kernel"), nn.initializers.lecun_normal(in_axes=3),
                        (k, k, in_ch, self.out_ch))                              # cvt.py:27-31
        b = self._param(scope + ("Conv_0", "bias"), nn.initializers.zeros, (self.out_ch,))
        x = F.conv2d_stream(inputs, w, b, k, self.strides, "SAME")
        B, Ho, Wo, C = x.shape
        x = F.reshape(x, (B, Ho * Wo, C))                                        # cvt.py:32
        scale = self._param(scope + ("LayerNorm_0", "scale"), nn.initializers.ones, (C,))
        bias = self._param(scope + ("LayerNorm_0", "bias"), nn.initializers.zeros, (C,))
        return F.layernorm_stream(x, scale, bias, 1e-6)                          # cvt.py:34


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


class Stage(nn.Module):
    size: int
    num_heads: int
    embed_dim: int
    embed_kernel_size: int
    embed_strides: int
    sa_kernel_size: int = 3
    use_bias: bool = False
    activation_fn: Callable = nn.activation.gelu
    bn_momentum: float = 0.9
    bn_epsilon: float = 1e-5
    expand_ratio: float = 4
    insert_cls: bool = False
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        x = ConvTokenEmbedBlock(out_ch=self.embed_dim, kernel_size=self.embed_kernel_size,
                                strides=self.embed_strides)(inputs, is_training=is_training)
        if self.insert_cls:                                                      # cvt.py:93-98
            cls = self.param('cls', nn.initializers.zeros, (1, 1, self.embed_dim))
            x = F.concat_tokens(F.tile_param_rows(cls, x.shape[0]), x)
        for _ in range(self.size):                                               # cvt.py:100-110
            x = StageBlock(num_heads=self.num_heads, embed_dim=self.embed_dim,
                           kernel_size=self.sa_kernel_size, use_bias=self.use_bias,
                           activation_fn=self.activation_fn, bn_momentum=self.bn_momentum,
                           bn_epsilon=self.bn_epsilon,
                           expand_ratio=self.expand_ratio)(x, is_training=is_training)
        return x


class CvT(nn.Module):
    num_classes: int
    stage_sizes: Tuple[int]
    num_heads: Tuple[int]
    embed_dim: Tuple[int]
    embed_kernel_size: Tuple[int] = (7, 3, 3)
    embed_strides: Tuple[int] = (4, 2, 2)
    sa_kernel_size: Tuple[int] = (3, 3, 3)
    use_bias: bool = False
    expand_ratio: float = 4
    activation_fn: Callable = nn.activation.gelu
    bn_momentum: float = 0.9
    bn_epsilon: float = 1e-5
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        x = inputs
        n = len(self.stage_sizes)
        for i in range(n):                                                       # cvt.py:133-163
            x = Stage(size=self.stage_sizes[i], num_heads=self.num_heads[i],
                      embed_dim=self.embed_dim[i], embed_kernel_size=self.embed_kernel_size[i],
                      embed_strides=self.embed_strides[i], sa_kernel_size=self.sa_kernel_size[i],
                      use_bias=self.use_bias, activation_fn=self.activation_fn,
                      bn_momentum=self.bn_momentum, bn_epsilon=self.bn_epsilon,
                      expand_ratio=self.expand_ratio, insert_cls=(i == n - 1))(x, is_training=is_training)
            if i < n - 1:
                B, L, C = x.shape
                S = int(round(L ** 0.5))                                         # cvt.py:147-149
                assert S * S == L
                x = F.reshape(x, (B, S, S, C))
        cls_token = F.stream_first_token(x)                                      # cvt.py:165
        return nn.Dense(features=self.num_classes, use_bias=True,
                        kernel_init=nn.initializers.zeros)(cls_token)            # cvt.py:166-170
