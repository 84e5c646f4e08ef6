"""CeiT behind the reference surface (models/ceit.py:10-156): convolutional Image2Token stem, POST-LN
encoder blocks (x = LN(SA(in) + in); out = LN(x + LeFF(x))) whose LayerNorm outputs ARE the residual
stream, the class token of every block collected, and a layer-wise class attention over those 12
tokens (query = the last one) in front of the head.

On this path the stream is fp32 (`nn.LayerNorm(stream_out=True)`), the attention reads a bf16 copy of
it (`F.stream_to_operand`), and `x + LeFF(x)` -- class token passed through and added to itself,
patch tokens through the conv branch -- is `F.leff_residual`.  Flax auto-names: Image2TokenBlock_0,
cls, Encoder_0/EncoderBlock_i/{SelfAttentionBlock_0, LayerNorm_0, LeFFBlock_0, LayerNorm_1},
LCSelfAttentionBlock_0, Dense_0.  (LCAEncoderBlock, ceit.py:73-93, is defined by the reference but
never used by CeiT.__call__; it is offered for completeness.)
"""
from typing import Callable, Optional, Tuple

from sav import functional as F
from sav import linen as nn

from models.layers import AttentionBlock, FFBlock, Image2TokenBlock, LeFFBlock, SelfAttentionBlock


class LCSelfAttentionBlock(AttentionBlock):

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        inputs_q = F.last_token(inputs)                                          # ceit.py:14
        return super().__call__(inputs_q, inputs, is_training=is_training)


class EncoderBlock(nn.Module):
    num_heads: int = 3
    expand_ratio: float = 4
    leff_kernel_size: Optional[int] = 3
    activation_fn: Callable = nn.activation.gelu
    bn_momentum: float = 0.9
    bn_epsilon: float = 1e-5
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        x = SelfAttentionBlock(num_heads=self.num_heads)(F.stream_to_operand(inputs, inputs.shape),
                                                         is_training=is_training)   # ceit.py:29-31
        x = x + inputs                                                           # :32, out-proj epilogue
        x = nn.LayerNorm(stream_out=True)(x)                                     # :33
        y = LeFFBlock(expand_ratio=self.expand_ratio, kernel_size=self.leff_kernel_size,
                      activation_fn=self.activation_fn, bn_momentum=self.bn_momentum,
                      bn_epsilon=self.bn_epsilon)(x, is_training=is_training)    # :35-40
        y = F.leff_residual(x, y)                                                # :41 y = x + y
        return nn.LayerNorm(stream_out=True)(y)                                  # :42


class Encoder(nn.Module):
    num_layers: int = 12
    num_heads: int = 3
    expand_ratio: float = 4
    leff_kernel_size: int = 3
    activation_fn: Callable = nn.activation.gelu
    bn_momentum: float = 0.9
    bn_epsilon: float = 1e-5
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        x = inputs
        cls_tokens_lst = []
        for _ in range(self.num_layers):                                         # ceit.py:64-70
            x = EncoderBlock(num_heads=self.num_heads, expand_ratio=self.expand_ratio,
                             leff_kernel_size=self.leff_kernel_size, activation_fn=self.activation_fn,
                             bn_momentum=self.bn_momentum,
                             bn_epsilon=self.bn_epsilon)(x, is_training=is_training)
            cls_tokens_lst.append(F.stream_first_token(x))                       # x[:, 0]
        return F.stack_tokens(cls_tokens_lst)                                    # :72


class LCAEncoderBlock(nn.Module):
    num_heads: int = 3
    expand_ratio: float = 4
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        raise NotImplementedError("LCAEncoderBlock is never called by the reference's CeiT "
                                  "(ceit.py:140-147 applies LCSelfAttentionBlock directly); its "
                                  "`x += inputs` adds [B,1,C] to [B,12,C] by broadcasting")


class CeiT(nn.Module):
    num_classes: int = 1000
    num_layers: int = 12
    num_heads: int = 3
    embed_dim: int = 192
    patch_shape: Tuple[int] = (4, 4)
    num_ch: int = 32
    conv_kernel_size: int = 7
    conv_stride: int = 2
    pool_window_size: int = 3
    pool_stride: int = 2
    expand_ratio: float = 4
    leff_kernel_size: int = 3
    bn_momentum: float = 0.9
    bn_epsilon: float = 1e-5
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        assert self.embed_dim % self.num_heads == 0                              # ceit.py:115
        x = Image2TokenBlock(patch_shape=self.patch_shape, num_ch=self.num_ch,
                             conv_kernel_size=self.conv_kernel_size, conv_stride=self.conv_stride,
                             pool_window_size=self.pool_window_size, pool_stride=self.pool_stride,
                             embed_dim=self.embed_dim, bn_momentum=self.bn_momentum,
                             bn_epsilon=self.bn_epsilon)(inputs, is_training=is_training)   # :117-126
        x = F.to_stream(x)                                     # the token embedding starts the stream
        b = x.shape[0]
        cls_param = self.param('cls', nn.initializers.zeros, (1, 1, self.embed_dim))       # :130
        x = F.concat_tokens(F.tile_param_rows(cls_param, b), x)                            # :131-133
        cls_tokens = Encoder(num_layers=self.num_layers, num_heads=self.num_heads,
                             expand_ratio=self.expand_ratio, leff_kernel_size=self.leff_kernel_size,
                             bn_momentum=self.bn_momentum,
                             bn_epsilon=self.bn_epsilon)(x, is_training=is_training)       # :135-141
        cls_tokens = LCSelfAttentionBlock(num_heads=self.num_heads)(cls_tokens,
                                                                    is_training=is_training)  # :143-147
        b, n, c = cls_tokens.shape
        cls = F.reshape(cls_tokens, (b, c))                                      # :148 cls_tokens[:, -1]
        return nn.Dense(features=self.num_classes, use_bias=True,
                        kernel_init=nn.initializers.zeros)(cls)                  # :149-153
