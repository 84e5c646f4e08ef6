"""CaiT behind the reference surface (models/cait.py:10-183): talking-heads self-attention blocks
with LayerScale + stochastic depth over the 196 patch tokens (no class token), then class-attention
blocks in which only the class token queries (Nq = 1), final LayerNorm over [cls; patches], head.

Kernel mapping: LayerScale and the per-sample drop-path factor live in the out-proj / Dense_1 GEMM
epilogues together with the residual add; talking-heads attention (head dim 48 for CaiT-S) and
class attention run through the materialised-logits path (sav_bgemm + sav_thsoftmax); the
[cls; patches] concat is an fp32 stream copy.  Parameter names follow Flax auto-naming:
PatchEmbedBlock_0, Encoder_0/{AddAbsPosEmbed_0, EncoderBlock_i/{LayerNorm_0, SelfAttentionBlock_0/
{queries,keys,values,DenseGeneral_0,TalkingHeadsBlock_0,TalkingHeadsBlock_1}, LayerScaleBlock_0,
LayerNorm_1, FFBlock_0, LayerScaleBlock_1}}, cls, CAEncoderBlock_i/{LayerNorm_0,
ClassSelfAttentionBlock_0, LayerScaleBlock_0, LayerNorm_1, FFBlock_0, LayerScaleBlock_1},
LayerNorm_0, Dense_0.  (The reference builds Encoder without dtype, cait.py:147-154 -- SURVEY D10;
here every block computes in bf16 storage / fp32 accumulation like the rest of the path.)
"""
from typing import Callable, Tuple

from sav import functional as F
from sav import linen as nn

from models.layers import (AddAbsPosEmbed, AttentionBlock, FFBlock, LayerScaleBlock,
                           PatchEmbedBlock, SelfAttentionBlock, StochasticDepthBlock)


class ClassSelfAttentionBlock(AttentionBlock):

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        inputs_q = F.first_token_keepdim(inputs)                               # cait.py:13
        return super().__call__(inputs_q, inputs, is_training=is_training)


class EncoderBlock(nn.Module):
    num_heads: int
    stoch_depth_rate: float
    layerscale_eps: float
    expand_ratio: float = 4
    attn_dropout_rate: float = 0.
    dropout_rate: float = 0.
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        x = nn.LayerNorm()(inputs)                                              # cait.py:30
        x = SelfAttentionBlock(num_heads=self.num_heads, talking_heads=True,
                               attn_dropout_rate=self.attn_dropout_rate,
                               out_dropout_rate=self.dropout_rate)(x, is_training=is_training)
        x = LayerScaleBlock(eps=self.layerscale_eps)(x, is_training=is_training)
        x = StochasticDepthBlock(drop_rate=self.stoch_depth_rate)(x, is_training=is_training)
        x = x + inputs                                                          # cait.py:40
        y = nn.LayerNorm()(x)
        y = FFBlock(expand_ratio=self.expand_ratio, dropout_rate=self.dropout_rate,
                    activation_fn=self.activation_fn)(y, is_training=is_training)
        y = LayerScaleBlock(eps=self.layerscale_eps)(y, is_training=is_training)
        y = StochasticDepthBlock(drop_rate=self.stoch_depth_rate)(y, is_training=is_training)
        return x + y                                                            # cait.py:52


class Encoder(nn.Module):
    num_layers: int
    num_heads: int
    stoch_depth_rate: float
    layerscale_eps: float
    expand_ratio: float = 4
    attn_dropout_rate: float = 0.
    dropout_rate: float = 0.
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        x = AddAbsPosEmbed()(inputs)                                            # cait.py:69
        x = nn.Dropout(rate=self.dropout_rate)(x, deterministic=not is_training)
        for _ in range(self.num_layers):
            x = EncoderBlock(num_heads=self.num_heads, expand_ratio=self.expand_ratio,
                             attn_dropout_rate=self.attn_dropout_rate,
                             dropout_rate=self.dropout_rate,
                             stoch_depth_rate=self.stoch_depth_rate,
                             layerscale_eps=self.layerscale_eps,
                             activation_fn=self.activation_fn)(x, is_training=is_training)
        return x


class CAEncoderBlock(nn.Module):
    num_heads: int
    stoch_depth_rate: float
    layerscale_eps: float
    expand_ratio: float = 4
    attn_dropout_rate: float = 0.
    dropout_rate: float = 0.
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, cls_token, is_training: bool):
        x = F.concat_tokens(cls_token, inputs)                                  # cait.py:97
        x = nn.LayerNorm()(x)
        x = ClassSelfAttentionBlock(num_heads=self.num_heads,
                                    attn_dropout_rate=self.attn_dropout_rate,
                                    out_dropout_rate=self.dropout_rate)(x, is_training=is_training)
        x = LayerScaleBlock(eps=self.layerscale_eps)(x, is_training=is_training)
        x = StochasticDepthBlock(drop_rate=self.stoch_depth_rate)(x, is_training=is_training)
        cls_token = cls_token + x                                               # cait.py:109
        y = nn.LayerNorm()(cls_token)
        y = FFBlock(expand_ratio=self.expand_ratio, dropout_rate=self.dropout_rate,
                    activation_fn=self.activation_fn)(y, is_training=is_training)
        y = LayerScaleBlock(eps=self.layerscale_eps)(y, is_training=is_training)
        y = StochasticDepthBlock(drop_rate=self.stoch_depth_rate)(y, is_training=is_training)
        return cls_token + y                                                    # cait.py:121


class CaiT(nn.Module):
    num_classes: int
    num_layers: int
    num_layers_token_only: int
    num_heads: int
    embed_dim: int
    patch_shape: Tuple[int]
    stoch_depth_rate: float
    layerscale_eps: float
    expand_ratio: float = 4
    attn_dropout_rate: float = 0.
    dropout_rate: float = 0.
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        x = PatchEmbedBlock(patch_shape=self.patch_shape, embed_dim=self.embed_dim)(inputs)
        x = Encoder(num_layers=self.num_layers, num_heads=self.num_heads,
                    expand_ratio=self.expand_ratio, attn_dropout_rate=self.attn_dropout_rate,
                    dropout_rate=self.dropout_rate, stoch_depth_rate=self.stoch_depth_rate,
                    layerscale_eps=self.layerscale_eps,
                    activation_fn=self.activation_fn)(x, is_training=is_training)
        b = x.shape[0]
        cls_param = self.param('cls', nn.initializers.zeros, (1, 1, self.embed_dim))
        cls_token = F.tile_param_rows(cls_param, b)                             # cait.py:157-160
        for _ in range(self.num_layers_token_only):
            cls_token = CAEncoderBlock(num_heads=self.num_heads, expand_ratio=self.expand_ratio,
                                       attn_dropout_rate=self.attn_dropout_rate,
                                       dropout_rate=self.dropout_rate,
                                       stoch_depth_rate=self.stoch_depth_rate,
                                       layerscale_eps=self.layerscale_eps,
                                       activation_fn=self.activation_fn)(
                                           x, cls_token, is_training=is_training)
        x = F.concat_tokens(cls_token, x)                                       # cait.py:175
        x = nn.LayerNorm()(x)
        cls_out = x[:, 0]
        return nn.Dense(features=self.num_classes, use_bias=True,
                        kernel_init=nn.initializers.zeros)(cls_out)             # cait.py:179-182
