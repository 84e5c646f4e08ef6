"""ViT / DeiT behind the reference surface (models/vit.py:9-99): pre-LN EncoderBlock, Encoder with
learned absolute position embedding + final LayerNorm, class token, Dense head on x[:, 0].

Parameter names follow Flax auto-naming exactly so a reference checkpoint tree loads:
PatchEmbedBlock_0/Dense_0, cls, Encoder_0/{AddAbsPosEmbed_0, EncoderBlock_i/{LayerNorm_0,
SelfAttentionBlock_0, LayerNorm_1, FFBlock_0}, LayerNorm_0}, Dense_0.
"""
from typing import Callable, Tuple

from sav import linen as nn

from models.layers import AddAbsPosEmbed, FFBlock, PatchEmbedBlock, SelfAttentionBlock


class EncoderBlock(nn.Module):
    num_heads: int
    expand_ratio: float = 4
    attn_dropout_rate: float = 0.
    dropout_rate: float = 0.
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        y = nn.LayerNorm()(inputs)                                              # vit.py:19
        y = SelfAttentionBlock(num_heads=self.num_heads,
                               attn_dropout_rate=self.attn_dropout_rate,
                               out_dropout_rate=self.dropout_rate)(y, is_training=is_training)
        x = y + inputs                     # vit.py:24 -> residual in the out-proj GEMM epilogue
        y = nn.LayerNorm()(x)                                                   # vit.py:26
        y = FFBlock(expand_ratio=self.expand_ratio, dropout_rate=self.dropout_rate,
                    activation_fn=self.activation_fn)(y, is_training=is_training)
        return x + y                       # vit.py:31 -> residual in the Dense_1 GEMM epilogue


class Encoder(nn.Module):
    num_layers: int
    num_heads: int
    expand_ratio: float = 4
    attn_dropout_rate: float = 0.
    dropout_rate: float = 0.
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool, cls=None):
        x = AddAbsPosEmbed()(inputs, cls=cls)                                   # vit.py:46
        x = nn.Dropout(rate=self.dropout_rate)(x, deterministic=not is_training)
        for _ in range(self.num_layers):                                        # vit.py:49-55
            x = EncoderBlock(num_heads=self.num_heads, expand_ratio=self.expand_ratio,
                             attn_dropout_rate=self.attn_dropout_rate,
                             dropout_rate=self.dropout_rate,
                             activation_fn=self.activation_fn)(x, is_training=is_training)
        return nn.LayerNorm()(x)                                                # vit.py:57


class ViT(nn.Module):
    num_classes: int
    num_layers: int
    num_heads: int
    embed_dim: int
    patch_shape: Tuple[int]
    expand_ratio: float = 4
    attn_dropout_rate: float = 0.
    dropout_rate: float = 0.
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        assert self.embed_dim % self.num_heads == 0                             # vit.py:75
        tok = PatchEmbedBlock(patch_shape=self.patch_shape, embed_dim=self.embed_dim)(inputs)
        cls_token = self.param('cls', nn.initializers.zeros, (1, 1, self.embed_dim))
        # vit.py:81-85 tiles + concatenates the class token; here it is fused with the pos-embed add
        x = Encoder(num_layers=self.num_layers, num_heads=self.num_heads,
                    expand_ratio=self.expand_ratio, attn_dropout_rate=self.attn_dropout_rate,
                    dropout_rate=self.dropout_rate,
                    activation_fn=self.activation_fn)(tok, is_training=is_training, cls=cls_token)
        cls_out = x[:, 0]                                                       # vit.py:95
        return nn.Dense(features=self.num_classes,
                        kernel_init=nn.initializers.zeros)(cls_out)             # vit.py:96-98
