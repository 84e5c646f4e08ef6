"""Transformer-in-Transformer behind the reference surface (models/tnt.py:10-182).

Two residual streams: the PIXEL stream [(b h w), 16, c] (every 16x16 patch as 4x4 sub-patches of
c = 24 / 40 channels; inner transformer with 4 heads of 6 / 10 channels -> csrc/attn_small.cu) and
the PATCH stream [b, 1 + 196, C] (outer transformer, 64-wide heads -> the fused tcgen05 attention).
Inner2OuterBlock flattens a patch's 16 pixel tokens, projects them to C and adds them to the patch
stream behind the class token.  Flax auto-names: PixelEmbedBlock_0/Dense_0, PatchEmbedBlock_0/Dense_0,
cls, AddAbsPosEmbed_{0,1}, Encoder_0/EncoderBlock_i/{LayerNorm_0..3, SelfAttentionBlock_{0,1},
FFBlock_{0,1}, Inner2OuterBlock_0/Dense_0}, Dense_0.
"""
from typing import Callable, Optional, Tuple

from sav import functional as F
from sav import linen as nn
from sav.train import HWCN

from models.layers import AddAbsPosEmbed, FFBlock, PatchEmbedBlock, SelfAttentionBlock


class PixelEmbedBlock(nn.Module):
    patch_shape: Tuple[int] = (16, 16)
    transformed_patch_shape: Tuple[int] = (4, 4)
    embed_dim: int = 0
    use_bias: bool = True
    dtype: object = None

    @nn.compact
    def __call__(self, inputs):
        ph, pw = self.patch_shape
        th, tw = self.transformed_patch_shape
        assert ph % th == 0                                                      # tnt.py:17
        assert pw % tw == 0                                                      # tnt.py:18
        if isinstance(inputs, HWCN):
            x = F.pixel_patchify(inputs.images, ph, pw, th, tw, hwcn=True)       # tnt.py:19-26
        else:
            x = F.pixel_patchify(inputs, ph, pw, th, tw, hwcn=False)
        return nn.Dense(features=self.embed_dim, use_bias=self.use_bias)(x)      # tnt.py:27-29


class Inner2OuterBlock(nn.Module):
    out_ch: Optional[int] = None
    dtype: object = None

    @nn.compact
    def __call__(self, patch_inputs, pixel_inputs):
        out_ch = self.out_ch or patch_inputs.shape[-1]
        n, t, d = pixel_inputs.shape
        x = F.stream_to_operand(pixel_inputs, (n, t * d))                        # tnt.py:40 '... n d -> ... (n d)'
        x = nn.Dense(features=out_ch)(x)                                         # tnt.py:41
        return F.inner_to_outer(patch_inputs, x)                                 # tnt.py:42-45: pad + add


class EncoderBlock(nn.Module):
    inner_num_heads: int = 4
    outer_num_heads: int = 6
    inner_expand_ratio: float = 4
    outer_expand_ratio: float = 4
    attn_dropout_rate: float = 0.
    dropout_rate: float = 0.
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, patch_inputs, pixel_inputs, is_training: bool):
        inner_x = nn.LayerNorm()(pixel_inputs)                                   # tnt.py:61
        inner_x = SelfAttentionBlock(num_heads=self.inner_num_heads,
                                     attn_dropout_rate=self.attn_dropout_rate,
                                     out_dropout_rate=self.dropout_rate)(inner_x, is_training=is_training)
        inner_x = inner_x + pixel_inputs                                         # tnt.py:67
        inner_y = nn.LayerNorm()(inner_x)
        inner_y = FFBlock(expand_ratio=self.inner_expand_ratio,
                          dropout_rate=self.dropout_rate)(inner_y, is_training=is_training)
        inner_output = inner_x + inner_y                                         # tnt.py:72

        outer_x = Inner2OuterBlock()(patch_inputs, inner_output)                 # tnt.py:74
        outer_x = nn.LayerNorm()(outer_x)
        outer_x = SelfAttentionBlock(num_heads=self.outer_num_heads,
                                     attn_dropout_rate=self.attn_dropout_rate,
                                     out_dropout_rate=self.dropout_rate)(outer_x, is_training=is_training)
        outer_x = outer_x + patch_inputs                                         # tnt.py:81 (patch_inputs!)
        outer_y = nn.LayerNorm()(outer_x)
        outer_y = FFBlock(expand_ratio=self.outer_expand_ratio,
                          dropout_rate=self.dropout_rate)(outer_y, is_training=is_training)
        return outer_x + outer_y, inner_output                                   # tnt.py:86-87


class Encoder(nn.Module):
    num_layers: int = 12
    inner_num_heads: int = 4
    outer_num_heads: int = 6
    inner_expand_ratio: float = 4
    outer_expand_ratio: float = 4
    attn_dropout_rate: float = 0.
    dropout_rate: float = 0.
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, patch_embeddings, pixel_embeddings, is_training: bool):
        for _ in range(self.num_layers):                                         # tnt.py:103-112
            patch_embeddings, pixel_embeddings = EncoderBlock(
                inner_num_heads=self.inner_num_heads, outer_num_heads=self.outer_num_heads,
                attn_dropout_rate=self.attn_dropout_rate, dropout_rate=self.dropout_rate,
                activation_fn=self.activation_fn)(patch_embeddings, pixel_embeddings,
                                                  is_training=is_training)
        return patch_embeddings


class TNT(nn.Module):
    num_classes: int = 1000
    num_layers: int = 12
    inner_num_heads: int = 4
    outer_num_heads: int = 6
    inner_embed_dim: int = 24
    outer_embed_dim: int = 384
    patch_shape: Tuple[int] = (16, 16)
    transformed_patch_shape: Tuple[int] = (4, 4)
    inner_expand_ratio: float = 4
    outer_expand_ratio: float = 4
    attn_dropout_rate: float = 0.
    dropout_rate: float = 0.
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        pixel_tok = PixelEmbedBlock(patch_shape=self.patch_shape,
                                    transformed_patch_shape=self.transformed_patch_shape,
                                    embed_dim=self.inner_embed_dim)(inputs)      # tnt.py:135-139
        patch_tok = PatchEmbedBlock(patch_shape=self.patch_shape, embed_dim=self.outer_embed_dim,
                                    use_bias=True)(inputs)                       # tnt.py:141-144
        cls_token = self.param('cls', nn.initializers.zeros, (1, 1, self.outer_embed_dim))
        # tnt.py:146-153: the class-token concat is fused with the patch position-embedding add
        pixel_embeddings = AddAbsPosEmbed()(pixel_tok)                           # tnt.py:155
        patch_embeddings = AddAbsPosEmbed()(patch_tok, cls=cls_token)            # tnt.py:156
        patch_embeddings = nn.Dropout(rate=self.dropout_rate)(patch_embeddings,
                                                              deterministic=not is_training)
        patch_embeddings = Encoder(num_layers=self.num_layers,
                                   inner_num_heads=self.inner_num_heads,
                                   outer_num_heads=self.outer_num_heads,
                                   attn_dropout_rate=self.attn_dropout_rate,
                                   dropout_rate=self.dropout_rate,
                                   activation_fn=self.activation_fn)(patch_embeddings, pixel_embeddings,
                                                                     is_training=is_training)
        cls_out = F.stream_first_token(patch_embeddings)                         # tnt.py:172 (no final LN)
        return nn.Dense(features=self.num_classes,
                        kernel_init=nn.initializers.zeros)(cls_out)              # tnt.py:173-177
