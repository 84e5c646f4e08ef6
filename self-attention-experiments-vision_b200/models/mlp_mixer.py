"""MLP-Mixer behind the reference surface (models/mlp_mixer.py:10-60).

MixerBlock: LayerNorm -> FFBlock over the TOKEN axis ('... l d -> ... d l', expand ratio 0.5) ->
residual; LayerNorm -> FFBlock over the channels (expand ratio 4) -> residual.  MLPMixer: patch
embedding with bias, the blocks, final LayerNorm, mean over tokens, Dense head.

The token-mixing FFBlock is two ordinary GEMMs over [B*C, L] rows between two transposes
(sav.functional.TokenMix): L = 196 or 49 tokens are padded to the GEMM's pitch unit with exact
zeros, and the transpose back is fused with the residual add into the fp32 stream.  Parameter names
follow Flax auto-naming: PatchEmbedBlock_0/Dense_0, MixerBlock_i/{LayerNorm_0, FFBlock_0/{Dense_0,
Dense_1}, LayerNorm_1, FFBlock_1/{Dense_0, Dense_1}}, LayerNorm_0, Dense_0.
"""
from typing import Callable, Tuple

from sav import functional as F
from sav import linen as nn

from models.layers import FFBlock, PatchEmbedBlock


class MixerBlock(nn.Module):
    tokens_expand_ratio: float = 0.5
    channels_expand_ratio: float = 4
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        x = nn.LayerNorm()(inputs)                                               # mlp_mixer.py:18
        x = F.swap_tokens_channels(x)                                            # :19
        x = FFBlock(expand_ratio=self.tokens_expand_ratio,
                    activation_fn=self.activation_fn)(x, is_training=is_training)
        x = F.swap_tokens_channels(x)                                            # :23
        x = x + inputs                                                           # :24

        y = nn.LayerNorm()(x)                                                    # :26
        y = FFBlock(expand_ratio=self.channels_expand_ratio,
                    activation_fn=self.activation_fn)(y, is_training=is_training)
        return x + y                                                             # :30


class MLPMixer(nn.Module):
    num_classes: int = 1000
    num_layers: int = 8
    embed_dim: int = 512
    patch_shape: Tuple[int] = (16, 16)
    tokens_expand_ratio: float = 0.5
    channels_expand_ratio: float = 4
    activation_fn: Callable = nn.activation.gelu
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        x = PatchEmbedBlock(patch_shape=self.patch_shape, embed_dim=self.embed_dim,
                            use_bias=True)(inputs)                               # :46-49
        x = F.to_stream(x)                 # the patch embedding itself is the residual stream
        for _ in range(self.num_layers):
            x = MixerBlock(tokens_expand_ratio=self.tokens_expand_ratio,
                           channels_expand_ratio=self.channels_expand_ratio,
                           activation_fn=self.activation_fn)(x, is_training=is_training)
        x = nn.LayerNorm()(x)                                                    # :57
        B, L, C = x.shape
        x = F.spatial_mean(F.reshape(x, (B, L, 1, C)))                           # :58 jnp.mean(x, axis=1)
        return nn.Dense(features=self.num_classes)(x)                            # :59
