"""BoTNet behind the reference surface (models/botnet.py): ResNet bottleneck stages
(BottleneckResNetBlock, botnet.py:17-67), bottleneck-transformer blocks (BoTBlock, :202-252) around
the multi-head self-attention with 2-D relative position logits (RelativeLogits / BoTMHSA, :70-199),
and the network (BoTNet, :255-331).

Conv stages: channel-last bf16 activations; every conv is the tcgen05 GEMM (1x1 directly, 3x3 / 7x7
through im2col, csrc/conv2d.cu); BatchNorm + swish is one pass (csrc/cnn.cu), the squeeze-excite
gate, the skip connection and the closing swish of a block are one pass.  BatchNorm keeps running
averages in 'batch_stats' (momentum 0.9, eps 1e-5) and switches to them when train=False.

The reference module is broken as written (undefined `self.head_dim`, softmax over
the last key axis only, V indexed by the query position).  This implements the intended BoTNet
math, which is also what the oracle restates: q, k, v = 1x1 convs without bias; q / sqrt(d);
    logits[q, k] = q.k + q.r^W[kx - qx + W - 1] + q.r^H[ky - qy + H - 1]
softmax over all H*W keys; O = P V.  Kernel mapping: the three 1x1 convs are GEMMs (q alone, k|v
fused); the per-query relative tables q.r^W | q.r^H are ONE skinny GEMM against the stacked
[2W-1 + 2H-1, d] table; the gather into the logits, the softmax and the scatter of the table
gradient are the row kernel of attn_mat.cu between two batched GEMMs.

Parameter names follow the reference: query/kernel, key/kernel, value/kernel ([1,1,C,h*d], flax
nn.Conv layout) and RelativeLogits_0/{rel_pos_emb_w, rel_pos_emb_h}.
"""
import math

import torch

from sav import functional as F
from sav import linen as nn


def he_uniform(gen, shape):
    """flax initializers.he_uniform (botnet.py:153): U(-l, l), l = sqrt(6 / fan_in)."""
    fan_in = 1
    for s in shape[:-1]:
        fan_in *= s
    lim = math.sqrt(6.0 / max(fan_in, 1))
    return (torch.rand(shape, generator=gen) * 2.0 - 1.0) * lim


class RelativeLogits(nn.Module):
    """Owns rel_pos_emb_w [2W-1, d] and rel_pos_emb_h [2H-1, d] (botnet.py:124-132), stored back to
    back so that q . [r^W | r^H] is a single GEMM.  Called with the key grid (H, W), returns the
    stacked table as a GEMM weight."""
    head_ch: int

    @nn.compact
    def __call__(self, H, W):
        std = self.head_ch ** -0.5
        _, table = self.stacked_params(('rel_pos_emb_w', 'rel_pos_emb_h'), (2 * W - 1, 2 * H - 1),
                                       self.head_ch, lambda g, shape: torch.randn(shape, generator=g) * std)
        return table


class BoTMHSA(nn.Module):
    num_heads: int
    head_ch: int
    dtype: object = None
    precision: object = None
    kernel_init: object = None
    bias_init: object = None

    @nn.compact
    def __call__(self, inputs):
        assert inputs.ndim == 4                              # [batch, height, width, dim]
        B, H, W, C = inputs.shape
        h, d = self.num_heads, self.head_ch
        hd = h * d
        init = self.kernel_init or he_uniform
        (wq, wk, wv), w_qkv = self.fused_params(("query", "key", "value"), "kernel", init, (C, hd),
                                                lead=(1, 1))                      # botnet.py:165-174
        table = RelativeLogits(head_ch=d)(H, W)                                   # botnet.py:191
        x = F.reshape(inputs, (B, H * W, C))
        gq = w_qkv.gw[:, :hd] if w_qkv.gw is not None else None
        gkv = w_qkv.gw[:, hd:] if w_qkv.gw is not None else None
        q = F.linear(x, F.Weight(w_qkv.w[:, :hd], gq, [wq]))
        kv = F.linear(x, F.Weight(w_qkv.w[:, hd:], gkv, [wk, wv]))
        ctx = F.mat_cross_attention_core(q, kv, h, d, rel_tab=table, rel_hw=(H, W))
        return F.reshape(ctx, (B, H, W, hd))                                      # botnet.py:198


# ------------------------------------------------------------------------------ conv stages
from functools import partial  # noqa: E402
from typing import Any, Callable, Tuple  # noqa: E402

from models.layers.squeeze_excite import SqueezeExciteBlock  # noqa: E402

ModuleDef = Any


class BottleneckResNetBlock(nn.Module):
    """ResNet bottleneck block (botnet.py:17-67): 1x1 -> 3x3 (strided) -> 1x1 convs, each followed by
    BatchNorm (+ swish for the first two), squeeze-excite, projection skip when the shape changes."""
    filters: int
    conv: ModuleDef
    norm: ModuleDef
    strides: Tuple[int]
    se_ratio: float = 0.0625
    projection_factor: int = 4
    activation_fn: Callable = nn.activation.swish
    dtype: object = None

    @nn.compact
    def __call__(self, inputs):
        act = self.activation_fn
        residual = inputs
        y = self.conv(self.filters, kernel_size=(1, 1))(inputs)                   # botnet.py:43
        y = self.norm()(y, act=act)                                              # :44-45
        y = self.conv(self.filters, kernel_size=(3, 3), strides=self.strides)(y)  # :46
        y = self.norm()(y, act=act)                                              # :47-48
        y = self.conv(self.filters * self.projection_factor, kernel_size=(1, 1))(y)   # :49-50
        y = self.norm(scale_init=nn.initializers.zeros)(y)                       # :51
        gate = None
        if self.se_ratio is not None:                                            # :53-56
            gate = SqueezeExciteBlock(se_ratio=self.se_ratio, activation_fn=act)(y, gate_only=True)
        if tuple(residual.shape) != tuple(y.shape):                              # :58-63
            residual = self.conv(self.filters * self.projection_factor, kernel_size=(1, 1),
                                 strides=self.strides)(residual)
            residual = self.norm()(residual, act=act)
        return F.gate_residual_act(y, gate, residual, nn._act_kind(act))         # :65-66


class BoTBlock(nn.Module):
    """Bottleneck transformer block (botnet.py:202-252).  The reference reads an undefined
    `self.config`; the attention geometry is carried as fields."""
    filters: int
    conv: ModuleDef
    norm: ModuleDef
    strides: Tuple[int, int]
    projection_factor: int = 4
    activation_fn: ModuleDef = nn.swish
    num_heads: int = 4
    head_dim: int = 128

    @nn.compact
    def __call__(self, inputs):
        act = self.activation_fn
        residual = inputs
        y = self.conv(self.filters, kernel_size=(1, 1))(inputs)                   # botnet.py:229
        y = self.norm()(y, act=act)                                              # :230-231
        y = BoTMHSA(num_heads=self.num_heads, head_ch=self.head_dim)(y)           # :232
        if tuple(self.strides) == (2, 2):                                        # :233-237
            raise NotImplementedError("BoTBlock with strides (2, 2) (stride_one=False): avg_pool")
        y = self.norm()(y, act=act)                                              # :238-239
        y = self.conv(self.filters * self.projection_factor, kernel_size=(1, 1))(y)   # :240-241
        y = self.norm(scale_init=nn.initializers.zeros)(y)                       # :242
        if tuple(residual.shape) != tuple(y.shape):                              # :244-249
            residual = self.conv(self.filters * self.projection_factor, kernel_size=(1, 1),
                                 strides=self.strides)(residual)
            residual = self.norm()(residual, act=act)
        return F.gate_residual_act(y, None, residual, nn._act_kind(act))         # :251


class BoTNet(nn.Module):
    """Bottleneck Transformer Network (botnet.py:255-331)."""
    num_classes: int
    stage_sizes: Tuple[int]
    stride_one: bool = True
    se_ratio: float = 0.0625
    activation_fn: ModuleDef = nn.swish
    num_heads: int = 4
    head_dim: int = 128
    initial_filters: int = 64
    projection_factor: int = 4
    bn_momentum: float = 0.9
    bn_epsilon: float = 1e-5
    dtype: object = None
    precision: Any = None
    kernel_init: Callable = he_uniform
    bias_init: Callable = None

    @nn.compact
    def __call__(self, inputs, train: bool = True, is_training: bool = None):
        if is_training is not None:            # botnet_test.py:22 passes is_training=
            train = is_training
        act = self.activation_fn
        conv = partial(nn.Conv, use_bias=False, kernel_init=self.kernel_init)    # botnet.py:280-285
        norm = partial(nn.BatchNorm, use_running_average=not train, momentum=self.bn_momentum,
                       epsilon=self.bn_epsilon)                                  # :286-290
        y = conv(self.initial_filters, kernel_size=(7, 7), strides=(2, 2),
                 padding=[(3, 3), (3, 3)])(inputs)                               # :292-295
        y = norm()(y, act=act)                                                   # :296-297
        y = F.maxpool3x3s2(y)                                                    # :298
        i = 0
        for i, block_size in enumerate(self.stage_sizes[:-1]):                   # :299-311
            for j in range(block_size):
                strides = (2, 2) if i > 0 and j == 0 else (1, 1)
                y = BottleneckResNetBlock(filters=self.initial_filters * 2 ** i, strides=strides,
                                          conv=conv, norm=norm, se_ratio=self.se_ratio,
                                          projection_factor=self.projection_factor,
                                          activation_fn=act)(y)
        for j in range(self.stage_sizes[-1]):                                    # :312-320
            strides = (2, 2) if j == 0 and self.stride_one is False else (1, 1)
            y = BoTBlock(filters=self.initial_filters * 2 ** (i + 1), strides=strides, conv=conv,
                         norm=norm, projection_factor=self.projection_factor, activation_fn=act,
                         num_heads=self.num_heads, head_dim=self.head_dim)(y)
        y = F.spatial_mean(y)                                                    # :321
        return nn.Dense(self.num_classes, kernel_init=self.kernel_init,
                        bias_init=self.bias_init or nn.initializers.normal(stddev=1e-6))(y)   # :322-325
