"""BoTNet multi-head self-attention with 2-D relative position logits behind the reference surface
(models/botnet.py:70-199).  Only the MHSA of the bottleneck-transformer block is on the hot path
(SURVEY 2a); the ResNet conv / BatchNorm / squeeze-excite stages are out of scope this round, so
`create_model('botnet_t*')` is not provided -- `BoTMHSA` is used on [B, H, W, C] activations.

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
