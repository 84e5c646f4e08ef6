"""Multi-head attention block behind the reference surface (models/layers/attentions/attention.py).

Same fields, call signature, assertions and parameter names (`queries`/`keys`/`values`/
`DenseGeneral_0` kernels of shape [C,H,d] / [H,d,C]) as AttentionBlock (attention.py:10-67), but the
three projections are ONE tcgen05 GEMM against a fused [C,3,H,d] weight block and
scale/QK^T/softmax/PV (attention.py:39-58) is one fused kernel that keeps S in TMEM.
"""
from typing import Optional

from sav import functional as F
from sav import linen as nn

from models.layers.attentions.talking_heads import TalkingHeadsBlock


class AttentionBlock(nn.Module):
    num_heads: int
    head_ch: Optional[int] = None
    out_ch: Optional[int] = None
    talking_heads: bool = False
    attn_dropout_rate: float = 0.
    out_dropout_rate: float = 0.
    use_bias: bool = False
    dtype: object = None

    @nn.compact
    def __call__(self, inputs_q, inputs_kv, is_training: bool):
        assert inputs_q.ndim == inputs_kv.ndim == 3          # attention.py:22
        in_ch = inputs_q.shape[-1]
        assert in_ch % self.num_heads == 0                   # attention.py:25
        H = self.num_heads
        d = self.head_ch or in_ch // H
        out_ch = self.out_ch or in_ch
        if self.use_bias:
            raise NotImplementedError("AttentionBlock(use_bias=True)")
        if (self.attn_dropout_rate or self.out_dropout_rate) and is_training:
            raise NotImplementedError("attention dropout with rate > 0")
        (wq, wk, wv), w_qkv = self.fused_params(("queries", "keys", "values"), "kernel",
                                                nn.initializers.lecun_normal(), (in_ch, H, d))
        scope = self._scope_path()
        w_out = self._param(scope + ("DenseGeneral_0", "kernel"),
                            nn.initializers.lecun_normal(in_axes=2), (H, d, out_ch))
        mix = None
        if self.talking_heads:                               # attention.py:44-46,50-52
            mix = (TalkingHeadsBlock(num_heads=H)(), TalkingHeadsBlock(num_heads=H)())
        # plain attention with d = 64 runs in the fused TMEM kernel; talking heads and other head
        # dims go through the materialised-logits path (batched GEMMs + one row kernel)
        # (a handful of query rows -- CeiT's layer-wise class attention, one query over 12 keys -- is
        # not worth a 128-row TMEM tile either)
        fused = mix is None and d == 64 and inputs_q.shape[1] >= 8
        if inputs_q is inputs_kv:
            qkv = F.linear(inputs_q, w_qkv)                  # attention.py:35-37, one GEMM
            if mix is None and d <= 16 and inputs_q.shape[1] <= 32:
                # TNT's inner transformer (tnt.py:60-69): 16 tokens, head dim 6 / 10
                ctx = F.small_self_attention_core(qkv, H, d)
            else:
                ctx = F.self_attention_core(qkv, H, d) if fused else \
                    F.mat_self_attention_core(qkv, H, d, mix)    # attention.py:39-58
        else:
            hd = H * d
            gq = w_qkv.gw[:, :hd] if w_qkv.gw is not None else None
            gkv = w_qkv.gw[:, hd:] if w_qkv.gw is not None else None
            q = F.linear(inputs_q, F.Weight(w_qkv.w[:, :hd], gq, [wq]))
            kv = F.linear(inputs_kv, F.Weight(w_qkv.w[:, hd:], gkv, [wk, wv]))
            ctx = F.cross_attention_core(q, kv, H, d) if fused else \
                F.mat_cross_attention_core(q, kv, H, d, mix)
        gw = w_out.grad.view(H * d, out_ch) if w_out.grad is not None else None
        return F.linear(ctx, F.Weight(w_out.shadow.view(H * d, out_ch), gw, [w_out],
                                      wm=w_out.master.view(H * d, out_ch)))          # :60-63


class SelfAttentionBlock(AttentionBlock):

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        return super().__call__(inputs, inputs, is_training=is_training)
