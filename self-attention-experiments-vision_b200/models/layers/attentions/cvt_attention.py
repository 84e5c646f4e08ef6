"""CvT attention behind the reference surface (models/layers/attentions/cvt_attention.py:12-120).

ConvProjectionBlock = depthwise 3x3 conv (stride s, SAME, no bias) -> BatchNorm -> 1x1 conv; the
q / k / v projections use strides (1, 2, 2), so Nk = Nq / 4.  Kernel mapping: dwconv + BatchNorm
statistics are one channel-last pass over the fp32 stream (conv.cu), BatchNorm apply is one
elementwise pass producing the bf16 GEMM operand, the 1x1 conv is the tcgen05 GEMM, and the
attention core is the fused TMEM kernel with independent Nq / Nk (head dim 64 at every CvT stage).

Parameter names follow Flax auto-naming: ConvProjectionBlock_{0,1,2}/{Conv_0/kernel [3,3,1,C],
BatchNorm_0/{scale,bias}, Conv_1/kernel [1,1,C,h*d] (+ Conv_1/bias with use_bias)},
DenseGeneral_0/kernel [h,d,C].  BatchNorm follows flax: batch statistics when is_training, running
averages otherwise (use_running_average=not is_training, cvt_attention.py:33-36); the running
averages live in the 'batch_stats' collection ({...BatchNorm_0: {mean, var}}), are created by init()
and updated in place (momentum 0.9) when apply() is called with mutable=['batch_stats'].  The
reference's train.py never passes mutable, where flax would raise; here the update is
then skipped so that its apply_fn(params, images, is_training=True) call shape keeps working.
"""
from typing import Optional, Tuple

from sav import functional as F
from sav import linen as nn


class ConvProjectionBlock(nn.Module):
    out_ch: int
    kernel_size: int = 3
    strides: int = 1
    use_bias: bool = True
    bn_momentum: float = 0.9
    bn_epsilon: float = 1e-5
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        assert inputs.ndim == 4
        in_ch = inputs.shape[-1]
        if self.kernel_size != 3:
            raise NotImplementedError("ConvProjectionBlock: only the 3x3 depthwise kernel is built")
        scope = self._scope_path()
        w_dw = self._param(scope + ("Conv_0", "kernel"), nn.initializers.lecun_normal(in_axes=3),
                           (3, 3, 1, in_ch))                               # cvt_attention.py:25-30
        gamma = self._param(scope + ("BatchNorm_0", "scale"), nn.initializers.ones, (in_ch,))
        beta = self._param(scope + ("BatchNorm_0", "bias"), nn.initializers.zeros, (in_ch,))
        w_pw = self._param(scope + ("Conv_1", "kernel"), nn.initializers.lecun_normal(in_axes=3),
                           (1, 1, in_ch, self.out_ch))                     # cvt_attention.py:37-39
        b_pw = self._param(scope + ("Conv_1", "bias"), nn.initializers.zeros, (self.out_ch,)) \
            if self.use_bias else None                                   # cvt_attention.py:39
        ra = self._batch_stats(scope + ("BatchNorm_0",), in_ch)
        z = F.dwconv_bn(inputs, w_dw, gamma, beta, self.strides, self.bn_epsilon, running=ra,
                        momentum=self.bn_momentum, training=is_training,
                        update=self.is_mutable_collection("batch_stats"))
        gw = w_pw.grad.view(in_ch, self.out_ch) if w_pw.grad is not None else None
        return F.linear(z, F.Weight(w_pw.shadow.view(in_ch, self.out_ch), gw, [w_pw],
                                    wm=w_pw.master.view(in_ch, self.out_ch)), b_pw)

    def _batch_stats(self, path, C):
        """flax nn.BatchNorm state: batch_stats/<...>/BatchNorm_0/{mean (zeros), var (ones)}."""
        import torch
        from sav import linen
        run = linen._run()
        out = []
        for name, fill in (("mean", 0.0), ("var", 1.0)):
            p = path + (name,)
            if p not in run.batch_stats:
                if run.mode == "apply":
                    if not run.batch_stats:
                        return None      # apply() without the collection: training-mode statistics only
                    raise KeyError(f"batch_stats/{'/'.join(p)} not found in variables")
                run.batch_stats[p] = torch.full((C,), fill, dtype=torch.float32, device=run.device)
            out.append(run.batch_stats[p])
        return tuple(out)


class CvTAttentionBlock(nn.Module):
    num_heads: int
    head_ch: Optional[int] = None
    out_ch: Optional[int] = None
    talking_heads: bool = False
    attn_dropout_rate: float = 0.
    out_dropout_rate: float = 0.
    kernel_size: int = 3
    strides: Tuple[int] = (1, 2, 2)
    use_bias: bool = False
    bn_momentum: float = 0.9
    bn_epsilon: float = 1e-5
    dtype: object = None

    @nn.compact
    def __call__(self, inputs_q, inputs_kv, is_training: bool):
        assert inputs_q.ndim == 4                              # cvt_attention.py:59-61
        assert inputs_kv.ndim == 4
        assert len(self.strides) == 3
        in_ch = inputs_q.shape[-1]
        assert in_ch % self.num_heads == 0                     # cvt_attention.py:65
        H = self.num_heads
        d = self.head_ch or in_ch // H
        out_ch = self.out_ch or in_ch
        if self.talking_heads:
            raise NotImplementedError("CvTAttentionBlock(talking_heads=True): no model uses it")
        if (self.attn_dropout_rate or self.out_dropout_rate) and is_training:
            raise NotImplementedError("attention dropout with rate > 0")
        proj = [ConvProjectionBlock(out_ch=H * d, kernel_size=self.kernel_size, strides=s,
                                    use_bias=self.use_bias, bn_momentum=self.bn_momentum,
                                    bn_epsilon=self.bn_epsilon)(x, is_training=is_training)
                for s, x in zip(self.strides, (inputs_q, inputs_kv, inputs_kv))]   # :77-79
        B = inputs_q.shape[0]
        q, k, v = (F.reshape(t, (B, t.shape[1] * t.shape[2], H * d)) for t in proj)   # :81-83
        ctx = F.attention3(q, k, v, H, d)                                          # :85-104
        scope = self._scope_path()
        w_out = self._param(scope + ("DenseGeneral_0", "kernel"),
                            nn.initializers.lecun_normal(in_axes=2), (H, d, out_ch))
        gw = w_out.grad.view(H * d, out_ch) if w_out.grad is not None else None
        return F.linear(ctx, F.Weight(w_out.shadow.view(H * d, out_ch), gw, [w_out],
                                      wm=w_out.master.view(H * d, out_ch)))        # :106-109


class CvTSelfAttentionBlock(CvTAttentionBlock):

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        return super().__call__(inputs, inputs, is_training)
