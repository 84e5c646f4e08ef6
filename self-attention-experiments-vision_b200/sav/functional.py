"""Var-level ops: forward launches + recorded backward closures, with define-by-run fusion.

Precision layout: the residual stream (embed output, every `x + branch(x)` result) and
its gradient are fp32 `Var`s; everything that feeds a tensor-core MMA (LayerNorm outputs, q/k/v,
attention context, MLP hidden, and their gradients) is bf16.  A residual-stream Var carries its
gradient twice: `g` (fp32, what flows on) and `g16` (bf16 copy written by the same LayerNorm-
backward kernel, the MMA operand of the dX / dW GEMMs).

`linear` returns a LazyLinear that absorbs a following gelu / LayerScale / drop-path scale /
residual add into its GEMM epilogue before it is materialised, so `x + FFBlock(LN(x))` written like
the reference (vit.py:24,31) runs as fused kernels; `layernorm` is lazy in the same way so that
`LayerNorm(x)[:, 0]` (vit.py:57,95) only normalises the class-token rows.  Abstract ('meta' device)
inputs only propagate shapes (the analogue of jax.eval_shape) -- used by the CPU shape tests; there
is no CPU compute.
"""
import math

import os

import torch

from . import ops
from .tape import Var, current_tape, record

BF16 = torch.bfloat16
F32 = torch.float32


def _meta(t):
    return t.is_meta


class Weight:
    """2-D bf16 weight view [K,N] (unit inner stride) + its fp32 gradient view + owning Params."""
    __slots__ = ("w", "gw", "params", "wm")

    def __init__(self, w, gw, params, wm=None):
        self.w, self.gw, self.params, self.wm = w, gw, params, wm

    def touch(self):
        for p in self.params:
            p.touch_grad()


# The out-proj / FC2 bias gradients (column sums of the bf16 stream-gradient copy) are accumulated
# by LayerNorm backward itself in register accumulators: +1.6 us on the 60 us kernel against a
# separate 14.4 us column-sum pass (profiles/r02d_membound.log; 13.29 -> 13.18 ms per DeiT-S step).
# SAV_FUSE_BIAS_COLSUM=0 restores the separate pass (A/B switch).
_FUSE_BIAS_COLSUM = os.environ.get("SAV_FUSE_BIAS_COLSUM", "1") == "1"
# The FC1 bias gradient (column sums of dH * gelu') is accumulated by the FC2-dX GEMM's epilogue
# (sav_gemm dgelu_colsum).  SAV_FUSE_DGELU_COLSUM=0 restores the separate column-sum pass (A/B).
_FUSE_DGELU_COLSUM = os.environ.get("SAV_FUSE_DGELU_COLSUM", "1") == "1"
# LayerNorm of the stream (vit.py:19,26) computed by the epilogue of the GEMM that writes the stream
# (out-projection / FC2 with bias + residual): the residual-fused Dense stays pending after `x + f(x)`
# and the LayerNorm that reads it first runs both in one launch.
_FUSE_LN_GEMM = os.environ.get("SAV_FUSE_LN_GEMM", "1") == "1"


def _grad16(var):
    """bf16 form of a Var's gradient (for MMA operands); consumes the bf16 copy if there is one."""
    g16 = var.take_grad16()
    if g16 is not None:
        return g16
    g = var.g
    if g is None:
        return None
    if g.dtype == BF16:
        return g
    out = torch.empty(g.shape, dtype=BF16, device=g.device)
    ops.cast_bf16(g.contiguous(), out)
    return out


def _as2d(t, K):
    """2-D [rows, K] view of an activation without copying: contiguous tensors reshape; [B,1,K]
    row selections (class-token rows, cait.py:13) and 2-D strided views pass through."""
    if t.is_contiguous():
        return t.reshape(-1, K)
    if t.dim() == 3 and t.shape[1] == 1:
        return t[:, 0]
    if t.dim() == 2:
        return t
    return t.contiguous().reshape(-1, K)


class LazyLinear(Var):
    __slots__ = ("bias_grad_from", "x", "weight", "bias", "act", "col_scale", "row_scale", "rows_per_scale",
                 "residual", "out_shape", "_done", "f32_out")

    def __init__(self, x, weight, bias, out_shape):
        super().__init__(None)
        self.x, self.weight, self.bias = x, weight, bias
        self.act = 0
        self.col_scale = self.row_scale = self.residual = None
        self.rows_per_scale = 1
        self.out_shape = tuple(out_shape)
        self._done = False
        self.f32_out = False      # written as the fp32 stream without a skip input (to_stream)

    @property
    def shape(self):
        return self.out_shape

    @property
    def ndim(self):
        return len(self.out_shape)

    @property
    def dtype(self):
        return F32 if self.residual is not None or self.f32_out else BF16

    @property
    def v(self):
        if not self._done:
            self._materialize()
        return self._v

    def ln_fusable(self):
        """The pending stream-producing Dense can take the following LayerNorm into its epilogue."""
        if not _FUSE_LN_GEMM or self._done or self.residual is None or self.act or \
                self.col_scale is not None or self.row_scale is not None:
            return False
        return not _meta(self.weight.w) and ops.gemm_ln_tiles(self.weight.w.shape[1]) != 0

    def _materialize(self, ln=None):
        self._done = True
        x, W = self.x, self.weight
        xin = x.v
        K, N = W.w.shape
        stream = self.residual is not None or self.f32_out
        if _meta(xin):
            self._v = torch.empty(self.out_shape, dtype=F32 if stream else BF16, device="meta")
            return
        if xin.dtype != BF16:
            raise TypeError("linear: the input of a Dense must be a bf16 activation (apply LayerNorm "
                            "to the fp32 residual stream first)")
        x2 = _as2d(xin, K)
        M = x2.shape[0]
        out = torch.empty((M, N), dtype=F32 if stream else BF16, device=x2.device)
        pre = torch.empty((M, N), dtype=BF16, device=x2.device) if self.act else None
        res = None
        if self.residual is not None:
            res = self.residual.v
            if res.dtype != F32:
                raise TypeError("residual add: the skip input must be the fp32 residual stream")
            res = res.reshape(M, N)
        got = ops.gemm(x2, W.w, b_mn=True, out=out,
                       bias=self.bias.master if self.bias is not None else None, act=self.act,
                       aux_out=pre,
                       col_scale=self.col_scale.master if self.col_scale is not None else None,
                       row_scale=self.row_scale, rows_per_scale=self.rows_per_scale, residual=res, ln=ln)
        self._v = out.view(self.out_shape)
        self.pre_act = pre
        record(self._backward)
        return got[1] if ln is not None else None

    def _backward(self):
        if self.g is None and self.g16 is None:
            return
        x, W = self.x, self.weight
        K, N = W.w.shape
        if self.residual is not None and self.residual.requires_grad:
            # the skip path passes the fp32 stream gradient on; the branch below uses the bf16 copy
            if self.residual.g is not None:
                # a skip input with another consumer whose gradient arrived first (CeiT: a block's
                # output also feeds the class-token list, ceit.py:66-70)
                _add_grad(self.residual, self.g.view(self.residual.v.shape))
            else:
                self.residual.g = self.g.view(self.residual.v.shape)
        scaled = self.col_scale is not None or self.row_scale is not None
        if scaled:
            # reverse mode of the LayerScale / drop-path epilogue: the branch gradient is the fp32
            # stream gradient times both scales (one pass, bf16 out = the MMA operand)
            if self.residual is None or self.g is None or self.g.dtype != F32:
                raise NotImplementedError("LayerScale / drop-path epilogue without a residual add")
            self.g16 = None
            dy2 = ops.scale_rows_cols(self.g.reshape(-1, N),
                                      self.col_scale.master if self.col_scale is not None else None,
                                      self.row_scale, self.rows_per_scale)
        else:
            dy2 = _grad16(self).reshape(-1, N)
        self.g = None
        if self.act and not self.g_is_pre:
            raise NotImplementedError("activated linear output must be consumed by a linear op")
        x2 = _as2d(x.v, K)
        # dW[K,N] += X^T dY : both operands token-major (MN-major), split-K over tokens, TMA
        # reduce-add into the (zeroed) fp32 gradient arena
        W.touch()
        ops.gemm(x2, dy2, a_mn=True, b_mn=True, out=W.gw, accumulate=True, split_k=0)
        if self.bias is not None:
            done = getattr(self, "bias_grad_from", None)
            self.bias_grad_from = None
            if scaled or done is None or done.data_ptr() != dy2.data_ptr() or done.numel() != dy2.numel():
                self.bias.touch_grad()
                ops.colsum(dy2, out=self.bias.grad, accumulate=True)
        if self.col_scale is not None:
            # d layerscale from the finished dW / dbias (elementwise.cu), no saved activation
            if W.wm is None:
                raise NotImplementedError("LayerScale backward needs the fp32 master weight view")
            self.col_scale.touch_grad()
            ops.layerscale_grad(W.wm, W.gw, self.bias.master if self.bias is not None else None,
                                self.bias.grad if self.bias is not None else None,
                                self.col_scale.master, self.col_scale.grad)
        if x.requires_grad:
            # dX[M,K] = dY W^T (x gelu'(pre) when x is a fused-GELU output)
            # When x is a fused-GELU Dense output, this GEMM's result IS that Dense's dY (the gradient
            # of its pre-activation), so its bias gradient -- the column sums of the result -- is
            # accumulated by the same epilogue (a 16-shuffle transpose-reduce per slab) instead of a
            # separate 33 us pass over the 155 MB tensor.
            fuse_bias = (_FUSE_DGELU_COLSUM and x.pre_act is not None and isinstance(x, LazyLinear)
                         and x.g is None and x.bias is not None and x.col_scale is None and x.row_scale is None)
            if fuse_bias:
                x.bias.touch_grad()
            dx = ops.gemm(dy2, W.w, b_mn=False, aux_in=x.pre_act,
                          dgelu_colsum=x.bias.grad if fuse_bias else None)
            _add_grad(x, dx.view(x.v.shape) if x.v.is_contiguous() or dx.numel() == x.v.numel() else dx)
            x.g_is_pre = x.pre_act is not None
            if fuse_bias:
                x.bias_grad_from = dx


def _add_grad(var, g):
    if var.g is None:
        var.g = g
    elif g.dtype == BF16 and var.g.dtype == BF16:
        var.g = ops.add(var.g.contiguous(), g.contiguous()).view(var.g.shape)
    elif g.dtype == F32 and var.g.dtype == F32:
        # fan-in of fp32 stream gradients (CaiT: the patch tokens feed every class-attention block)
        C = g.shape[-1]
        ops.copy2d_f32(g.reshape(-1, C), var.g.reshape(-1, C), accumulate=True)
        var.g16 = None
    else:
        raise NotImplementedError("gradient fan-in with mixed precisions")


def linear(x, weight, bias=None, out_features_shape=None):
    lead = tuple(x.shape[:-1])
    N = weight.w.shape[1]
    return LazyLinear(x, weight, bias, lead + (N,))


def gelu(x):
    """jax.nn.gelu (tanh form) -- only as the epilogue of the producing linear (ff.py:24-25)."""
    if isinstance(x, TokenMix) and not x._done and x.layers and x.layers[-1][2] == 0:
        x.layers[-1][2] = 1
        return x
    if isinstance(x, LazyLinear) and not x._done and x.act == 0 and x.residual is None:
        x.act = 1
        return x
    # anything else (an activation that is not a pending Dense output): the standalone kernel pair
    # on the materialised bf16 tensor (sav_act_fwd / sav_act_bwd, kind 4)
    if x.dtype != BF16:
        raise TypeError("gelu: a bf16 activation expected (apply LayerNorm to the fp32 residual stream first)")
    return activation(x, "gelu")


def to_stream(x):
    """Start the fp32 residual stream from a Dense output that has no skip input (MLP-Mixer: the
    patch embedding itself is the stream, mlp_mixer.py:46-49): the pending GEMM writes fp32."""
    if isinstance(x, LazyLinear) and not x._done and x.act == 0 and x.residual is None \
            and x.col_scale is None and x.row_scale is None:
        x.f32_out = True
        return x
    raise NotImplementedError("to_stream: apply it directly to a Dense output")


def scale_channels(x, scale_param):
    """LayerScale (layerscale.py:20-23) folded into the producing GEMM."""
    if isinstance(x, LazyLinear) and not x._done and x.col_scale is None and x.residual is None:
        x.col_scale = scale_param
        return x
    raise NotImplementedError("standalone LayerScale: apply it directly to a Dense output")


def scale_rows(x, per_sample_scale, rows_per_sample):
    """Stochastic depth (stochastic_depth.py:16-27) folded into the producing GEMM."""
    if isinstance(x, LazyLinear) and not x._done and x.row_scale is None and x.residual is None:
        x.row_scale, x.rows_per_scale = per_sample_scale, rows_per_sample
        return x
    raise NotImplementedError("standalone drop-path: apply it directly to a Dense output")


def add(a, b):
    """Residual add (vit.py:24,31): folded into the pending linear's epilogue (fp32 stream out)."""
    for lazy, other in ((a, b), (b, a)):
        if isinstance(lazy, TokenMix) and not lazy._done and lazy.closed and isinstance(other, Var) \
                and tuple(other.shape) == tuple(lazy.shape):
            return lazy.run(other)     # mlp_mixer.py:24: the transpose back carries the add
    for lazy, other in ((a, b), (b, a)):
        if isinstance(lazy, LazyLinear) and not lazy._done and lazy.residual is None and \
                lazy.act == 0 and isinstance(other, Var) and tuple(other.shape) == tuple(lazy.shape):
            lazy.residual = other
            if not lazy.ln_fusable():
                lazy.v  # materialise now so program order is preserved
            elif current_tape() is not None:
                current_tape().defer(lazy)
            # else: stays pending until its first reader -- a LayerNorm takes it into one launch,
            # anything else (`.v`) materialises it as before; either way it is recorded on the tape
            # before that reader
            return lazy
    return _plain_add(a, b)


def _plain_add(a, b):
    """a + b of two materialised Vars of one shape and dtype (no pending Dense to fold the add into):
    one elementwise kernel; the gradient fans out to both inputs unchanged."""
    if not (isinstance(a, Var) and isinstance(b, Var)) or tuple(a.shape) != tuple(b.shape):
        raise TypeError("add: two activations of the same shape expected")
    ta, tb = a.v, b.v
    if ta.dtype != tb.dtype:
        raise TypeError("add: mixed precisions (cast the fp32 residual stream with LayerNorm / "
                        "stream_to_operand first)")
    if _meta(ta):
        return Var(torch.empty(ta.shape, dtype=ta.dtype, device="meta"))
    if ta.dtype == BF16:
        out = Var(ops.add(ta.contiguous(), tb.contiguous()).view(ta.shape))
    else:
        y = torch.empty(ta.shape, dtype=F32, device=ta.device)
        C = ta.shape[-1]
        ops.copy2d_f32(ta.contiguous().view(-1, C), y.view(-1, C))
        ops.copy2d_f32(tb.contiguous().view(-1, C), y.view(-1, C), accumulate=True)
        out = Var(y)

    def bwd():
        g = out.take_grad()
        out.g16 = None
        if g is None:
            return
        for var in (a, b):
            if var.requires_grad:
                if var.g is None and g.dtype == F32:
                    # fp32 gradients are accumulated in place by later consumers: hand out a copy
                    gc = torch.empty(g.shape, dtype=F32, device=g.device)
                    ops.copy2d_f32(g.contiguous().view(-1, g.shape[-1]), gc.view(-1, g.shape[-1]))
                    var.g = gc
                else:
                    _add_grad(var, g)
    record(bwd)
    return out


class LazyLayerNorm(Var):
    """LayerNorm over the fp32 residual stream -> bf16.  Materialised over all rows when `.v` is
    read; `x[:, 0]` on a still-pending instance normalises only the class-token rows."""
    __slots__ = ("x", "scale", "bias", "eps", "_done")

    def __init__(self, x, scale, bias, eps):
        super().__init__(None)
        self.x, self.scale, self.bias, self.eps = x, scale, bias, eps
        self._done = False

    @property
    def shape(self):
        return tuple(self.x.shape)

    @property
    def ndim(self):
        return len(self.x.shape)

    @property
    def dtype(self):
        return BF16

    @property
    def v(self):
        if not self._done:
            self._materialize()
        return self._v

    def _check(self, xv):
        if xv.dtype != F32:
            raise TypeError("layernorm: input must be the fp32 residual stream")

    def _materialize(self):
        self._done = True
        x = self.x
        fused = None
        if isinstance(x, LazyLinear) and x.ln_fusable() and not _meta(x.x.v):
            fused = x._materialize(ln=(self.scale.master, self.bias.master, self.eps))
        xv = x.v
        if _meta(xv):
            self._v = torch.empty(xv.shape, dtype=BF16, device="meta")
            return
        self._check(xv)
        xc = xv.contiguous()
        C = xc.shape[-1]
        x2 = xc.view(-1, C)
        if fused is not None:
            y, mean, rstd = fused
        else:
            y, mean, rstd = ops.layernorm_fwd(x2, self.scale.master, self.bias.master, self.eps)
        self._v = y.view(xc.shape)
        scale, bias = self.scale, self.bias

        def bwd():
            dy = self.take_grad()
            if dy is None:
                return
            scale.touch_grad()
            bias.touch_grad()
            prev = x.take_grad() if x.requires_grad else None
            x.g16 = None
            # When the stream was written by a residual-fused Linear (out-proj / FC2), the bf16 copy
            # produced here IS that Linear's dY, so its bias gradient (column sums of dY) is
            # accumulated by this kernel instead of a separate pass over dY.
            fuse = (_FUSE_BIAS_COLSUM and x.requires_grad and isinstance(x, LazyLinear) and x.bias is not None
                    and x.col_scale is None and x.row_scale is None and x.v.is_contiguous())
            if fuse:
                x.bias.touch_grad()
            dx, dx16 = ops.layernorm_bwd(dy.contiguous().view(-1, C), x2, scale.master, mean, rstd,
                                         scale.grad, bias.grad,
                                         dx_add=prev.contiguous().view(-1, C) if prev is not None else None,
                                         dx16_colsum=x.bias.grad.reshape(C) if fuse else None)
            if x.requires_grad:
                x.g, x.g16 = dx.view(xc.shape), dx16.view(xc.shape)
                if fuse:
                    x.bias_grad_from = dx16
        record(bwd)

    def first_token(self):
        """LayerNorm(x)[:, 0] without normalising the other tokens (vit.py:57,95)."""
        self._done = True
        x, xv = self.x, self.x.v
        B, N, C = xv.shape
        if _meta(xv):
            return Var(torch.empty((B, C), dtype=BF16, device="meta"))
        self._check(xv)
        xc = xv.contiguous()
        rows = xc.view(B * N, C)[::N]          # [B, C] view, row pitch N*C
        y, mean, rstd = ops.layernorm_fwd(rows, self.scale.master, self.bias.master, self.eps)
        out = Var(y)
        scale, bias = self.scale, self.bias

        def bwd():
            dy = out.take_grad()
            if dy is None or not x.requires_grad:
                return
            if x.g is not None:
                raise NotImplementedError("class-token LayerNorm on a stream Var with another consumer")
            scale.touch_grad()
            bias.touch_grad()
            # the stream gradient is zero except on the class-token rows
            g = torch.empty((B * N, C), dtype=F32, device=xc.device)
            g16 = torch.empty((B * N, C), dtype=BF16, device=xc.device)
            ops.fill(g, 0.0)
            ops.fill(g16.view(F32), 0.0)
            ops.layernorm_bwd(dy.contiguous(), rows, scale.master, mean, rstd, scale.grad, bias.grad,
                              dx=g[::N], dx16=g16[::N])
            x.g, x.g16 = g.view(B, N, C), g16.view(B, N, C)
        record(bwd)
        return out


def layernorm(x, scale, bias, eps=1e-6):
    return LazyLayerNorm(x, scale, bias, eps)


def self_attention_core(qkv, num_heads, head_ch):
    """attention.py:39-58 on a fused QKV activation [B,N,3*H*d] -> [B,N,H*d]."""
    t = qkv.v
    B, N, _ = t.shape
    H, D = num_heads, head_ch
    if _meta(t):
        return Var(torch.empty((B, N, H * D), dtype=BF16, device="meta"))
    t5 = t.view(B, N, 3, H, D)
    q, k, v = t5[:, :, 0], t5[:, :, 1], t5[:, :, 2]
    scale = 1.0 / math.sqrt(D)
    o, lse = ops.attn_fwd(q, k, v, scale)
    out = Var(o[0].view(B, N, H * D))

    def bwd():
        d_o = out.take_grad()
        if d_o is None:
            return
        dqkv = torch.empty_like(t5)
        ops.attn_bwd(q, k, v, o, d_o.contiguous().view(B, N, H, D), lse, dqkv[:, :, 0],
                     dqkv[:, :, 1], dqkv[:, :, 2], scale)
        _add_grad(qkv, dqkv.view(B, N, 3 * H * D))
    record(bwd)
    return out


def cross_attention_core(q, kv, num_heads, head_ch):
    """q [B,Nq,H*d]; kv fused [B,Nk,2*H*d] (keys then values) -> [B,Nq,H*d]."""
    tq, tkv = q.v, kv.v
    B, Nq, _ = tq.shape
    Nk = tkv.shape[1]
    H, D = num_heads, head_ch
    if _meta(tq):
        return Var(torch.empty((B, Nq, H * D), dtype=BF16, device="meta"))
    q4 = tq.contiguous().view(B, Nq, H, D)
    kv5 = tkv.contiguous().view(B, Nk, 2, H, D)
    k, v = kv5[:, :, 0], kv5[:, :, 1]
    scale = 1.0 / math.sqrt(D)
    o, lse = ops.attn_fwd(q4, k, v, scale)
    out = Var(o[0].view(B, Nq, H * D))

    def bwd():
        d_o = out.take_grad()
        if d_o is None:
            return
        dq = torch.empty_like(q4)
        dkv = torch.empty_like(kv5)
        ops.attn_bwd(q4, k, v, o, d_o.contiguous().view(B, Nq, H, D), lse, dq, dkv[:, :, 0],
                     dkv[:, :, 1], scale)
        _add_grad(q, dq.view(B, Nq, H * D))
        _add_grad(kv, dkv.view(B, Nk, 2 * H * D))
    record(bwd)
    return out


def patchify(images, ph, pw, hwcn):
    """Raw fp32 image batch -> bf16 patch rows [B, n, ph*pw*c] (no gradient)."""
    if hwcn:
        H, W, c, B = images.shape
    else:
        B, H, W, c = images.shape
    n = (H // ph) * (W // pw)
    if _meta(images):
        return Var(torch.empty((B, n, ph * pw * c), dtype=BF16, device="meta"), requires_grad=False)
    out = ops.patchify(images.contiguous(), ph, pw, hwcn)
    return Var(out.view(B, n, ph * pw * c), requires_grad=False)


def embed(tok, cls, pos):
    """[cls; tok] + pos (vit.py:81-85, position_embed.py:54-56) -> fp32 residual stream;
    cls may be None (cait.py:69)."""
    t = tok.v
    B, n, C = t.shape
    has_cls = cls is not None
    if _meta(t):
        return Var(torch.empty((B, n + int(has_cls), C), dtype=F32, device="meta"))
    x = ops.embed_fwd(t.contiguous(), cls.master if has_cls else None, pos.master, B, n, C)
    out = Var(x)

    def bwd():
        dx = out.take_grad()
        out.g16 = None
        if dx is None:
            return
        pos.touch_grad()
        if has_cls:
            cls.touch_grad()
        dtok = ops.embed_bwd(dx.contiguous(), cls.grad if has_cls else None, pos.grad, B, n, C,
                             accumulate=True)
        if tok.requires_grad:
            _add_grad(tok, dtok.view(B, n, C))
    record(bwd)
    return out


def getitem(x, idx):
    """Only the class-token select x[:, 0] (vit.py:95) is needed on the hot path."""
    if not (isinstance(idx, tuple) and len(idx) == 2 and idx[0] == slice(None) and idx[1] == 0):
        raise NotImplementedError("Var indexing supports x[:, 0] only")
    if isinstance(x, LazyLayerNorm) and not x._done:
        return x.first_token()
    t = x.v
    B, N, C = t.shape
    if _meta(t):
        return Var(torch.empty((B, C), dtype=t.dtype, device="meta"))
    if t.dtype != BF16:
        raise NotImplementedError("x[:, 0] on the fp32 stream: apply LayerNorm first")
    out = Var(t[:, 0])  # strided [B,C] view (row stride N*C): a valid GEMM A operand, no copy

    def bwd():
        g = out.take_grad()
        if g is None:
            return
        full = torch.empty((B, N, C), dtype=BF16, device=t.device)
        ops.fill(full.view(torch.float32), 0.0)
        ops.copy2d(g.reshape(B, C), full[:, 0])
        _add_grad(x, full)
    record(bwd)
    return out


class LossVar:
    """Scalar loss living on the device; .seed() installs dloss/dlogits, .item() syncs."""

    def __init__(self, logits, dlogits, stats, batch):
        self.logits, self.dlogits, self.stats, self.batch = logits, dlogits, stats, batch

    def seed(self):
        self.logits.g = self.dlogits

    def item(self):
        return float(self.stats[0].item()) / self.batch

    def top1(self):
        return float(self.stats[1].item())


def softmax_cross_entropy(logits, labels, label_smoothing=0.0, world_size=1, mix_labels=None, ratio=None):
    """train.py:83-91 (plain mean over the global batch); mix_labels / ratio are the
    optional mixup keys of the batch (train.py:84-87)."""
    lv = logits.v
    B = lv.shape[0]
    dl, stats = ops.softmax_xent(lv.contiguous(), labels, label_smoothing, 1.0 / (B * world_size),
                                 mix_labels, ratio)
    return LossVar(logits, dl, stats, B)


# ------------------------------------------------------------------ CaiT / BoTNet additions
def tile_param_rows(param, B):
    """jnp.tile(cls, [b,1,1]) (cait.py:157-160): fp32 parameter [1,1,C] -> stream Var [B,1,C]."""
    C = param.master.shape[-1]
    m = param.master.reshape(1, C)
    if _meta(m):
        return Var(torch.empty((B, 1, C), dtype=F32, device="meta"))
    out = torch.empty((B, 1, C), dtype=F32, device=m.device)
    ops.copy2d_f32(m.expand(B, C), out.view(B, C))
    res = Var(out)

    def bwd():
        g = res.take_grad()
        res.g16 = None
        if g is None:
            return
        param.touch_grad()
        g16 = torch.empty((B, C), dtype=BF16, device=g.device)
        ops.cast_bf16(g.contiguous().view(B, C), g16)
        ops.colsum(g16, out=param.grad.reshape(C), accumulate=True)
    record(bwd)
    return res


def concat_tokens(cls_tok, x):
    """jnp.concatenate([cls_token, x], axis=1) on the fp32 stream (cait.py:97,175)."""
    tc, tx = cls_tok.v, x.v
    B, N, C = tx.shape
    if _meta(tx):
        return Var(torch.empty((B, N + 1, C), dtype=F32, device="meta"))
    if tc.dtype != F32 or tx.dtype != F32:
        raise TypeError("concat_tokens: both inputs must be fp32 residual-stream tensors")
    out = torch.empty((B, N + 1, C), dtype=F32, device=tx.device)
    o2 = out.view(B, (N + 1) * C)
    ops.copy2d_f32(tc.contiguous().view(B, C), o2[:, :C])
    ops.copy2d_f32(tx.contiguous().view(B, N * C), o2[:, C:])
    res = Var(out)

    def bwd():
        g = res.take_grad()
        res.g16 = None
        if g is None:
            return
        g2 = g.contiguous().view(B, (N + 1) * C)
        for var, sl, shape in ((cls_tok, g2[:, :C], (B, 1, C)), (x, g2[:, C:], (B, N, C))):
            if not var.requires_grad:
                continue
            if var.g is None:
                fresh = torch.empty(shape, dtype=F32, device=g.device)
                ops.copy2d_f32(sl, fresh.view(B, -1))
                var.g, var.g16 = fresh, None
            else:
                ops.copy2d_f32(sl, var.g.view(B, -1), accumulate=True)
                var.g16 = None
    record(bwd)
    return res


def first_token_keepdim(x):
    """inputs[:, 0:1, :] of a bf16 activation (cait.py:13) as a strided view -- no copy."""
    t = x.v
    B, N, C = t.shape
    if _meta(t):
        return Var(torch.empty((B, 1, C), dtype=t.dtype, device="meta"))
    if t.dtype != BF16:
        raise NotImplementedError("first_token_keepdim on the fp32 stream: apply LayerNorm first")
    out = Var(t[:, 0:1])

    def bwd():
        g = out.take_grad()
        if g is None:
            return
        full = torch.empty((B, N, C), dtype=BF16, device=t.device)
        ops.fill(full.view(torch.float32), 0.0)
        ops.copy2d(g.reshape(B, C), full[:, 0])
        _add_grad(x, full)
    record(bwd)
    return out


def _mat_attention(q4, k4, v4, scale, mix, rel_tab, rel_hw, out_shape, grads_cb):
    """Materialised-logits attention (attn_mat.cu): S = q k^T (batched GEMM) -> mix / relative
    logits / softmax / mix rows -> O = P v (batched GEMM).  q4/k4/v4: [B,N,H,D] bf16 views."""
    B, Nq, H, D = q4.shape
    mix_pre = mix[0].master if mix is not None else None
    mix_post = mix[1].master if mix is not None else None
    rel = None
    if rel_tab is not None:
        # per-query relative logits q . r for all 2W-1 + 2Hh-1 offsets: one skinny GEMM
        # (botnet.py:96-111, the einsum 'b h H W d, r d -> b h H W r' for both axes at once)
        rel = ops.gemm(q4.reshape(B * Nq * H, D), rel_tab.w, b_mn=False, out_fp32=True)
        rel = rel.view(B, Nq, H, rel_tab.w.shape[0])
    o, S, P = ops.attn_mat_fwd(q4, k4, v4, scale, mix_pre, mix_post, rel, rel_hw)
    out = Var(o.view(out_shape))

    def bwd():
        d_o = out.take_grad()
        if d_o is None:
            return
        dq4, dk4, dv4 = grads_cb()
        dmix_pre = dmix_post = None
        if mix is not None:
            mix[0].touch_grad()
            mix[1].touch_grad()
            dmix_pre, dmix_post = mix[0].grad, mix[1].grad
        drel = ops.attn_mat_bwd(q4, k4, v4, S, P, d_o.contiguous().view(B, Nq, H, D), dq4, dk4, dv4,
                                scale, mix_pre, mix_post, dmix_pre, dmix_post, rel, rel_hw)
        if drel is not None:
            R = rel_tab.w.shape[0]
            drel16 = torch.empty((B * Nq * H, R), dtype=BF16, device=drel.device)
            ops.cast_bf16(drel.view(-1), drel16.view(-1))
            # d table[R, D] += drel^T q ;  dq += drel table
            rel_tab.touch()
            ops.gemm(drel16, q4.reshape(B * Nq * H, D), a_mn=True, b_mn=True, out=rel_tab.gw,
                     accumulate=True, split_k=0)
            dq_rel = ops.gemm(drel16, rel_tab.w, b_mn=True)
            dq_flat = dq4.reshape(-1)
            return ops.add(dq_flat, dq_rel.view(-1))   # only the contiguous-q (cross) form has rel
        return None
    return out, bwd


def mat_self_attention_core(qkv, num_heads, head_ch, mix=None):
    """attention.py:39-58 with talking heads (attention.py:44-52) or a head dim other than 64, on a
    fused QKV activation [B,N,3*H*d]."""
    t = qkv.v
    B, N, _ = t.shape
    H, D = num_heads, head_ch
    if _meta(t):
        return Var(torch.empty((B, N, H * D), dtype=BF16, device="meta"))
    t5 = t.view(B, N, 3, H, D)
    hold = {}

    def grads():
        hold["dqkv"] = torch.empty_like(t5)
        d = hold["dqkv"]
        return d[:, :, 0], d[:, :, 1], d[:, :, 2]

    out, bwd = _mat_attention(t5[:, :, 0], t5[:, :, 1], t5[:, :, 2], 1.0 / math.sqrt(D), mix, None,
                              None, (B, N, H * D), grads)

    def bwd2():
        bwd()
        if "dqkv" in hold:
            _add_grad(qkv, hold.pop("dqkv").view(B, N, 3 * H * D))
    record(bwd2)
    return out


def mat_cross_attention_core(q, kv, num_heads, head_ch, mix=None, rel_tab=None, rel_hw=None):
    """q [B,Nq,H*d]; kv fused [B,Nk,2*H*d] -> [B,Nq,H*d] through the materialised path (class
    attention cait.py:10-15; BoTNet MHSA with 2-D relative logits botnet.py:158-199)."""
    tq, tkv = q.v, kv.v
    B, Nq, _ = tq.shape
    Nk = tkv.shape[1]
    H, D = num_heads, head_ch
    if _meta(tq):
        return Var(torch.empty((B, Nq, H * D), dtype=BF16, device="meta"))
    q4 = _as2d(tq, H * D).view(B, Nq, H, D) if not tq.is_contiguous() else tq.view(B, Nq, H, D)
    if not q4.is_contiguous():
        q4 = q4.contiguous()
    kv5 = tkv.contiguous().view(B, Nk, 2, H, D)
    hold = {}

    def grads():
        hold["dq"] = torch.empty((B, Nq, H, D), dtype=BF16, device=tq.device)
        hold["dkv"] = torch.empty_like(kv5)
        return hold["dq"], hold["dkv"][:, :, 0], hold["dkv"][:, :, 1]

    out, bwd = _mat_attention(q4, kv5[:, :, 0], kv5[:, :, 1], 1.0 / math.sqrt(D), mix, rel_tab,
                              rel_hw, (B, Nq, H * D), grads)

    def bwd2():
        dq_total = bwd()
        if "dq" in hold:
            dq = hold.pop("dq")
            if dq_total is not None:
                dq = dq_total.view(B, Nq, H, D)
            _add_grad(q, dq.view(B, Nq, H * D))
            _add_grad(kv, hold.pop("dkv").view(B, Nk, 2 * H * D))
    record(bwd2)
    return out


def reshape(x, shape):
    """View a contiguous activation under another shape (e.g. 'b H W c -> b (H W) c',
    botnet.py:175-183); the gradient is viewed back."""
    t = x.v
    if _meta(t):
        return Var(torch.empty(t.shape, dtype=t.dtype, device="meta").reshape(shape))
    old = tuple(t.shape)
    out = Var(t.contiguous().view(shape))

    def bwd():
        g = out.take_grad()
        if g is not None:
            _add_grad(x, g.contiguous().view(old))
    record(bwd)
    return out


# ------------------------------------------------------------------ CvT additions
def dwconv_bn(x, w, gamma, beta, stride, eps, running=None, momentum=0.9, training=True, update=False):
    """Depthwise 3x3 conv (SAME, no bias) + BatchNorm on the fp32 stream [B,H,W,C] -> bf16
    [B,Ho,Wo,C] (cvt_attention.py:25-36).  training: batch statistics, and -- when `update` (the
    'batch_stats' collection is mutable) -- running = (ra_mean, ra_var) <- momentum * ra +
    (1 - momentum) * batch statistic (flax nn.BatchNorm).  Eval: normalise with the running averages."""
    t = x.v
    B, H, W, C = t.shape
    if _meta(t):
        ho, wo = -(-H // stride), -(-W // stride)
        return Var(torch.empty((B, ho, wo, C), dtype=BF16, device="meta"))
    if t.dtype != F32:
        raise TypeError("dwconv_bn: input must be the fp32 residual stream")
    tc = t.contiguous()
    w9 = w.master.reshape(9, C)
    y, stats = ops.dwconv3x3_fwd(tc, w9, stride)
    if not training:
        if running is None:
            raise ValueError("BatchNorm in eval mode needs the 'batch_stats' running averages")
        # use_running_average=True (cvt_attention.py:33): no statistics, no state update
        return Var(ops.batchnorm_eval(y, running[0], running[1], gamma.master, beta.master, eps),
                   requires_grad=False)
    z, mrv = ops.batchnorm_fwd(y, stats, gamma.master, beta.master, eps)
    if update and running is not None:
        ops.bn_update_running(mrv[0], mrv[2], running[0], running[1], momentum)
    out = Var(z)

    def bwd():
        dz = out.take_grad()
        if dz is None:
            return
        for p in (w, gamma, beta):
            p.touch_grad()
        dy = ops.batchnorm_bwd(dz.contiguous(), y, mrv, gamma.master, gamma.grad, beta.grad)
        if x.requires_grad:
            fresh = x.g is None
            if fresh:
                x.g = torch.empty_like(tc)
            dx = x.g if x.g.is_contiguous() else None
            if dx is None:
                raise NotImplementedError("dwconv backward into a strided stream gradient")
            ops.dwconv3x3_bwd(tc, w9, dy, dx, w.grad.reshape(9, C), stride, not fresh)
            x.g16 = None
        else:
            scratch = torch.empty_like(tc)
            ops.dwconv3x3_bwd(tc, w9, dy, scratch, w.grad.reshape(9, C), stride, False)
    record(bwd)
    return out


def attention3(q, k, v, num_heads, head_ch):
    """attention.py:39-58 / cvt_attention.py:85-104 on three separate projections q [B,Nq,H*d],
    k, v [B,Nk,H*d] (Nq != Nk when the key / value projections are strided)."""
    tq, tk, tv = q.v, k.v, v.v
    B, Nq, _ = tq.shape
    Nk = tk.shape[1]
    H, D = num_heads, head_ch
    if _meta(tq):
        return Var(torch.empty((B, Nq, H * D), dtype=BF16, device="meta"))
    q4, k4, v4 = (t.contiguous().view(B, -1, H, D) for t in (tq, tk, tv))
    scale = 1.0 / math.sqrt(D)
    fused = D == 64
    if fused:
        o, lse = ops.attn_fwd(q4, k4, v4, scale)
        out = Var(o[0].view(B, Nq, H * D))
    else:
        o, S, P = ops.attn_mat_fwd(q4, k4, v4, scale)
        out = Var(o.view(B, Nq, H * D))

    def bwd():
        d_o = out.take_grad()
        if d_o is None:
            return
        dq, dk, dv = torch.empty_like(q4), torch.empty_like(k4), torch.empty_like(v4)
        d4 = d_o.contiguous().view(B, Nq, H, D)
        if fused:
            ops.attn_bwd(q4, k4, v4, o, d4, lse, dq, dk, dv, scale)
        else:
            ops.attn_mat_bwd(q4, k4, v4, S, P, d4, dq, dk, dv, scale)
        _add_grad(q, dq.view(B, Nq, H * D))
        _add_grad(k, dk.view(B, Nk, H * D))
        _add_grad(v, dv.view(B, Nk, H * D))
    record(bwd)
    return out


def pad_tokens_to_square(x):
    """cvt.py:10-16 zero_pad_and_reshape on the fp32 stream: [B,L,C] -> ([B,S*S,C], S) with
    S = ceil(sqrt(L)), zero tokens appended."""
    t = x.v
    B, L, C = t.shape
    S = int(math.ceil(math.sqrt(L)))
    if S * S == L:
        return x, S
    if _meta(t):
        return Var(torch.empty((B, S * S, C), dtype=t.dtype, device="meta")), S
    if t.dtype != F32:
        raise TypeError("pad_tokens_to_square: fp32 residual stream expected")
    out = torch.empty((B, S * S, C), dtype=F32, device=t.device)
    ops.fill(out, 0.0)
    ops.copy2d_f32(t.contiguous().view(B, L * C), out.view(B, S * S * C)[:, :L * C])
    res = Var(out)

    def bwd():
        g = res.take_grad()
        res.g16 = None
        if g is None or not x.requires_grad:
            return
        sl = g.contiguous().view(B, S * S * C)[:, :L * C]
        if x.g is None:
            x.g = torch.empty((B, L, C), dtype=F32, device=g.device)
            ops.copy2d_f32(sl, x.g.view(B, L * C))
        else:
            ops.copy2d_f32(sl, x.g.view(B, L * C), accumulate=True)
        x.g16 = None
    record(bwd)
    return res, S


# ------------------------------------------------------------------ rotary helpers
def rotary(x, seq_axis=0):
    """apply_rotary_pos_emb (position_embed.py:17-20) with sin / cos of the FixedPositionalEmbedding
    frequencies; the gradient is rotated back (the rotation is orthogonal)."""
    t = x.v
    if _meta(t):
        return Var(torch.empty(t.shape, dtype=t.dtype, device="meta"))
    out = Var(ops.rotary(t.contiguous(), seq_axis))

    def bwd():
        g = out.take_grad()
        if g is not None and x.requires_grad:
            _add_grad(x, ops.rotary(g.contiguous(), seq_axis, inverse=True))
    record(bwd)
    return out


# ------------------------------------------------------------------ dense convolutions / conv stems
def _image_geometry(x):
    """(tensor, B, H, W, C, element strides (b,h,w,c), producing Var or None) of a conv input:
    train.HWCN images [H,W,C,N], a raw [B,H,W,C] tensor, or a contiguous Var [B,H,W,C]."""
    from .train import HWCN
    if isinstance(x, HWCN):
        t = x.images
        H, W, Cc, B = t.shape
        return t, B, H, W, Cc, (1, W * Cc * B, Cc * B, B), None
    if isinstance(x, Var):
        t = x.v
        if not _meta(t):
            t = t.contiguous()
        B, H, W, Cc = t.shape
        return t, B, H, W, Cc, (H * W * Cc, W * Cc, Cc, 1), (x if x.requires_grad else None)
    t = x if _meta(x) else x.contiguous()
    B, H, W, Cc = t.shape
    return t, B, H, W, Cc, (H * W * Cc, W * Cc, Cc, 1), None


def conv2d_stream(x, kernel, bias, ksize, stride, padding="SAME"):
    """nn.Conv (cvt.py:27-31) as im2col + tcgen05 GEMM: x [B,H,W,Cin] (fp32 stream Var, raw images,
    or HWCN images) -> fp32 Var [B, Ho, Wo, Cout].  kernel: Param [kh,kw,cin,cout]; bias: Param or
    None.  Reverse mode: dW = cols^T dY (split-K GEMM into the gradient arena), dbias = column sums,
    dX = fold(dY W^T) (GEMM + gather col2im)."""
    t, B, H, W, Cin, xs, src = _image_geometry(x)
    Cout = kernel.shadow.shape[-1]
    K = ksize * ksize * Cin
    if _meta(t):
        if padding == "SAME":
            Ho, Wo = -(-H // stride), -(-W // stride)
        else:
            Ho = (H + 2 * int(padding) - ksize) // stride + 1
            Wo = (W + 2 * int(padding) - ksize) // stride + 1
        return Var(torch.empty((B, Ho, Wo, Cout), dtype=F32, device="meta"))
    g = ops.conv2d_geom(B, H, W, Cin, ksize, stride, padding, xs)
    cols = ops.im2col(t, g)
    A = cols[:, :K]
    W2 = kernel.shadow.reshape(K, Cout)
    y = ops.gemm(A, W2, b_mn=True, out_fp32=True, bias=bias.master if bias is not None else None)
    out = Var(y.view(B, g.Ho, g.Wo, Cout))

    def bwd():
        dy = out.take_grad()
        dy16 = out.take_grad16()
        if dy is None and dy16 is None:
            return
        if dy16 is None:
            dy16 = torch.empty((B * g.Ho * g.Wo, Cout), dtype=BF16, device=t.device)
            ops.cast_bf16(dy.contiguous().view(-1), dy16.view(-1))
        dy16 = dy16.reshape(-1, Cout)
        kernel.touch_grad()
        ops.gemm(A, dy16, a_mn=True, b_mn=True, out=kernel.grad.reshape(K, Cout), accumulate=True, split_k=0)
        if bias is not None:
            bias.touch_grad()
            ops.colsum(dy16, out=bias.grad, accumulate=True)
        if src is not None:
            dcols = torch.empty((B * g.Ho * g.Wo, g.ld), dtype=BF16, device=t.device)
            ops.gemm(dy16, W2, b_mn=False, out=dcols[:, :K])
            dx = torch.empty((B, H, W, Cin), dtype=F32, device=t.device)
            ops.col2im(dcols, g, dx)
            if src.g is None:
                src.g, src.g16 = dx.view(src.v.shape), None
            else:
                _add_grad(src, dx.view(src.v.shape))
    record(bwd)
    return out


def layernorm_stream(x, scale, bias, eps=1e-6):
    """LayerNorm whose OUTPUT is the fp32 residual stream (cvt.py:34: the token embedding's LayerNorm
    feeds StageBlock's `x + attention(x)` directly).  fp32 in, fp32 out."""
    xv = x.v
    if _meta(xv):
        return Var(torch.empty(xv.shape, dtype=F32, device="meta"))
    if xv.dtype != F32:
        raise TypeError("layernorm_stream: fp32 input expected")
    xc = xv.contiguous()
    C = xc.shape[-1]
    x2 = xc.view(-1, C)
    y, mean, rstd = ops.layernorm_fwd_f32(x2, scale.master, bias.master, eps)
    out = Var(y.view(xc.shape))

    def bwd():
        dy = out.take_grad()
        out.g16 = None
        if dy is None:
            return
        scale.touch_grad()
        bias.touch_grad()
        dx, dx16 = ops.layernorm_bwd_f32(dy.contiguous().view(-1, C), x2, scale.master, mean, rstd,
                                         scale.grad, bias.grad)
        if x.requires_grad:
            if x.g is not None:
                raise NotImplementedError("layernorm_stream input with a second consumer")
            x.g, x.g16 = dx.view(xc.shape), dx16.view(xc.shape)
    record(bwd)
    return out


def stream_first_token(x):
    """x[:, 0] of the fp32 residual stream as a bf16 GEMM operand (cvt.py:165-170: the CvT head reads
    the class token without a LayerNorm in between)."""
    t = x.v
    B, N, C = t.shape
    if _meta(t):
        return Var(torch.empty((B, C), dtype=BF16, device="meta"))
    if t.dtype != F32:
        raise TypeError("stream_first_token: fp32 residual stream expected")
    tc = t.contiguous()
    rows = tc.view(B * N, C)[::N]
    out = Var(ops.scale_rows_cols(rows, None, None, 1))

    def bwd():
        g = out.take_grad()
        if g is None or not x.requires_grad:
            return
        g2 = g.reshape(B, C).contiguous()
        if x.g is not None:
            # the stream has later consumers whose gradients arrived first (CeiT: every block's
            # class token is collected while the stream goes on to the next block, ceit.py:66-70)
            if not x.g.is_contiguous():
                x.g = x.g.contiguous()
            ops.cast2d_bf16_f32(g2, x.g.view(B * N, C)[::N], accumulate=True)
            x.g16 = None
            return
        full = torch.empty((B * N, C), dtype=F32, device=t.device)
        full16 = torch.empty((B * N, C), dtype=BF16, device=t.device)
        ops.fill(full, 0.0)
        ops.fill(full16.view(torch.float32), 0.0)
        g2 = g.reshape(B, C).contiguous()
        ops.cast2d_bf16_f32(g2, full[::N])
        ops.copy2d(g2, full16[::N])
        x.g, x.g16 = full.view(B, N, C), full16.view(B, N, C)
    record(bwd)
    return out


# ------------------------------------------------------------------ BoTNet conv stages (bf16 NHWC)
def conv2d(x, kernel, ksize, stride, padding="SAME", bias=None):
    """nn.Conv(use_bias=False) of the BoTNet stem / bottleneck blocks (botnet.py:280-285,47-64):
    x bf16 Var [B,H,W,Cin] (or raw / HWCN fp32 images for the stem) -> bf16 Var [B,Ho,Wo,Cout].
    1x1 stride-1 convs are the GEMM itself; everything else goes through im2col / col2im."""
    t, B, H, W, Cin, xs, src = _image_geometry(x)
    Cout = kernel.shadow.shape[-1]
    K = ksize * ksize * Cin
    if _meta(t):
        if padding == "SAME":
            Ho, Wo = -(-H // stride), -(-W // stride)
        else:
            Ho = (H + 2 * int(padding) - ksize) // stride + 1
            Wo = (W + 2 * int(padding) - ksize) // stride + 1
        return Var(torch.empty((B, Ho, Wo, Cout), dtype=BF16, device="meta"))
    W2 = kernel.shadow.reshape(K, Cout)
    direct = ksize == 1 and stride == 1 and t.dtype == BF16 and isinstance(x, Var)
    if direct:
        g = None
        A = t.view(B * H * W, Cin)
        Ho, Wo = H, W
    else:
        g = ops.conv2d_geom(B, H, W, Cin, ksize, stride, padding, xs)
        A = ops.im2col(t, g)[:, :K]
        Ho, Wo = g.Ho, g.Wo
    y = ops.gemm(A, W2, b_mn=True, bias=bias.master if bias is not None else None)
    out = Var(y.view(B, Ho, Wo, Cout))

    def bwd():
        dy = out.take_grad()
        if dy is None:
            return
        dy2 = dy.contiguous().view(-1, Cout)
        kernel.touch_grad()
        if bias is not None:
            bias.touch_grad()
            ops.colsum(dy2, out=bias.grad, accumulate=True)
        ops.gemm(A, dy2, a_mn=True, b_mn=True, out=kernel.grad.reshape(K, Cout), accumulate=True, split_k=0)
        if src is None:
            return
        if direct:
            dx = ops.gemm(dy2, W2, b_mn=False)
        elif K % 8:
            raise NotImplementedError("conv2d input gradient with kh*kw*cin not a multiple of 8 (only "
                                      "the 3-channel image stems have that, and images carry no gradient)")
        else:
            dcols = torch.empty((B * Ho * Wo, g.ld), dtype=BF16, device=t.device)
            ops.gemm(dy2, W2, b_mn=False, out=dcols[:, :K])
            dx = torch.empty((B, H, W, Cin), dtype=t.dtype, device=t.device)
            ops.col2im(dcols, g, dx)
        _add_grad(src, dx.view(src.v.shape))
    record(bwd)
    return out


def batchnorm_act(y, gamma, beta, eps, act="none", running=None, momentum=0.9, training=True,
                  update=False):
    """act(nn.BatchNorm(y)) on a bf16 [.., C] Var (botnet.py:286-290: batch statistics in training,
    running averages otherwise; momentum update of `running` = (ra_mean, ra_var) when `update`)."""
    t = y.v
    if _meta(t):
        return Var(torch.empty(t.shape, dtype=BF16, device="meta"))
    tc = t.contiguous()
    C = tc.shape[-1]
    rows = tc.numel() // C
    if not training:
        if running is None:
            raise ValueError("BatchNorm in eval mode needs the 'batch_stats' running averages")
        ra_mean, ra_var = running
        z = ops.batchnorm_eval(tc, ra_mean, ra_var, gamma.master, beta.master, eps)
        if act != "none":
            z = ops.act_fwd(z, act)
        return Var(z, requires_grad=False)
    mrv = ops.bn_finalize(ops.bn_stats(tc), rows, eps)
    if update and running is not None:
        ops.bn_update_running(mrv[0], mrv[2], running[0], running[1], momentum)
    out = Var(ops.bn_act_fwd(tc, mrv[0], mrv[1], gamma.master, beta.master, act))

    def bwd():
        dz = out.take_grad()
        if dz is None:
            return
        gamma.touch_grad()
        beta.touch_grad()
        dy = ops.bn_act_bwd(dz.contiguous(), tc, mrv[0], mrv[1], gamma.master, beta.master, gamma.grad,
                            beta.grad, act)
        if y.requires_grad:
            _add_grad(y, dy.view(t.shape))
    record(bwd)
    return out


def gate_residual_act(x, gate=None, residual=None, act="none"):
    """act(residual + x * gate[b, c]) in one pass: the squeeze-excite gate, the skip connection and
    the closing activation of a bottleneck block (botnet.py:54-67, squeeze_excite.py:37)."""
    t = x.v
    if _meta(t):
        return Var(torch.empty(t.shape, dtype=BF16, device="meta"))
    tc = t.contiguous()
    gt = gate.v.contiguous() if gate is not None else None
    rt = residual.v.contiguous() if residual is not None else None
    out = Var(ops.gate_res_act_fwd(tc, gt, rt, act))

    def bwd():
        d = out.take_grad()
        if d is None:
            return
        want_res = residual is not None and residual.requires_grad
        dx, dres, dgate = ops.gate_res_act_bwd(d.contiguous(), tc, gt, rt, act, want_res)
        if x.requires_grad:
            _add_grad(x, dx.view(t.shape))
        if want_res:
            _add_grad(residual, dres.view(residual.v.shape))
        if gate is not None and gate.requires_grad:
            g16 = torch.empty(dgate.shape, dtype=BF16, device=dgate.device)
            ops.cast_bf16(dgate.view(-1), g16.view(-1))
            _add_grad(gate, g16.view(gate.v.shape))
    record(bwd)
    return out


def spatial_mean(x):
    """jnp.mean(x, axis=(1, 2)) of a bf16 [B,H,W,C] Var -> [B,C] (squeeze_excite.py:32, botnet.py:322)."""
    t = x.v
    if _meta(t):
        return Var(torch.empty((t.shape[0], t.shape[-1]), dtype=BF16, device="meta"))
    tc = t.contiguous()
    out = Var(ops.spatial_mean_fwd(tc))

    def bwd():
        d = out.take_grad()
        if d is not None and x.requires_grad:
            _add_grad(x, ops.spatial_mean_bwd(d.contiguous(), tuple(tc.shape)).view(t.shape))
    record(bwd)
    return out


def maxpool3x3s2(x, valid=False):
    """nn.max_pool(x, (3, 3), strides=(2, 2), padding='SAME') (botnet.py:297); valid=True is flax's
    default 'VALID' padding (image_to_token.py:33-37)."""
    t = x.v
    if _meta(t):
        B, H, W, C = t.shape
        if valid:
            return Var(torch.empty((B, (H - 3) // 2 + 1, (W - 3) // 2 + 1, C), dtype=BF16, device="meta"))
        return Var(torch.empty((B, -(-H // 2), -(-W // 2), C), dtype=BF16, device="meta"))
    tc = t.contiguous()
    y, arg = ops.maxpool3x3s2_fwd(tc, valid)
    out = Var(y)

    def bwd():
        d = out.take_grad()
        if d is not None and x.requires_grad:
            _add_grad(x, ops.maxpool3x3s2_bwd(d.contiguous(), arg, tuple(tc.shape), valid).view(t.shape))
    record(bwd)
    return out


def activation(x, kind):
    """Standalone swish / sigmoid / relu on a (small) bf16 Var: the SE hidden layer and gate."""
    t = x.v
    if _meta(t):
        return Var(torch.empty(t.shape, dtype=BF16, device="meta"))
    tc = t.contiguous()
    out = Var(ops.act_fwd(tc, kind))

    def bwd():
        d = out.take_grad()
        if d is not None and x.requires_grad:
            _add_grad(x, ops.act_bwd(tc, d.contiguous(), kind).view(t.shape))
    record(bwd)
    return out


# ------------------------------------------------------------------ MLP-Mixer token mixing
def _pad8(n):
    return (n + 7) // 8 * 8


class TokenMix(Var):
    """The token-mixing branch of a MixerBlock (mlp_mixer.py:18-23), collected lazily:

        x = rearrange(LN(inputs), '... l d -> ... d l'); x = FFBlock(...)(x)
        x = rearrange(x, '... d l -> ... l d'); x = x + inputs

    `swap_tokens_channels` on the LayerNorm output opens it, the two Dense layers and the GELU of the
    FFBlock append to it, the second swap closes it and the residual add runs it: transpose (zero
    padded to a multiple of 8 tokens) -> GEMM + bias + GELU -> GEMM + bias -> transpose back fused
    with the residual add into the fp32 stream.  Shapes seen by module code are the logical ones
    ([B, C, L] / [B, C, hidden]); the padding never leaves this class."""
    __slots__ = ("src", "layers", "closed", "_done", "logical")

    def __init__(self, src):
        super().__init__(None)
        self.src = src
        self.layers = []          # [kernel Param, bias Param or None, act]
        self.closed = False
        self._done = False
        B, L, C = src.shape
        self.logical = (B, C, L)

    @property
    def shape(self):
        if self.closed:
            B, C, L = self.logical
            return (B, L, C)
        return self.logical

    @property
    def ndim(self):
        return 3

    @property
    def dtype(self):
        return self._v.dtype if self._done else BF16

    @property
    def v(self):
        if not self._done:
            self.run(None)
        return self._v

    def dense(self, kernel, bias, features):
        if self.closed or self._done:
            raise NotImplementedError("token mixing: Dense after the closing rearrange")
        self.layers.append([kernel, bias, 0])
        self.logical = self.logical[:2] + (features,)
        return self

    def run(self, residual):
        """Launch the branch; with `residual` (fp32 stream Var [B, L, C]) the result is the new
        fp32 stream, without it the bf16 [B, L, C] activation."""
        self._done = True
        src = self.src
        xv = src.v
        B, L, C = xv.shape
        if len(self.layers) != 2 or self.layers[0][2] != 1 or self.layers[1][2] != 0 or \
                not self.closed or self.logical[2] != L:
            raise NotImplementedError("token mixing: expected Dense -> gelu -> Dense back to the "
                                      "token count (FFBlock, ff.py:24-27)")
        (k0, b0, _), (k1, b1, _) = self.layers
        H = k0.master.shape[1]
        if _meta(xv):
            self._v = torch.empty((B, L, C), dtype=F32 if residual is not None else BF16, device="meta")
            return self
        dev = xv.device
        Lp, Hp = _pad8(L), _pad8(H)
        M = B * C

        def padded_weight(kernel, rows, cols, rp, cp):
            # bf16 W^T [cols -> cp rows, rows -> rp pitch] from the fp32 master [rows, cols]
            wt = torch.zeros((1, cp, rp), dtype=BF16, device=dev)
            ops.transpose_pad(kernel.master.view(1, rows, cols), rows, cols, wt, rp)
            return wt[0]

        def padded_bias(bias, n, npad):
            bp = torch.zeros((npad,), dtype=F32, device=dev)
            if bias is not None:     # (a 1 x n "transpose": no alignment demands on the arena view)
                ops.transpose_pad(bias.master.view(1, 1, n), 1, n, bp.view(1, npad, 1), 1)
            return bp

        w0t = padded_weight(k0, L, H, Lp, Hp)        # [Hp, Lp]  = B operand [N, K] of Dense_0
        w1t = padded_weight(k1, H, L, Hp, Lp)        # [Lp, Hp]
        b0p, b1p = padded_bias(b0, H, Hp), padded_bias(b1, L, Lp)
        xt = torch.empty((B, C, Lp), dtype=BF16, device=dev)
        ops.transpose_pad(xv.contiguous(), L, C, xt, Lp)
        xt2 = xt.view(M, Lp)
        pre = torch.empty((M, Hp), dtype=BF16, device=dev)
        h = ops.gemm(xt2, w0t, bias=b0p, act=1, aux_out=pre)
        y = ops.gemm(h, w1t, bias=b1p)               # [M, Lp]
        res = None
        if residual is not None:
            res = residual.v
            if res.dtype != F32:
                raise TypeError("residual add: the skip input must be the fp32 residual stream")
            res = res.contiguous()
        out = torch.empty((B, L, C), dtype=F32 if res is not None else BF16, device=dev)
        ops.transpose_pad(y.view(B, C, Lp), C, L, out, C, residual=res)
        self._v = out

        def bwd():
            g = self.take_grad()
            g16 = self.take_grad16()
            if g is None and g16 is None:
                return
            if residual is not None and residual.requires_grad:
                if residual.g is not None:
                    raise NotImplementedError("residual-stream Var with a second consumer before its "
                                              "LayerNorm backward")
                residual.g = g
            gin = g16 if g16 is not None else g
            dy = torch.empty((B, C, Lp), dtype=BF16, device=dev)
            ops.transpose_pad(gin.contiguous().view(B, L, C), L, C, dy, Lp)
            dy2 = dy.view(M, Lp)

            def grads_into(kernel, bias, x2, d2, rows, cols, rp, cp, bias_tmp=None):
                # dW [rows, cols] += (x2^T d2)[:rows, :cols]; dbias += colsum(d2)[:cols].  The GEMM
                # writes the padded transpose (d2^T x2) and the un-padding transpose adds it into the
                # arena view in place (any size, any alignment).
                gwt = torch.zeros((1, cp, rp), dtype=F32, device=dev)
                ops.gemm(d2, x2, a_mn=True, b_mn=True, out=gwt[0], accumulate=True, split_k=0)
                kernel.touch_grad()
                kg = kernel.grad.view(1, rows, cols)
                ops.transpose_pad(gwt, cols, rows, kg, cols, residual=kg)
                if bias is not None:
                    if bias_tmp is None:
                        bias_tmp = torch.zeros((cp,), dtype=F32, device=dev)
                        ops.colsum(d2, out=bias_tmp, accumulate=True)
                    bias.touch_grad()
                    bg = bias.grad.view(1, cols, 1)
                    ops.transpose_pad(bias_tmp.view(1, 1, cp), 1, cols, bg, 1, residual=bg)

            grads_into(k1, b1, h, dy2, H, L, Hp, Lp)
            db0 = torch.zeros((Hp,), dtype=F32, device=dev)
            dh = ops.gemm(dy2, w1t, b_mn=True, aux_in=pre, dgelu_colsum=db0)     # [M, Hp], x gelu'(pre)
            grads_into(k0, b0, xt2, dh, L, H, Lp, Hp, bias_tmp=db0)
            if src.requires_grad:
                dxt = ops.gemm(dh, w0t, b_mn=True)                                # [M, Lp]
                dx = torch.empty((B, L, C), dtype=BF16, device=dev)
                ops.transpose_pad(dxt.view(B, C, Lp), C, L, dx, C)
                _add_grad(src, dx)
        record(bwd)
        return self


def swap_tokens_channels(x):
    """einops.rearrange(x, '... l d -> ... d l') around the token-mixing FFBlock (mlp_mixer.py:19,23)."""
    if isinstance(x, TokenMix) and not x._done:
        if x.closed:
            raise NotImplementedError("token mixing: rearranged twice after the FFBlock")
        x.closed = True
        return x
    if x.ndim != 3:
        raise NotImplementedError("swap_tokens_channels: [B, L, C] activations only")
    return TokenMix(x)


# ------------------------------------------------------------------ TNT (models/tnt.py)
def pixel_patchify(images, ph, pw, th, tw, hwcn):
    """'b (h p1) (w p2) c -> (b h w) p1 p2 c' + 'n (p1 t1) (p2 t2) c -> n (p1 p2) (c t1 t2)'
    (tnt.py:19-26) of the fp32 input images; no gradient."""
    if _meta(images):
        if hwcn:
            H, W, C, N = images.shape
        else:
            N, H, W, C = images.shape
        return Var(torch.empty((N * (H // ph) * (W // pw), (ph // th) * (pw // tw), C * th * tw),
                               dtype=BF16, device="meta"), requires_grad=False)
    return Var(ops.pixel_patchify(images.contiguous(), ph, pw, th, tw, hwcn), requires_grad=False)


def small_self_attention_core(qkv, num_heads, head_ch):
    """attention.py:39-58 for TNT's inner transformer (16 tokens, head dim 6 / 10): the FP32
    CUDA-core kernel of csrc/attn_small.cu on the fused QKV activation [B,N,3*H*d] -> [B,N,H*d]."""
    t = qkv.v
    B, N, _ = t.shape
    H, D = num_heads, head_ch
    if _meta(t):
        return Var(torch.empty((B, N, H * D), dtype=BF16, device="meta"))
    t5 = t.contiguous().view(B, N, 3, H, D)
    scale = 1.0 / math.sqrt(D)
    out = Var(ops.attn_small_fwd(t5, scale).view(B, N, H * D))

    def bwd():
        d_o = out.take_grad()
        if d_o is None:
            return
        dqkv = ops.attn_small_bwd(t5, d_o.contiguous().view(B, N, H, D), scale)
        _add_grad(qkv, dqkv.view(B, N, 3 * H * D))
    record(bwd)
    return out


def stream_to_operand(x, shape):
    """The fp32 residual stream as a bf16 GEMM operand under another shape (tnt.py:37: the inner
    transformer's output '... n d -> ... (n d)' feeds Inner2OuterBlock's Dense) while the stream itself
    goes on to the next block: the operand's gradient is added into the stream's fp32 gradient."""
    t = x.v
    if _meta(t):
        return Var(torch.empty(t.shape, dtype=BF16, device="meta").reshape(shape))
    if t.dtype != F32:
        raise TypeError("stream_to_operand: fp32 residual stream expected")
    tc = t.contiguous()
    y = torch.empty(tc.shape, dtype=BF16, device=t.device)
    ops.cast_bf16(tc, y)
    out = Var(y.view(shape))

    def bwd():
        g = out.take_grad()
        if g is None or not x.requires_grad:
            return
        C = tc.shape[-1]
        g2 = g.contiguous().view(-1, C)
        if x.g is None:
            full = torch.empty(tc.shape, dtype=F32, device=t.device)
            ops.cast2d_bf16_f32(g2, full.view(-1, C))
            x.g = full
            x.g16 = None
            return
        ops.cast2d_bf16_f32(g2, x.g.view(-1, C), accumulate=True)
        # This consumer runs AFTER the next block's LayerNorm backward in reverse order (tnt.py: the
        # pixel stream feeds Inner2OuterBlock of block i, then block i+1).  If that LayerNorm already
        # folded the producing Dense's bias gradient (column sums of the bf16 stream-gradient copy,
        # `bias_grad_from`), add this contribution's column sums too and hand the Dense a fresh bf16
        # copy of the total so that it neither recomputes nor misses anything.
        folded = getattr(x, "bias_grad_from", None) if isinstance(x, LazyLinear) else None
        if folded is not None:
            ops.colsum(g2, out=x.bias.grad.reshape(C), accumulate=True)
            g16 = torch.empty(tc.shape, dtype=BF16, device=t.device)
            ops.cast_bf16(x.g.contiguous(), g16)
            x.g16 = g16
            x.bias_grad_from = g16
        else:
            x.g16 = None
    record(bwd)
    return out


def inner_to_outer(patch_inputs, inner):
    """Inner2OuterBlock's tail (tnt.py:39-42): x = pad(inner, one zero row in front of every
    sequence) + patch_inputs.  patch_inputs: fp32 stream [B, 1 + L, C]; inner: bf16 [B*L, C] (the
    Dense output) -> new fp32 stream."""
    pv, iv = patch_inputs.v, inner.v
    B, L1, C = pv.shape
    L = L1 - 1
    if _meta(pv):
        return Var(torch.empty((B, L1, C), dtype=F32, device="meta"))
    if pv.dtype != F32:
        raise TypeError("inner_to_outer: the patch embeddings must be the fp32 residual stream")
    out_t = torch.empty((B, L1, C), dtype=F32, device=pv.device)
    ops.copy2d_f32(pv.contiguous().view(B, L1 * C), out_t.view(B, L1 * C))
    ops.cast2d_bf16_f32(iv.contiguous().view(B, L * C), out_t.view(B, L1 * C)[:, C:], accumulate=True)
    out = Var(out_t)

    def bwd():
        g = out.take_grad()
        out.g16 = None
        if g is None:
            return
        g = g.contiguous()
        if inner.requires_grad:
            tmp = torch.empty((B, L * C), dtype=F32, device=g.device)
            ops.copy2d_f32(g.view(B, L1 * C)[:, C:], tmp)
            d16 = torch.empty((B * L, C), dtype=BF16, device=g.device)
            ops.cast_bf16(tmp, d16)
            _add_grad(inner, d16.view(iv.shape))
        if patch_inputs.requires_grad:
            _add_grad(patch_inputs, g)
    record(bwd)
    return out


# ------------------------------------------------------------------ CeiT (models/ceit.py)
def patchify_act(x, ph, pw):
    """'b (h ph) (w pw) c -> b (h w) (ph pw c)' of a bf16 activation (image_to_token.py:38-43)."""
    t = x.v
    B, H, W, C = t.shape
    if _meta(t):
        return Var(torch.empty((B, (H // ph) * (W // pw), ph * pw * C), dtype=BF16, device="meta"))
    out = Var(ops.patchify_act(t.contiguous(), B, H, W, C, ph, pw))

    def bwd():
        g = out.take_grad()
        if g is not None and x.requires_grad:
            _add_grad(x, ops.patchify_act(g.contiguous(), B, H, W, C, ph, pw, inverse=True))
    record(bwd)
    return out


def stream_tokens_operand(x):
    """inputs[:, 1:] of the fp32 stream as a bf16 operand (leff.py:22: LeFF works on the patch tokens,
    the class token bypasses it); its gradient is added into rows 1.. of the stream's gradient."""
    t = x.v
    B, L1, C = t.shape
    L = L1 - 1
    if _meta(t):
        return Var(torch.empty((B, L, C), dtype=BF16, device="meta"))
    if t.dtype != F32:
        raise TypeError("stream_tokens_operand: fp32 residual stream expected")
    tc = t.contiguous()
    out = Var(ops.scale_rows_cols(tc.view(B, L1 * C)[:, C:], None, None, 1).view(B, L, C))

    def bwd():
        g = out.take_grad()
        if g is None or not x.requires_grad:
            return
        if x.g is None:
            full = torch.empty((B, L1 * C), dtype=F32, device=t.device)
            ops.fill(full, 0.0)
            x.g = full.view(B, L1, C)
        ops.cast2d_bf16_f32(g.contiguous().view(B, L * C), x.g.view(B, L1 * C)[:, C:], accumulate=True)
        x.g16 = None
    record(bwd)
    return out


def leff_residual(x, tokens):
    """y = x + concat([x[:, :1], tokens]) (leff.py:63-65 + ceit.py:42): the class token passes LeFF
    unchanged and is then added to itself, the patch tokens get the LeFF branch.  x: fp32 stream
    [B, 1 + L, C]; tokens: bf16 [B, L, C] -> new fp32 stream."""
    xv, tv = x.v, tokens.v
    B, L1, C = xv.shape
    L = L1 - 1
    if _meta(xv):
        return Var(torch.empty((B, L1, C), dtype=F32, device="meta"))
    if xv.dtype != F32:
        raise TypeError("leff_residual: the skip input must be the fp32 residual stream")
    xc = xv.contiguous().view(B, L1 * C)
    y = torch.empty((B, L1 * C), dtype=F32, device=xv.device)
    ops.copy2d_f32(xc, y)
    ops.copy2d_f32(xc[:, :C], y[:, :C], accumulate=True)
    ops.cast2d_bf16_f32(tv.contiguous().view(B, L * C), y[:, C:], accumulate=True)
    out = Var(y.view(B, L1, C))

    def bwd():
        g = out.take_grad()
        out.g16 = None
        if g is None:
            return
        g2 = g.contiguous().view(B, L1 * C)
        if tokens.requires_grad:
            tmp = torch.empty((B, L * C), dtype=F32, device=g.device)
            ops.copy2d_f32(g2[:, C:], tmp)
            d16 = torch.empty((B, L, C), dtype=BF16, device=g.device)
            ops.cast_bf16(tmp, d16)
            _add_grad(tokens, d16.view(tv.shape))
        if x.requires_grad:
            gx = torch.empty((B, L1 * C), dtype=F32, device=g.device)
            ops.copy2d_f32(g2, gx)
            ops.copy2d_f32(g2[:, :C], gx[:, :C], accumulate=True)
            _add_grad(x, gx.view(B, L1, C))
    record(bwd)
    return out


def stack_tokens(rows):
    """jnp.concatenate([expand_dims(r, 1) for r in rows], axis=1) of bf16 [B, C] Vars -> [B, n, C]
    (ceit.py:66-72: the class tokens of every encoder block)."""
    n = len(rows)
    B, C = rows[0].shape
    if _meta(rows[0].v):
        return Var(torch.empty((B, n, C), dtype=BF16, device="meta"))
    buf = torch.empty((B, n * C), dtype=BF16, device=rows[0].v.device)
    for i, r in enumerate(rows):
        ops.copy2d(r.v.reshape(B, C), buf[:, i * C:(i + 1) * C])
    out = Var(buf.view(B, n, C))

    def bwd():
        g = out.take_grad()
        if g is None:
            return
        g2 = g.contiguous().view(B, n * C)
        for i, r in enumerate(rows):
            if r.requires_grad:
                gi = torch.empty((B, C), dtype=BF16, device=g2.device)
                ops.copy2d(g2[:, i * C:(i + 1) * C], gi)
                _add_grad(r, gi.view(r.v.shape))
    record(bwd)
    return out


def last_token(x):
    """inputs[:, -1:, :] of a bf16 [B, n, C] Var (ceit.py:14: the query of the layer-wise class
    attention) -> [B, 1, C]."""
    t = x.v
    B, n, C = t.shape
    if _meta(t):
        return Var(torch.empty((B, 1, C), dtype=BF16, device="meta"))
    q = torch.empty((B, C), dtype=BF16, device=t.device)
    ops.copy2d(t.contiguous().view(B, n * C)[:, (n - 1) * C:], q)
    out = Var(q.view(B, 1, C))

    def bwd():
        g = out.take_grad()
        if g is None or not x.requires_grad:
            return
        full = torch.empty((B, n * C), dtype=BF16, device=t.device)
        ops.fill(full.view(torch.float32), 0.0)
        ops.copy2d(g.contiguous().view(B, C), full[:, (n - 1) * C:])
        _add_grad(x, full.view(B, n, C))
    record(bwd)
    return out
