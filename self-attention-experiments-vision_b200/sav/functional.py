"""Var-level ops: forward launches + recorded backward closures, with define-by-run fusion.

`linear` returns a LazyLinear that absorbs a following gelu / LayerScale / drop-path scale /
residual add into its GEMM epilogue before it is materialised, so `x + FFBlock(LN(x))` written like
the reference (vit.py:24,31) runs as fused kernels.  Abstract ('meta' device) inputs only propagate
shapes (the analogue of jax.eval_shape) -- used by the CPU shape tests; there is no CPU compute.
"""
import math

import torch

from . import ops
from .tape import Var, record

BF16 = torch.bfloat16


def _meta(t):
    return t.is_meta


class Weight:
    """2-D bf16 weight view [K,N] (unit inner stride) + its fp32 gradient view + owning Params."""
    __slots__ = ("w", "gw", "params")

    def __init__(self, w, gw, params):
        self.w, self.gw, self.params = w, gw, params

    def accumulate_flag(self):
        flags = [p.grad_accumulate_flag() for p in self.params]
        return all(flags)


class LazyLinear(Var):
    __slots__ = ("x", "weight", "bias", "act", "col_scale", "row_scale", "rows_per_scale",
                 "residual", "out_shape", "_done")

    def __init__(self, x, weight, bias, out_shape):
        super().__init__(None)
        self.x, self.weight, self.bias = x, weight, bias
        self.act = 0
        self.col_scale = self.row_scale = self.residual = None
        self.rows_per_scale = 1
        self.out_shape = tuple(out_shape)
        self._done = False

    @property
    def shape(self):
        return self.out_shape

    @property
    def ndim(self):
        return len(self.out_shape)

    @property
    def v(self):
        if not self._done:
            self._materialize()
        return self._v

    def _materialize(self):
        self._done = True
        x, W = self.x, self.weight
        xin = x.v
        K, N = W.w.shape
        if _meta(xin):
            self._v = torch.empty(self.out_shape, dtype=BF16, device="meta")
            return
        x2 = xin.reshape(-1, K) if xin.is_contiguous() else xin  # strided 2-D views pass through
        M = x2.shape[0]
        out = torch.empty((M, N), dtype=BF16, device=x2.device)
        pre = torch.empty((M, N), dtype=BF16, device=x2.device) if self.act else None
        res = self.residual.v.reshape(M, N) if self.residual is not None else None
        ops.gemm(x2, W.w, b_mn=True, out=out,
                 bias=self.bias.master if self.bias is not None else None, act=self.act,
                 aux_out=pre,
                 col_scale=self.col_scale.master if self.col_scale is not None else None,
                 row_scale=self.row_scale, rows_per_scale=self.rows_per_scale, residual=res)
        self._v = out.view(self.out_shape)
        self.pre_act = pre
        record(self._backward)

    def _backward(self):
        dy = self.take_grad()
        if dy is None:
            return
        x, W = self.x, self.weight
        K, N = W.w.shape
        dy2 = dy.reshape(-1, N)
        if self.residual is not None and self.residual.requires_grad:
            _add_grad(self.residual, dy.view(self.residual.v.shape))
        if self.col_scale is not None or self.row_scale is not None:
            raise NotImplementedError("backward through LayerScale / drop-path epilogue")
        if self.act and not self.g_is_pre:
            raise NotImplementedError("activated linear output must be consumed by a linear op")
        x2 = x.v.reshape(-1, K) if x.v.is_contiguous() else x.v
        # dW[K,N] = X^T dY : both operands token-major (MN-major), split-K over tokens, fp32 out
        ops.gemm(x2, dy2, a_mn=True, b_mn=True, out=W.gw, accumulate=W.accumulate_flag(), split_k=0)
        if self.bias is not None:
            b = self.bias
            ops.colsum(dy2, out=b.grad, accumulate=b.grad_accumulate_flag())
        if x.requires_grad:
            # dX[M,K] = dY W^T (+ gelu'(pre) when x is a fused-GELU output) (+ x.g fan-in)
            prev = x.take_grad()
            dx = ops.gemm(dy2, W.w, b_mn=False, aux_in=x.pre_act,
                          residual=prev.reshape(-1, K) if prev is not None else None)
            x.g = dx.view(x.v.shape) if x.v.is_contiguous() else dx
            x.g_is_pre = x.pre_act is not None


def _add_grad(var, g):
    if var.g is None:
        var.g = g
    else:
        var.g = ops.add(var.g.contiguous(), g.contiguous()).view(var.g.shape)


def linear(x, weight, bias=None, out_features_shape=None):
    lead = tuple(x.shape[:-1])
    N = weight.w.shape[1]
    return LazyLinear(x, weight, bias, lead + (N,))


def gelu(x):
    """jax.nn.gelu (tanh form) -- only as the epilogue of the producing linear (ff.py:24-25)."""
    if isinstance(x, LazyLinear) and not x._done and x.act == 0 and x.residual is None:
        x.act = 1
        return x
    raise NotImplementedError("standalone gelu: apply it directly to a Dense output")


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
    """Residual add (vit.py:24,31): folded into a pending linear when possible."""
    for lazy, other in ((a, b), (b, a)):
        if isinstance(lazy, LazyLinear) and not lazy._done and lazy.residual is None and \
                isinstance(other, Var) and tuple(other.shape) == tuple(lazy.shape):
            lazy.residual = other
            lazy.v  # materialise now so program order is preserved
            return lazy
    av, bv = a.v, b.v
    if _meta(av):
        return Var(torch.empty(av.shape, dtype=BF16, device="meta"))
    out = Var(ops.add(av.contiguous(), bv.contiguous()).view(av.shape))

    def bwd():
        g = out.take_grad()
        if g is not None:
            _add_grad(a, g)
            _add_grad(b, g)
    record(bwd)
    return out


def layernorm(x, scale, bias, eps=1e-6):
    xv = x.v
    if _meta(xv):
        return Var(torch.empty(xv.shape, dtype=BF16, device="meta"))
    xc = xv.contiguous()
    y, mean, rstd = ops.layernorm_fwd(xc, scale.master, bias.master, eps)
    out = Var(y)

    def bwd():
        dy = out.take_grad()
        if dy is None:
            return
        acc = scale.grad_accumulate_flag()
        bias.grad_accumulate_flag()
        prev = x.take_grad() if x.requires_grad else None
        dx = ops.layernorm_bwd(dy.contiguous(), xc, scale.master, mean, rstd, scale.grad, bias.grad,
                               dx_add=prev.contiguous() if prev is not None else None,
                               accumulate=acc)
        if x.requires_grad:
            x.g = dx
    record(bwd)
    return out


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
    out = Var(o.view(B, N, H * D))

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
    out = Var(o.view(B, Nq, H * D))

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
    """[cls; tok] + pos (vit.py:81-85, position_embed.py:54-56); cls may be None (cait.py:69)."""
    t = tok.v
    B, n, C = t.shape
    has_cls = cls is not None
    if _meta(t):
        return Var(torch.empty((B, n + int(has_cls), C), dtype=BF16, device="meta"))
    x = ops.embed_fwd(t.contiguous(), cls.master if has_cls else None, pos.master, B, n, C)
    out = Var(x)

    def bwd():
        dx = out.take_grad()
        if dx is None:
            return
        acc = pos.grad_accumulate_flag()
        if has_cls:
            cls.grad_accumulate_flag()
        dtok = ops.embed_bwd(dx.contiguous(), cls.grad if has_cls else None, pos.grad, B, n, C,
                             accumulate=acc)
        if tok.requires_grad:
            _add_grad(tok, dtok.view(B, n, C))
    record(bwd)
    return out


def getitem(x, idx):
    """Only the class-token select x[:, 0] (vit.py:95) is needed on the hot path."""
    if not (isinstance(idx, tuple) and len(idx) == 2 and idx[0] == slice(None) and idx[1] == 0):
        raise NotImplementedError("Var indexing supports x[:, 0] only")
    t = x.v
    B, N, C = t.shape
    if _meta(t):
        return Var(torch.empty((B, C), dtype=BF16, device="meta"))
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


def softmax_cross_entropy(logits, labels, label_smoothing=0.0, world_size=1):
    """train.py:83-91 (plain mean over the global batch)."""
    lv = logits.v
    B = lv.shape[0]
    dl, stats = ops.softmax_xent(lv.contiguous(), labels, label_smoothing, 1.0 / (B * world_size))
    return LossVar(logits, dl, stats, B)
