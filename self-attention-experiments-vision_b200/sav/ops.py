"""Functional forward/backward ops over torch device buffers; each is one C-ABI call.

torch is used only as the device allocator / stream owner.  All arithmetic runs in
libsav_b200.so; nothing here falls back to torch ops.
"""
import ctypes as C
import math
import os

import torch

from . import _abi
from ._abi import AttnBwdArgs, AttnFwdArgs, GemmArgs, check, lib

BF16 = torch.bfloat16
F32 = torch.float32


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _req(t, dtype, name):
    if t.dtype != dtype or not t.is_cuda:
        raise _abi.SavError(f"{name}: expected CUDA {dtype}, got {t.dtype} on {t.device}")


_ws_cache = {}


def _workspace(nbytes, device, tag="gemm"):
    """Grow-only scratch per (device, tag, stream); reused across calls on the same stream."""
    key = (device, tag, torch.cuda.current_stream().cuda_stream)
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=device)
        _ws_cache[key] = buf
    return buf


def gemm(A, B, *, a_mn=False, b_mn=False, out=None, out_fp32=False, bias=None, act=0,
         aux_out=None, aux_in=None, col_scale=None, row_scale=None, rows_per_scale=1,
         residual=None, accumulate=False, split_k=1, dgelu_colsum=None, ln=None):
    """D[M,N] = epi(sum_k A[m,k] B[n,k]).  A: [M,K] (or [K,M] if a_mn); B: [N,K] (or [K,N] if b_mn).

    2-D bf16 tensors with unit inner stride; row strides are taken from the tensors, so column
    slices of a wider buffer are fine.
    """
    _req(A, BF16, "A"); _req(B, BF16, "B")
    assert A.dim() == 2 and B.dim() == 2 and A.stride(1) == 1 and B.stride(1) == 1
    M, K = (A.shape[1], A.shape[0]) if a_mn else (A.shape[0], A.shape[1])
    N, Kb = (B.shape[1], B.shape[0]) if b_mn else (B.shape[0], B.shape[1])
    if K != Kb:
        raise _abi.SavError(f"gemm: contraction mismatch {K} vs {Kb}")
    if out is None:
        out = torch.empty((M, N), dtype=F32 if out_fp32 else BF16, device=A.device)
    assert out.shape == (M, N) and out.stride(1) == 1
    a = GemmArgs()
    a.A, a.B, a.D = A.data_ptr(), B.data_ptr(), out.data_ptr()
    a.M, a.N, a.K = M, N, K
    a.lda, a.ldb, a.ldd = A.stride(0), B.stride(0), out.stride(0)
    a.a_mn, a.b_mn = int(a_mn), int(b_mn)
    a.d_fp32 = int(out.dtype == F32)
    a.split_k = split_k
    a.bias = bias.data_ptr() if bias is not None else None
    a.act = act
    a.accumulate = int(accumulate)
    for name, t in (("aux_out", aux_out), ("aux_in", aux_in)):
        if t is not None:
            assert t.shape == (M, N) and t.stride(0) == out.stride(0) and t.dtype == BF16
            setattr(a, name, t.data_ptr())
    if col_scale is not None:
        _req(col_scale, F32, "col_scale"); a.col_scale = col_scale.data_ptr()
    if row_scale is not None:
        _req(row_scale, F32, "row_scale"); a.row_scale = row_scale.data_ptr()
    a.rows_per_scale = rows_per_scale
    if residual is not None:
        _req(residual, F32, "residual")
        assert residual.shape == (M, N) and residual.stride(1) == 1
        a.residual, a.ldr = residual.data_ptr(), residual.stride(0)
    ln_res = None
    if ln is not None:
        # fused LayerNorm of the fp32 stream this GEMM produces: ln = (gamma, beta, eps)
        gamma, beta, eps = ln
        y = torch.empty((M, N), dtype=BF16, device=A.device)
        mean = torch.empty(M, dtype=F32, device=A.device)
        rstd = torch.empty(M, dtype=F32, device=A.device)
        a.ln_out, a.ldy = y.data_ptr(), y.stride(0)
        a.ln_gamma, a.ln_beta = gamma.data_ptr(), beta.data_ptr()
        a.ln_mean, a.ln_rstd, a.ln_eps = mean.data_ptr(), rstd.data_ptr(), float(eps)
        ln_res = (y, mean, rstd)
    if dgelu_colsum is not None:
        _req(dgelu_colsum, F32, "dgelu_colsum")
        assert aux_in is not None and dgelu_colsum.numel() == N and dgelu_colsum.is_contiguous()
        a.dgelu_colsum = dgelu_colsum.data_ptr()
    check(lib().sav_gemm(C.byref(a), _stream()))
    if ln_res is not None:
        return out, ln_res
    return out


def gemm_ln_tiles(N):
    """Non-zero when a LayerNorm over N columns can ride in sav_gemm's fp32-stream epilogue."""
    return int(lib().sav_gemm_ln_tiles(N))


_COLSUM_WS = {}


def colsum(X, out=None, accumulate=False):
    """out[n] (+)= sum_m X[m,n] through the two-pass entry point (row-split partials in a cached
    per-device workspace, no same-address atomics)."""
    _req(X, BF16, "X")
    M, N = X.shape
    if out is None:
        out = torch.empty(N, dtype=F32, device=X.device)
    need = int(lib().sav_colsum_workspace_bytes(M, N))
    ws = _COLSUM_WS.get(X.device)
    if ws is None or ws.numel() < need:
        # allocated outside CUDA-graph capture by the eager warm-up step; the stream orders its reuse
        ws = _COLSUM_WS[X.device] = torch.empty(max(need, 16 << 20), dtype=torch.uint8, device=X.device)
    check(lib().sav_colsum_bf16_ws(_p(X), M, N, X.stride(0), _p(out), int(accumulate), _p(ws),
                                   ws.numel(), _stream()))
    return out


def layernorm_fwd(x, scale, bias, eps=1e-6):
    """x: fp32 [rows, C] view of the residual stream (unit inner stride, any row pitch)."""
    _req(x, F32, "x")
    assert x.dim() == 2 and x.stride(1) == 1
    rows, C_ = x.shape
    y = torch.empty((rows, C_), dtype=BF16, device=x.device)
    mean = torch.empty(rows, dtype=F32, device=x.device)
    rstd = torch.empty(rows, dtype=F32, device=x.device)
    check(lib().sav_layernorm_fwd(_p(x), x.stride(0), _p(scale), _p(bias), _p(y), y.stride(0),
                                  _p(mean), _p(rstd), rows, C_, eps, _stream()))
    return y, mean, rstd


def layernorm_bwd(dy, x, scale, mean, rstd, dscale, dbias, dx_add=None, dx=None, dx16=None,
                  dx16_colsum=None):
    """dy bf16 [rows,C]; x fp32 [rows,C] view; returns (dx fp32, dx16 bf16) with x's row pitch.
    dscale / dbias are accumulated."""
    _req(dy, BF16, "dy"); _req(x, F32, "x")
    assert dy.dim() == 2 and x.dim() == 2 and dy.stride(1) == 1 and x.stride(1) == 1
    rows, C_ = x.shape
    if dx is None:
        assert x.is_contiguous()
        dx = torch.empty_like(x)
        dx16 = torch.empty((rows, C_), dtype=BF16, device=x.device)
    assert dx.stride(0) == x.stride(0) and (dx16 is None or dx16.stride(0) == x.stride(0))
    assert dx_add is None or (dx_add.dtype == F32 and dx_add.stride(0) == x.stride(0))
    if dx16_colsum is not None:
        # accumulates the column sums of dx16 (the bias gradient of the Linear that wrote x)
        _req(dx16_colsum, F32, "dx16_colsum")
        assert dx16_colsum.numel() == C_ and dx16 is not None
    check(lib().sav_layernorm_bwd_colsum(_p(dy), dy.stride(0), _p(x), x.stride(0), _p(scale), _p(mean),
                                         _p(rstd), _p(dx_add), _p(dx), _p(dx16), _p(dscale), _p(dbias),
                                         _p(dx16_colsum), rows, C_, _stream()))
    return dx, dx16


def patchify(images, ph, pw, hwcn=True):
    """images fp32 [H,W,C,N] (hwcn) or [N,H,W,C] -> bf16 [N*gh*gw, ph*pw*C]."""
    _req(images, F32, "images")
    assert images.is_contiguous()
    if hwcn:
        H, W, Cc, N = images.shape
    else:
        N, H, W, Cc = images.shape
    out = torch.empty((N * (H // ph) * (W // pw), ph * pw * Cc), dtype=BF16, device=images.device)
    check(lib().sav_patchify(_p(images), _p(out), N, H, W, Cc, ph, pw, int(hwcn), _stream()))
    return out


def embed_fwd(tok, cls, pos, B, n_tok, C_):
    has_cls = cls is not None
    x = torch.empty((B, n_tok + int(has_cls), C_), dtype=F32, device=tok.device)
    check(lib().sav_embed_fwd(_p(tok), _p(cls), _p(pos), _p(x), B, n_tok, C_, int(has_cls), _stream()))
    return x


def embed_bwd(dx, dcls, dpos, B, n_tok, C_, accumulate=False):
    has_cls = dcls is not None
    dtok = torch.empty((B * n_tok, C_), dtype=BF16, device=dx.device)
    check(lib().sav_embed_bwd(_p(dx), _p(dtok), _p(dcls), _p(dpos), B, n_tok, C_, int(has_cls),
                              int(accumulate), _stream()))
    return dtok


def _bnhd_strides(t):
    """t: [B,N,H,D] view with D contiguous and H stride == D."""
    assert t.dim() == 4 and t.stride(3) == 1 and t.stride(2) == t.shape[3]
    return t.stride(0), t.stride(1)


_ATTN_SAVE_LO = os.environ.get("SAV_ATTN_SAVE_LO", "1") == "1"


def attn_fwd(q, k, v, scale=None, save_lo=None):
    """q [B,Nq,H,D], k/v [B,Nk,H,D] bf16 (strided views ok) -> o [2,B,Nq,H,D] (o[0] = output,
    o[1] = its low half for backward when save_lo), lse [B,H,Nq]."""
    if save_lo is None:
        save_lo = _ATTN_SAVE_LO
    B, Nq, H, D = q.shape
    Nk = k.shape[1]
    o2 = torch.empty((2 if save_lo else 1, B, Nq, H, D), dtype=BF16, device=q.device)
    o = o2[0]
    lse = torch.empty((B, H, Nq), dtype=F32, device=q.device)
    a = AttnFwdArgs()
    a.q, a.k, a.v, a.o, a.lse = q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), lse.data_ptr()
    a.o_lo = o2[1].data_ptr() if save_lo else None
    a.B, a.H, a.Nq, a.Nk, a.D = B, H, Nq, Nk, D
    a.q_batch_stride, a.q_row_stride = _bnhd_strides(q)
    a.k_batch_stride, a.k_row_stride = _bnhd_strides(k)
    a.v_batch_stride, a.v_row_stride = _bnhd_strides(v)
    a.o_batch_stride, a.o_row_stride = _bnhd_strides(o)
    a.scale = scale if scale is not None else 1.0 / math.sqrt(D)
    check(lib().sav_attn_fwd(C.byref(a), _stream()))
    return o2, lse


def attn_bwd(q, k, v, o2, d_o, lse, dq, dk, dv, scale=None):
    """Writes dq/dk/dv (views with the same strides as q/k/v).  o2: what attn_fwd returned."""
    B, Nq, H, D = q.shape
    Nk = k.shape[1]
    o = o2[0]
    assert d_o.stride() == o.stride()
    assert dq.stride() == q.stride() and dk.stride() == k.stride() and dv.stride() == v.stride()
    delta = torch.empty((B, H, Nq), dtype=F32, device=q.device)
    # sequences beyond 256 queries run as query groups: fp32 accumulators of dK and dV (zeroed by
    # the callee), large enough for the development build's older kernel (dQ accumulator) as well
    dq_acc = torch.empty((max(2 * B * Nk, B * Nq), H, D), dtype=F32, device=q.device) if Nq > 256 else None
    a = AttnBwdArgs()
    a.q, a.k, a.v, a.o, a.d_o, a.lse = (q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(),
                                        d_o.data_ptr(), lse.data_ptr())
    a.o_lo = o2[1].data_ptr() if o2.shape[0] > 1 else None
    a.dq, a.dk, a.dv = dq.data_ptr(), dk.data_ptr(), dv.data_ptr()
    a.delta, a.dq_accum = delta.data_ptr(), (dq_acc.data_ptr() if dq_acc is not None else None)
    a.B, a.H, a.Nq, a.Nk, a.D = B, H, Nq, Nk, D
    a.q_batch_stride, a.q_row_stride = _bnhd_strides(q)
    a.k_batch_stride, a.k_row_stride = _bnhd_strides(k)
    a.v_batch_stride, a.v_row_stride = _bnhd_strides(v)
    a.o_batch_stride, a.o_row_stride = _bnhd_strides(o)
    a.scale = scale if scale is not None else 1.0 / math.sqrt(D)
    check(lib().sav_attn_bwd(C.byref(a), _stream()))


def softmax_xent(logits, labels, label_smoothing=0.0, grad_scale=1.0, mix_labels=None, ratio=None):
    """logits bf16 [B,classes], labels int32 [B] (+ optional mixup partner labels int32 [B] and
    ratio fp32 [B], train.py:84-87) -> (dlogits bf16, stats fp32 [loss_sum, top1])."""
    _req(logits, BF16, "logits")
    B, classes = logits.shape
    if labels.dtype != torch.int32 or (mix_labels is not None and mix_labels.dtype != torch.int32):
        raise _abi.SavError("softmax_xent: labels must be int32")
    if (mix_labels is None) != (ratio is None):
        raise _abi.SavError("softmax_xent: mix_labels and ratio go together")
    if ratio is not None:
        _req(ratio, F32, "ratio")
    dl = torch.empty_like(logits)
    stats = torch.empty(2, dtype=F32, device=logits.device)
    check(lib().sav_softmax_xent_mix(_p(logits), logits.stride(0), _p(labels), _p(mix_labels), _p(ratio),
                                     B, classes, label_smoothing, grad_scale, _p(dl), _p(stats), _stream()))
    return dl, stats


def schedule_step(step_dev, hyper, mode, init_value, peak_value, warmup_steps, decay_steps, end_value,
                  b1, b2):
    """Advance the device step counter (int64 [1]) and write {lr, 1-b1^t, 1-b2^t} into hyper."""
    assert step_dev.dtype == torch.int64 and hyper.dtype == F32 and hyper.numel() >= 3
    check(lib().sav_schedule_step(_p(step_dev), _p(hyper), int(mode), float(init_value), float(peak_value),
                                  int(warmup_steps), int(decay_steps), float(end_value), float(b1),
                                  float(b2), _stream()))


def drop_path_factors(seed, step_dev, step_host, call_index, B, keep, out=None, device=None):
    """floor(keep + U)/keep per sample from the device Philox stream (stochastic_depth.py:16-27)."""
    if out is None:
        out = torch.empty(B, dtype=F32, device=device)
    check(lib().sav_drop_path_factors(int(seed) & 0xFFFFFFFFFFFFFFFF, _p(step_dev), int(step_host),
                                      int(call_index), B, float(keep), _p(out), _stream()))
    return out


def cast_bf16(src, dst):
    check(lib().sav_cast_f32_bf16(_p(src), _p(dst), src.numel(), _stream()))


def sumsq(g, out):
    check(lib().sav_sumsq(_p(g), g.numel(), _p(out), _stream()))


def dp_allreduce_sum(multicast_ptr, peer_ptrs_dev, signal_pad_ptrs_dev, signal_pad_bytes, offset, n,
                     rank, world, blocks=0):
    """In-place SUM over the ranks of floats [offset, offset + n) of a symmetric allocation
    (csrc/dp.cu); the pointer tables are device addresses (ints)."""
    check(lib().sav_dp_allreduce_sum_f32(int(multicast_ptr), int(peer_ptrs_dev), int(signal_pad_ptrs_dev),
                                         int(signal_pad_bytes), int(offset), int(n), int(rank),
                                         int(world), int(blocks), _stream()))


def fill(t, v):
    check(lib().sav_fill_f32(_p(t), float(v), t.numel(), _stream()))


def adam_step(p, g, m, v, p_bf16, sumsq_buf, max_norm, lr, b1, b2, eps, wd, t, hyper=None):
    check(lib().sav_adam_step(_p(p), _p(g), _p(m), _p(v), _p(p_bf16), p.numel(), _p(sumsq_buf),
                              max_norm, lr, b1, b2, eps, wd, t, _p(hyper), _stream()))


def add(a, b):
    y = torch.empty_like(a)
    check(lib().sav_add_bf16(_p(a), _p(b), _p(y), a.numel(), _stream()))
    return y


def copy2d(src, dst):
    """dst[r, :] = src[r, :] for 2-D bf16 views with unit inner stride."""
    rows, cols = src.shape
    assert dst.shape == src.shape and src.stride(1) == 1 and dst.stride(1) == 1
    check(lib().sav_copy2d_bf16(_p(src), src.stride(0), _p(dst), dst.stride(0), rows, cols, _stream()))


# ------------------------------------------------------------------ materialised-logits attention
def bgemm(A, B, D, M, N, K, a_mn=False, b_mn=False):
    """Batched GEMM over 4-D views [outer, inner, rows, cols] (cols contiguous; any other strides).
    A: rows = M (or K if a_mn), B: rows = N (or K if b_mn), D: [outer, inner, M, N] bf16 or fp32."""
    from ._abi import BGemmArgs
    for t in (A, B, D):
        assert t.dim() == 4 and t.stride(3) == 1
    a = BGemmArgs()
    a.A, a.B, a.D = A.data_ptr(), B.data_ptr(), D.data_ptr()
    a.M, a.N, a.K = M, N, K
    a.batch_outer, a.batch_inner = A.shape[0], A.shape[1]
    a.a_stride_outer, a.a_stride_inner, a.lda = A.stride(0), A.stride(1), A.stride(2)
    a.b_stride_outer, a.b_stride_inner, a.ldb = B.stride(0), B.stride(1), B.stride(2)
    a.d_stride_outer, a.d_stride_inner, a.ldd = D.stride(0), D.stride(1), D.stride(2)
    a.a_mn, a.b_mn, a.d_fp32 = int(a_mn), int(b_mn), int(D.dtype == F32)
    check(lib().sav_bgemm(C.byref(a), _stream()))
    return D


def _thsoftmax_args(S, Nk, scale, mix_pre, mix_post, rel, rel_hw):
    from ._abi import ThSoftmaxArgs
    B, H, Nq, ld = S.shape
    a = ThSoftmaxArgs()
    a.S = S.data_ptr()
    a.B, a.H, a.Nq, a.Nk, a.ld = B, H, Nq, Nk, ld
    a.scale = scale
    if mix_pre is not None:
        a.mix_pre, a.mix_post = mix_pre.data_ptr(), mix_post.data_ptr()
    if rel is not None:
        assert rel.is_contiguous() and rel.dtype == F32 and rel.dim() == 4
        a.rel, a.ldr = rel.data_ptr(), rel.shape[3]
        a.rel_h, a.rel_w = rel_hw
    return a


def attn_mat_fwd(q, k, v, scale, mix_pre=None, mix_post=None, rel=None, rel_hw=None):
    """q [B,Nq,H,D], k/v [B,Nk,H,D] bf16 views -> (o [B,Nq,H,D], S fp32 [B,H,Nq,ld], P bf16).
    Any D % 8 == 0; optional talking-heads transforms [H,H] fp32 and relative logits
    rel fp32 [B,Nq,H,ldr] on an rel_hw = (Hh, W) key grid."""
    B, Nq, H, D = q.shape
    Nk = k.shape[1]
    ld = (Nk + 7) // 8 * 8
    S = torch.empty((B, H, Nq, ld), dtype=F32, device=q.device)
    P = torch.empty((B, H, Nq, ld), dtype=BF16, device=q.device)
    qv, kv, vv = q.permute(0, 2, 1, 3), k.permute(0, 2, 1, 3), v.permute(0, 2, 1, 3)
    bgemm(qv, kv, S[..., :Nk], Nq, Nk, D)
    a = _thsoftmax_args(S, Nk, scale, mix_pre, mix_post, rel, rel_hw)
    a.P = P.data_ptr()
    check(lib().sav_thsoftmax_fwd(C.byref(a), _stream()))
    o = torch.empty((B, Nq, H, D), dtype=BF16, device=q.device)
    bgemm(P[..., :Nk], vv, o.permute(0, 2, 1, 3), Nq, D, Nk, b_mn=True)
    return o, S, P


def attn_mat_bwd(q, k, v, S, P, d_o, dq, dk, dv, scale, mix_pre=None, mix_post=None,
                 dmix_pre=None, dmix_post=None, rel=None, rel_hw=None):
    """Reverse mode of attn_mat_fwd; writes dq/dk/dv (views like q/k/v), accumulates dmix_*,
    returns drel (fp32, like rel) or None."""
    B, Nq, H, D = q.shape
    Nk = k.shape[1]
    ld = S.shape[3]
    qv, kv, vv = q.permute(0, 2, 1, 3), k.permute(0, 2, 1, 3), v.permute(0, 2, 1, 3)
    dov = d_o.permute(0, 2, 1, 3)
    Pv = P[..., :Nk]
    bgemm(Pv, dov, dv.permute(0, 2, 1, 3), Nk, D, Nq, a_mn=True, b_mn=True)       # dV = P^T dO
    dP = torch.empty((B, H, Nq, ld), dtype=F32, device=q.device)
    bgemm(dov, vv, dP[..., :Nk], Nq, Nk, D)                                        # dP = dO V^T
    dS = torch.empty((B, H, Nq, ld), dtype=BF16, device=q.device)
    a = _thsoftmax_args(S, Nk, scale, mix_pre, mix_post, rel, rel_hw)
    a.dP, a.dS = dP.data_ptr(), dS.data_ptr()
    drel = None
    if mix_pre is not None:
        a.dmix_pre, a.dmix_post = dmix_pre.data_ptr(), dmix_post.data_ptr()
    if rel is not None:
        drel = torch.empty_like(rel)
        a.drel = drel.data_ptr()
    check(lib().sav_thsoftmax_bwd(C.byref(a), _stream()))
    dSv = dS[..., :Nk]
    bgemm(dSv, kv, dq.permute(0, 2, 1, 3), Nq, D, Nk, b_mn=True)                  # dQ = dS K
    bgemm(dSv, qv, dk.permute(0, 2, 1, 3), Nk, D, Nq, a_mn=True, b_mn=True)       # dK = dS^T Q
    return drel


def scale_rows_cols(g, col_scale, row_scale, rows_per_scale, out=None):
    """bf16(g * col_scale[n] * row_scale[m // rows_per_scale]) for fp32 g [rows, cols]."""
    _req(g, F32, "g")
    rows, cols = g.shape
    if out is None:
        out = torch.empty((rows, cols), dtype=BF16, device=g.device)
    check(lib().sav_scale_rows_cols(_p(g), g.stride(0), _p(col_scale), _p(row_scale),
                                    rows_per_scale, _p(out), out.stride(0), rows, cols, _stream()))
    return out


def layerscale_grad(W, dW, bias, dbias, col_scale, dcol_scale):
    K, N = W.shape
    check(lib().sav_layerscale_grad(_p(W), W.stride(0), _p(dW), dW.stride(0), K, N, _p(bias),
                                    _p(dbias), _p(col_scale), _p(dcol_scale), _stream()))


def copy2d_f32(src, dst, accumulate=False):
    rows, cols = src.shape
    assert dst.shape == src.shape and src.stride(1) == 1 and dst.stride(1) == 1
    check(lib().sav_copy2d_f32(_p(src), src.stride(0), _p(dst), dst.stride(0), rows, cols,
                               int(accumulate), _stream()))


# ------------------------------------------------------------------ CvT convolutional projection
def dwconv_out_hw(H, W, stride):
    ho, wo = C.c_int64(0), C.c_int64(0)
    check(lib().sav_dwconv3x3_out_hw(H, W, stride, C.byref(ho), C.byref(wo)))
    return int(ho.value), int(wo.value)


def dwconv3x3_fwd(x, w, stride):
    """x fp32 [B,H,W,C], w fp32 [3,3,C] -> (y bf16 [B,Ho,Wo,C], sum [C], sumsq [C])."""
    _req(x, F32, "x")
    B, H, W, Cc = x.shape
    Ho, Wo = dwconv_out_hw(H, W, stride)
    y = torch.empty((B, Ho, Wo, Cc), dtype=BF16, device=x.device)
    stats = torch.empty((2, Cc), dtype=F32, device=x.device)
    check(lib().sav_dwconv3x3_fwd(_p(x), _p(w), _p(y), _p(stats[0]), _p(stats[1]), B, H, W, Cc, stride,
                                  _stream()))
    return y, stats


def dwconv3x3_bwd(x, w, dy, dx, dw, stride, accumulate_dx):
    B, H, W, Cc = x.shape
    check(lib().sav_dwconv3x3_bwd(_p(x), _p(w), _p(dy), _p(dx), _p(dw), B, H, W, Cc, stride,
                                  int(accumulate_dx), _stream()))


def batchnorm_fwd(y, stats, gamma, beta, eps):
    """y bf16 [..., C] + (sum, sumsq) -> (z bf16, mean, rstd, var)."""
    Cc = y.shape[-1]
    rows = y.numel() // Cc
    z = torch.empty_like(y)
    mrv = torch.empty((3, Cc), dtype=F32, device=y.device)
    check(lib().sav_batchnorm_fwd(_p(y), _p(stats[0]), _p(stats[1]), _p(gamma), _p(beta), eps, _p(z),
                                  _p(mrv[0]), _p(mrv[1]), _p(mrv[2]), rows, Cc, _stream()))
    return z, mrv


def batchnorm_bwd(dz, y, mrv, gamma, dgamma, dbeta):
    Cc = y.shape[-1]
    rows = y.numel() // Cc
    dy = torch.empty_like(y)
    scratch = torch.empty(2 * Cc, dtype=F32, device=y.device)
    check(lib().sav_batchnorm_bwd(_p(dz), _p(y), _p(mrv[0]), _p(mrv[1]), _p(gamma), _p(dgamma),
                                  _p(dbeta), _p(scratch), _p(dy), rows, Cc, _stream()))
    return dy


# ------------------------------------------------------------------ rotary helpers
def rotary(x, seq_axis, inverse=False):
    """Rotate the (2i, 2i+1) pairs of the last axis of a contiguous bf16 / fp32 tensor by
    t * inv_freq[i], t = index along `seq_axis` (position_embed.py:8-20)."""
    assert x.is_contiguous() and x.dtype in (BF16, F32)
    seq_axis %= x.dim()
    assert seq_axis < x.dim() - 1
    outer = 1
    for s_ in x.shape[:seq_axis]:
        outer *= s_
    inner = 1
    for s_ in x.shape[seq_axis + 1:-1]:
        inner *= s_
    y = torch.empty_like(x)
    check(lib().sav_rotary(_p(x), _p(y), outer, x.shape[seq_axis], inner, x.shape[-1],
                           int(x.dtype == F32), int(inverse), _stream()))
    return y


def rotary_table(n, dim, device):
    emb = torch.empty((n, dim), dtype=F32, device=device)
    check(lib().sav_rotary_table(_p(emb), n, dim, _stream()))
    return emb


# ------------------------------------------------------------------ dense convolutions (im2col + GEMM)
def conv_same_geom(n, k, stride):
    """flax padding='SAME': (out, pad_lo)."""
    out, lo = C.c_int64(0), C.c_int32(0)
    check(lib().sav_conv_same_geom(n, k, stride, C.byref(out), C.byref(lo)))
    return int(out.value), int(lo.value)


def conv2d_geom(B, H, W, Cin, k, stride, padding="SAME", x_strides=None):
    """Geometry of a k x k / stride conv over [B,H,W,Cin]; padding 'SAME' or an explicit low-side
    amount (botnet.py:294: padding=[(3,3),(3,3)])."""
    from ._abi import Conv2dGeom
    g = Conv2dGeom()
    g.B, g.H, g.W, g.C = B, H, W, Cin
    g.kh = g.kw = k
    g.sh = g.sw = stride
    if padding == "SAME":
        g.Ho, g.pad_top = conv_same_geom(H, k, stride)
        g.Wo, g.pad_left = conv_same_geom(W, k, stride)
    else:
        p = int(padding)
        g.pad_top = g.pad_left = p
        g.Ho = (H + 2 * p - k) // stride + 1
        g.Wo = (W + 2 * p - k) // stride + 1
    if x_strides is None:
        x_strides = (H * W * Cin, W * Cin, Cin, 1)
    g.xs_b, g.xs_h, g.xs_w, g.xs_c = x_strides
    K = k * k * Cin
    g.ld = (K + 7) // 8 * 8
    return g


def im2col(x, g):
    """x fp32 / bf16 (addressed through g's strides) -> bf16 [B*Ho*Wo, ld]."""
    assert x.dtype in (F32, BF16)
    cols = torch.empty((g.B * g.Ho * g.Wo, g.ld), dtype=BF16, device=x.device)
    check(lib().sav_im2col(_p(x), int(x.dtype == F32), _p(cols), C.byref(g), _stream()))
    return cols


def col2im(dcols, g, dx, accumulate=False):
    """dx [B,H,W,C] contiguous fp32 / bf16 (+)= fold(dcols bf16 [B*Ho*Wo, ld])."""
    from ._abi import Conv2dGeom
    assert dcols.dtype == BF16 and dcols.stride(0) == g.ld and dx.is_contiguous()
    gc = Conv2dGeom.from_buffer_copy(g)
    gc.xs_b, gc.xs_h, gc.xs_w, gc.xs_c = g.H * g.W * g.C, g.W * g.C, g.C, 1
    check(lib().sav_col2im(_p(dcols), _p(dx), int(dx.dtype == F32), int(accumulate), C.byref(gc), _stream()))
    return dx


def bn_stats(y):
    """Per-channel (sum, sumsq) fp32 [2, C] of a bf16 [..., C] tensor."""
    _req(y, BF16, "y")
    Cc = y.shape[-1]
    stats = torch.empty((2, Cc), dtype=F32, device=y.device)
    check(lib().sav_bn_stats(_p(y), y.numel() // Cc, Cc, _p(stats[0]), _p(stats[1]), _stream()))
    return stats


def bn_update_running(mean, var, ra_mean, ra_var, momentum):
    check(lib().sav_bn_update_running(_p(mean), _p(var), _p(ra_mean), _p(ra_var), float(momentum),
                                      mean.numel(), _stream()))


def batchnorm_eval(y, ra_mean, ra_var, gamma, beta, eps):
    _req(y, BF16, "y")
    Cc = y.shape[-1]
    z = torch.empty_like(y)
    check(lib().sav_batchnorm_eval(_p(y), _p(ra_mean), _p(ra_var), _p(gamma), _p(beta), float(eps), _p(z),
                                   y.numel() // Cc, Cc, _stream()))
    return z


def layernorm_fwd_f32(x, scale, bias, eps=1e-6):
    """fp32 [rows, C] -> fp32 [rows, C] (LayerNorm into the residual stream, cvt.py:34)."""
    _req(x, F32, "x")
    rows, C_ = x.shape
    y = torch.empty((rows, C_), dtype=F32, device=x.device)
    mean = torch.empty(rows, dtype=F32, device=x.device)
    rstd = torch.empty(rows, dtype=F32, device=x.device)
    check(lib().sav_layernorm_fwd_f32(_p(x), x.stride(0), _p(scale), _p(bias), _p(y), y.stride(0),
                                      _p(mean), _p(rstd), rows, C_, eps, _stream()))
    return y, mean, rstd


def layernorm_bwd_f32(dy, x, scale, mean, rstd, dscale, dbias):
    """dy fp32 (stream gradient) -> (dx fp32, dx16 bf16)."""
    _req(dy, F32, "dy"); _req(x, F32, "x")
    rows, C_ = x.shape
    dx = torch.empty_like(x)
    dx16 = torch.empty((rows, C_), dtype=BF16, device=x.device)
    check(lib().sav_layernorm_bwd_f32(_p(dy), dy.stride(0), _p(x), x.stride(0), _p(scale), _p(mean),
                                      _p(rstd), None, _p(dx), _p(dx16), _p(dscale), _p(dbias), rows, C_,
                                      _stream()))
    return dx, dx16


def cast2d_bf16_f32(src, dst, accumulate=False):
    rows, cols = src.shape
    assert dst.shape == src.shape and src.stride(1) == 1 and dst.stride(1) == 1
    check(lib().sav_cast2d_bf16_f32(_p(src), src.stride(0), _p(dst), dst.stride(0), rows, cols,
                                    int(accumulate), _stream()))


def pixel_patchify(images, ph, pw, th, tw, hwcn):
    """tnt.py:19-26: fp32 images -> bf16 [N*(H/ph)*(W/pw), (ph/th)*(pw/tw), C*th*tw]."""
    _req(images, F32, "images")
    if hwcn:
        H, W, Cc, N = images.shape
    else:
        N, H, W, Cc = images.shape
    out = torch.empty((N * (H // ph) * (W // pw), (ph // th) * (pw // tw), Cc * th * tw), dtype=BF16,
                      device=images.device)
    check(lib().sav_pixel_patchify(_p(images), _p(out), N, H, W, Cc, ph, pw, th, tw, int(hwcn), _stream()))
    return out


def attn_small_fwd(qkv5, scale):
    """qkv5 bf16 contiguous [B, N, 3, H, D] -> o [B, N, H, D] (N <= 32, D <= 16)."""
    _req(qkv5, BF16, "qkv")
    B, N, _, H, D = qkv5.shape
    o = torch.empty((B, N, H, D), dtype=BF16, device=qkv5.device)
    check(lib().sav_attn_small_fwd(_p(qkv5), _p(o), B, N, H, D, float(scale), _stream()))
    return o


def attn_small_bwd(qkv5, d_o, scale):
    _req(qkv5, BF16, "qkv"); _req(d_o, BF16, "d_o")
    B, N, _, H, D = qkv5.shape
    dqkv = torch.empty_like(qkv5)
    check(lib().sav_attn_small_bwd(_p(qkv5), _p(d_o), _p(dqkv), B, N, H, D, float(scale), _stream()))
    return dqkv


def transpose_pad(src, R, C, out, out_cols, residual=None):
    """out[b, c, r] = src[b, r, c] (+ residual[b, c, r]); columns R..out_cols-1 of `out` zeroed.
    src: [B, R, >=C] (bf16 or fp32, pitch = stride(1)); out: [B, C, >=out_cols]; batches dense."""
    B = src.shape[0]
    assert src.dim() == 3 and out.dim() == 3 and src.stride(2) == 1 and out.stride(2) == 1
    assert B == 1 or (src.stride(0) == R * src.stride(1) and out.stride(0) == C * out.stride(1))
    assert src.shape[1] >= R and out.shape[1] >= C and out.shape[0] == B
    assert src.dtype in (BF16, F32) and out.dtype in (BF16, F32)
    if residual is not None:
        _req(residual, F32, "residual")
        assert residual.stride(2) == 1 and (B == 1 or residual.stride(0) == C * residual.stride(1))
    check(lib().sav_transpose_pad(_p(src), int(src.dtype == F32), src.stride(1), _p(out),
                                  int(out.dtype == F32), out.stride(1),
                                  _p(residual) if residual is not None else None,
                                  residual.stride(1) if residual is not None else 0,
                                  B, R, C, out_cols, _stream()))
    return out


# ------------------------------------------------------------------ BoTNet conv-stage pieces (cnn.cu)
ACT = {"none": 0, "swish": 1, "sigmoid": 2, "relu": 3, "gelu": 4}


def act_fwd(x, kind):
    _req(x, BF16, "x")
    y = torch.empty_like(x)
    check(lib().sav_act_fwd(_p(x), _p(y), x.numel(), ACT[kind], _stream()))
    return y


def act_bwd(x, dy, kind):
    dx = torch.empty_like(x)
    check(lib().sav_act_bwd(_p(x), _p(dy), _p(dx), x.numel(), ACT[kind], _stream()))
    return dx


def bn_finalize(stats, rows, eps):
    """(sum, sumsq) [2, C] -> mrv fp32 [3, C] = (mean, rstd, biased var)."""
    Cc = stats.shape[1]
    mrv = torch.empty((3, Cc), dtype=F32, device=stats.device)
    check(lib().sav_bn_finalize(_p(stats[0]), _p(stats[1]), rows, float(eps), _p(mrv[0]), _p(mrv[1]),
                                _p(mrv[2]), Cc, _stream()))
    return mrv


def bn_act_fwd(y, mean, rstd, gamma, beta, kind):
    _req(y, BF16, "y")
    Cc = y.shape[-1]
    z = torch.empty_like(y)
    check(lib().sav_bn_act_fwd(_p(y), _p(mean), _p(rstd), _p(gamma), _p(beta), _p(z), y.numel() // Cc, Cc,
                               ACT[kind], _stream()))
    return z


def bn_act_bwd(dz, y, mean, rstd, gamma, beta, dgamma, dbeta, kind):
    Cc = y.shape[-1]
    dy = torch.empty_like(y)
    scratch = torch.empty(2 * Cc, dtype=F32, device=y.device)
    check(lib().sav_bn_act_bwd(_p(dz), _p(y), _p(mean), _p(rstd), _p(gamma), _p(beta), _p(dgamma), _p(dbeta),
                               _p(scratch), _p(dy), y.numel() // Cc, Cc, ACT[kind], _stream()))
    return dy


def gate_res_act_fwd(x, gate, res, kind):
    """x bf16 [B, ..., C]; gate bf16 [B, C] or None; res like x or None."""
    B, Cc = x.shape[0], x.shape[-1]
    out = torch.empty_like(x)
    check(lib().sav_gate_res_act_fwd(_p(x), _p(gate), _p(res), _p(out), B, x.numel() // (B * Cc), Cc,
                                     ACT[kind], _stream()))
    return out


def gate_res_act_bwd(dout, x, gate, res, kind, want_dres):
    B, Cc = x.shape[0], x.shape[-1]
    dx = torch.empty_like(x)
    dres = torch.empty_like(x) if want_dres else None
    dgate = torch.empty((B, Cc), dtype=F32, device=x.device) if gate is not None else None
    check(lib().sav_gate_res_act_bwd(_p(dout), _p(x), _p(gate), _p(res), _p(dx), _p(dres), _p(dgate), B,
                                     x.numel() // (B * Cc), Cc, ACT[kind], _stream()))
    return dx, dres, dgate


def spatial_mean_fwd(x):
    B, Cc = x.shape[0], x.shape[-1]
    out = torch.empty((B, Cc), dtype=BF16, device=x.device)
    check(lib().sav_spatial_mean_fwd(_p(x), _p(out), B, x.numel() // (B * Cc), Cc, _stream()))
    return out


def spatial_mean_bwd(dout, shape):
    B, Cc = shape[0], shape[-1]
    dx = torch.empty(shape, dtype=BF16, device=dout.device)
    check(lib().sav_spatial_mean_bwd(_p(dout), _p(dx), B, dx.numel() // (B * Cc), Cc, _stream()))
    return dx


def maxpool3x3s2_fwd(x, valid=False):
    B, H, W, Cc = x.shape
    if valid:
        Ho, Wo = (H - 3) // 2 + 1, (W - 3) // 2 + 1
    else:
        Ho, _ = conv_same_geom(H, 3, 2)
        Wo, _ = conv_same_geom(W, 3, 2)
    y = torch.empty((B, Ho, Wo, Cc), dtype=BF16, device=x.device)
    arg = torch.empty((B, Ho, Wo, Cc), dtype=torch.uint8, device=x.device)
    fn = lib().sav_maxpool3x3s2_valid_fwd if valid else lib().sav_maxpool3x3s2_fwd
    check(fn(_p(x), _p(y), _p(arg), B, H, W, Cc, _stream()))
    return y, arg


def maxpool3x3s2_bwd(dy, arg, shape, valid=False):
    B, H, W, Cc = shape
    dx = torch.empty(shape, dtype=BF16, device=dy.device)
    fn = lib().sav_maxpool3x3s2_valid_bwd if valid else lib().sav_maxpool3x3s2_bwd
    check(fn(_p(dy), _p(arg), _p(dx), B, H, W, Cc, _stream()))
    return dx


def patchify_act(src, B, H, W, Cc, ph, pw, inverse=False):
    """bf16 [B,H,W,C] -> [B, (H/ph)(W/pw), ph*pw*C] (image_to_token.py:38-43), or back."""
    _req(src, BF16, "src")
    shape = (B, H, W, Cc) if inverse else (B, (H // ph) * (W // pw), ph * pw * Cc)
    dst = torch.empty(shape, dtype=BF16, device=src.device)
    check(lib().sav_patchify_act(_p(src), _p(dst), B, H, W, Cc, ph, pw, int(inverse), _stream()))
    return dst
