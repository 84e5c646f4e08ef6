"""ORACLE (test infrastructure, NOT product code) -- fp32 CPU restatement of the CaiT path
(models/cait.py) and of the BoTNet MHSA with 2-D relative position logits (models/botnet.py:70-199).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline legs may import this module.

PARITY UNPINNED for the same reason as oracle/vit_oracle.py: the reference ships no golden values
(cait_test.py / botnet_test.py assert logits shape only) and jax / flax are not installable here.
Validated by float64 finite differences (tests/test_variants_oracle_cpu.py), by the index identity
abs[i, j] = rel[i, j - i + l - 1] of `_to_absolute_logits` (botnet.py:78-93, restated literally
below as pad / reshape / slice and compared with the direct gather), and by the zero-head KAT.

BoTNet is broken as written; `bot_mhsa` follows the intended math: softmax over all
H*W keys, O = P V with V indexed by key, relative terms computed from the scaled query.
"""
import math

import torch

from oracle.vit_oracle import (attention, ff_block, layer_norm, loss_fn, hwcn_to_nhwc, patch_embed,
                               softmax_last, tree_map)


# ----------------------------------------------------------------------------- CaiT
def drop_path(x, factor):
    """StochasticDepthBlock (stochastic_depth.py:16-27) with the per-sample factor
    floor(keep + U) / keep supplied explicitly; None = eval / rate 0."""
    if factor is None:
        return x
    return x * factor.view(-1, *([1] * (x.dim() - 1)))


def cait_encoder_block(x, p, num_heads, factors=(None, None)):
    """cait.py EncoderBlock.__call__ (cait.py:29-53)."""
    sa = p["SelfAttentionBlock_0"]
    talking = (sa["TalkingHeadsBlock_0"]["talking_heads_transform"],
               sa["TalkingHeadsBlock_1"]["talking_heads_transform"])
    y = layer_norm(x, p["LayerNorm_0"]["scale"], p["LayerNorm_0"]["bias"])        # cait.py:30
    y = attention(y, y, sa, num_heads, talking=talking)                            # cait.py:31-35
    y = y * p["LayerScaleBlock_0"]["layerscale"]                                   # cait.py:36
    x = drop_path(y, factors[0]) + x                                               # cait.py:38-40
    y = layer_norm(x, p["LayerNorm_1"]["scale"], p["LayerNorm_1"]["bias"])        # cait.py:42
    y = ff_block(y, p["FFBlock_0"])                                                # cait.py:43-46
    y = y * p["LayerScaleBlock_1"]["layerscale"]                                   # cait.py:47
    return x + drop_path(y, factors[1])                                            # cait.py:49-52


def cait_ca_block(x, cls_token, p, num_heads, factors=(None, None)):
    """CAEncoderBlock.__call__ (cait.py:96-122); ClassSelfAttentionBlock = q from row 0 only
    (cait.py:10-15)."""
    z = torch.cat([cls_token, x], dim=1)                                           # cait.py:97
    z = layer_norm(z, p["LayerNorm_0"]["scale"], p["LayerNorm_0"]["bias"])        # cait.py:98
    a = attention(z[:, 0:1, :], z, p["ClassSelfAttentionBlock_0"], num_heads)     # cait.py:99-103
    a = a * p["LayerScaleBlock_0"]["layerscale"]                                   # cait.py:105
    cls_token = cls_token + drop_path(a, factors[0])                               # cait.py:107-109
    y = layer_norm(cls_token, p["LayerNorm_1"]["scale"], p["LayerNorm_1"]["bias"])
    y = ff_block(y, p["FFBlock_0"])
    y = y * p["LayerScaleBlock_1"]["layerscale"]                                   # cait.py:116
    return cls_token + drop_path(y, factors[1])                                    # cait.py:121


def cait_forward(params, images_nhwc, cfg, factors=None):
    """CaiT.__call__ (cait.py:141-183) + Encoder.__call__ (cait.py:68-83).  factors: optional
    list of per-sample drop-path factor tensors [B], two per block in execution order."""
    ph, pw = cfg["patch_shape"]
    it = iter(factors) if factors is not None else None
    nxt = (lambda: (next(it), next(it))) if it is not None else (lambda: (None, None))
    x = patch_embed(images_nhwc, params["PatchEmbedBlock_0"]["Dense_0"]["kernel"], ph, pw)  # :143
    enc = params["Encoder_0"]
    x = x + enc["AddAbsPosEmbed_0"]["pos_embed"]                                   # cait.py:69
    for i in range(cfg["num_layers"]):                                             # cait.py:72-80
        x = cait_encoder_block(x, enc[f"EncoderBlock_{i}"], cfg["num_heads"], nxt())
    cls_token = params["cls"].expand(x.shape[0], -1, -1)                           # cait.py:157-160
    for i in range(cfg["num_layers_token_only"]):                                  # cait.py:162-173
        cls_token = cait_ca_block(x, cls_token, params[f"CAEncoderBlock_{i}"], cfg["num_heads"],
                                  nxt())
    z = torch.cat([cls_token, x], dim=1)                                           # cait.py:175
    z = layer_norm(z, params["LayerNorm_0"]["scale"], params["LayerNorm_0"]["bias"])
    return z[:, 0] @ params["Dense_0"]["kernel"] + params["Dense_0"]["bias"]       # cait.py:178-182


def cait_config(name):
    cfgs = {
        "cait_xxs_24": dict(num_layers=24, num_heads=4, embed_dim=192, layerscale_eps=1e-5),
        "cait_xs_24": dict(num_layers=24, num_heads=6, embed_dim=288, layerscale_eps=1e-5),
        "cait_s_24": dict(num_layers=24, num_heads=8, embed_dim=384, layerscale_eps=1e-6),
    }
    c = dict(cfgs[name])
    c.update(num_layers_token_only=2, patch_shape=(16, 16), num_classes=1000, expand_ratio=4)
    return c


def _orthogonal(g, n, dtype):
    a = torch.randn((n, n), generator=g, dtype=torch.float64)
    q, r = torch.linalg.qr(a)
    return (q * torch.sign(torch.diagonal(r))).to(dtype).contiguous()


def synthetic_cait_params(cfg, img_size=224, seed=42, layerscale=None, dtype=torch.float32):
    """SURVEY 8d synthetic (non-default) init with the Flax auto-names of cait.py; talking-heads
    transforms are random orthogonal plus N(0, 0.1) so they are not exactly orthogonal; layerscale
    = cfg eps unless given (the parity tests also use 0.1, SURVEY 7 'Tolerance')."""
    g = torch.Generator().manual_seed(seed)
    C, H = cfg["embed_dim"], cfg["num_heads"]
    d = C // H
    ph, pw = cfg["patch_shape"]
    n_tok = (img_size // ph) * (img_size // pw)
    hid = int(cfg["expand_ratio"] * C)
    ls = cfg["layerscale_eps"] if layerscale is None else layerscale

    def normal(shape, std):
        return (torch.randn(shape, generator=g, dtype=torch.float64) * std).to(dtype)

    def kern(shape, fan_in):
        return normal(shape, 1.0 / math.sqrt(fan_in))

    def ln():
        return {"scale": 1.0 + normal((C,), 0.02), "bias": normal((C,), 0.02)}

    def attn(talking):
        p = {"queries": {"kernel": kern((C, H, d), C)}, "keys": {"kernel": kern((C, H, d), C)},
             "values": {"kernel": kern((C, H, d), C)},
             "DenseGeneral_0": {"kernel": kern((H, d, C), C)}}
        if talking:
            for j in range(2):
                p[f"TalkingHeadsBlock_{j}"] = {
                    "talking_heads_transform": (_orthogonal(g, H, dtype) + normal((H, H), 0.1)).contiguous()}
        return p

    def lscale():
        return {"layerscale": torch.full((C,), ls, dtype=dtype) * (1.0 + normal((C,), 0.1))}

    def ffn():
        return {"Dense_0": {"kernel": kern((C, hid), C), "bias": normal((hid,), 0.02)},
                "Dense_1": {"kernel": kern((hid, C), hid), "bias": normal((C,), 0.02)}}

    enc = {"AddAbsPosEmbed_0": {"pos_embed": normal((1, n_tok, C), 0.02)}}
    for i in range(cfg["num_layers"]):
        enc[f"EncoderBlock_{i}"] = {"LayerNorm_0": ln(), "SelfAttentionBlock_0": attn(True),
                                    "LayerScaleBlock_0": lscale(), "LayerNorm_1": ln(),
                                    "FFBlock_0": ffn(), "LayerScaleBlock_1": lscale()}
    params = {
        "PatchEmbedBlock_0": {"Dense_0": {"kernel": kern((ph * pw * 3, C), ph * pw * 3)}},
        "Encoder_0": enc,
        "cls": normal((1, 1, C), 0.02),
    }
    for i in range(cfg["num_layers_token_only"]):
        params[f"CAEncoderBlock_{i}"] = {"LayerNorm_0": ln(), "ClassSelfAttentionBlock_0": attn(False),
                                         "LayerScaleBlock_0": lscale(), "LayerNorm_1": ln(),
                                         "FFBlock_0": ffn(), "LayerScaleBlock_1": lscale()}
    params["LayerNorm_0"] = ln()
    nc = cfg["num_classes"]
    params["Dense_0"] = {"kernel": normal((C, nc), 0.02), "bias": normal((nc,), 0.02)}
    return params


def cait_loss_and_grads(params, images_hwcn, labels, cfg, label_smoothing=0.0, factors=None):
    """train.py:79-95 applied to CaiT."""
    def req(t):
        return t.detach().clone().requires_grad_(True)
    p = tree_map(req, params)
    logits = cait_forward(p, hwcn_to_nhwc(images_hwcn), cfg, factors)
    loss = loss_fn(logits, labels, cfg["num_classes"], label_smoothing)
    loss.backward()
    grads = tree_map(lambda t: t.grad if t.grad is not None else torch.zeros_like(t), p)
    return loss.detach(), logits.detach(), grads


# ----------------------------------------------------------------------------- BoTNet MHSA
def to_absolute_logits(rel):
    """RelativeLogits._to_absolute_logits (botnet.py:78-93), restated literally: rel
    [b, h, l, 2l-1] -> abs [b, h, l, l] by zero-pad one column, flatten, pad l-1, reshape to
    [l+1, 2l-1], keep [:l, l-1:]."""
    b, h, l, _ = rel.shape
    x = torch.cat([rel, rel.new_zeros(b, h, l, 1)], dim=3)                         # botnet.py:81-82
    x = x.reshape(b, h, l * 2 * l)                                                 # botnet.py:84
    x = torch.cat([x, rel.new_zeros(b, h, l - 1)], dim=2)                          # botnet.py:85-86
    x = x.reshape(b, h, l + 1, 2 * l - 1)                                          # botnet.py:88-89
    return x[:, :, :l, l - 1:]                                                     # botnet.py:90-92


def relative_logits_1d(q, rel_k):
    """RelativeLogits._relative_logits_1d (botnet.py:96-111) for q [b, h, H, W, d] along W:
    returns [b, h, H, W, W] with out[..., x, kx] = q[..., x, :] . rel_k[kx - x + W - 1]."""
    b, h, H, W, d = q.shape
    rel = torch.einsum("bhxyd,md->bhxym", q, rel_k)                                # botnet.py:105
    rel = rel.reshape(b, h * H, W, 2 * W - 1)                                      # botnet.py:106
    return to_absolute_logits(rel).reshape(b, h, H, W, W)                          # botnet.py:107-108


def relative_logits_2d(q, rel_w, rel_h):
    """RelativeLogits.__call__ (botnet.py:114-141): [b, h, H, W, H, W] (query y,x ; key y,x)."""
    b, h, H, W, d = q.shape
    lw = relative_logits_1d(q, rel_w)                                              # [b,h,H,W,Wk]
    lh = relative_logits_1d(q.permute(0, 1, 3, 2, 4), rel_h)                       # [b,h,W,H,Hk]
    lw = lw[:, :, :, :, None, :].expand(b, h, H, W, H, W)                          # tile over key rows
    lh = lh.permute(0, 1, 3, 2, 4)[:, :, :, :, :, None].expand(b, h, H, W, H, W)   # over key cols
    return lw + lh


def bot_mhsa(x, p, num_heads, head_ch):
    """BoTMHSA.__call__ intended semantics (botnet.py:158-199).  x [b, H, W, C];
    p: {'query','key','value': {'kernel' [1, 1, C, h*d]} (1x1 nn.Conv, no bias, botnet.py:165-174),
    'RelativeLogits_0': {'rel_pos_emb_w' [2W-1, d], 'rel_pos_emb_h' [2H-1, d]}} (botnet.py:124-132)."""
    b, H, W, C = x.shape
    h, d = num_heads, head_ch

    def proj(name):
        y = x.reshape(b, H * W, C) @ p[name]["kernel"].reshape(C, h * d)           # 1x1 conv
        return y.reshape(b, H, W, h, d).permute(0, 3, 1, 2, 4)                     # b h H W d
    q, k, v = proj("query"), proj("key"), proj("value")                            # botnet.py:175-183
    q = q / math.sqrt(d)                                                           # botnet.py:185
    logits = torch.einsum("bhxyd,bhpqd->bhxypq", q, k)                             # botnet.py:187-190
    rl = p["RelativeLogits_0"]
    logits = logits + relative_logits_2d(q, rl["rel_pos_emb_w"], rl["rel_pos_emb_h"])      # botnet.py:191-192
    w = softmax_last(logits.reshape(b, h, H, W, H * W))                            # intended (D2)
    o = torch.einsum("bhxyk,bhkd->bhxyd", w, v.reshape(b, h, H * W, d))            # O = P V
    return o.permute(0, 2, 3, 1, 4).reshape(b, H, W, h * d)                        # botnet.py:198


def bot_mhsa_direct(x, p, num_heads, head_ch):
    """Same result by direct gather logits += q.r^W[kx-qx+W-1] + q.r^H[ky-qy+H-1] --
    the form the kernel implements; used to pin the pad/reshape restatement above."""
    b, H, W, C = x.shape
    h, d = num_heads, head_ch
    xf = x.reshape(b, H * W, C)
    q = (xf @ p["query"]["kernel"].reshape(C, h * d)).reshape(b, H * W, h, d) / math.sqrt(d)
    k = (xf @ p["key"]["kernel"].reshape(C, h * d)).reshape(b, H * W, h, d)
    v = (xf @ p["value"]["kernel"].reshape(C, h * d)).reshape(b, H * W, h, d)
    s = torch.einsum("bqhd,bkhd->bhqk", q, k)
    rl = p["RelativeLogits_0"]
    rw = torch.einsum("bqhd,md->bhqm", q, rl["rel_pos_emb_w"])
    rh = torch.einsum("bqhd,md->bhqm", q, rl["rel_pos_emb_h"])
    idx = torch.arange(H * W)
    qy, qx = idx // W, idx % W
    iw = (qx[None, :] - qx[:, None]) + W - 1          # [q, k] -> kx - qx + W - 1
    ih = (qy[None, :] - qy[:, None]) + H - 1
    s = s + torch.gather(rw, 3, iw.expand(b, h, -1, -1)) + torch.gather(rh, 3, ih.expand(b, h, -1, -1))
    w = softmax_last(s)
    o = torch.einsum("bhqk,bkhd->bqhd", w, v)
    return o.reshape(b, H, W, h * d)


def synthetic_bot_params(C, num_heads, head_ch, H, W, seed=42, dtype=torch.float32):
    """1x1-conv kernels N(0, 1/C); relative tables N(0, d^-1/2) (botnet.py:124-132)."""
    g = torch.Generator().manual_seed(seed)

    def normal(shape, std):
        return (torch.randn(shape, generator=g, dtype=torch.float64) * std).to(dtype)
    hd = num_heads * head_ch
    return {
        "query": {"kernel": normal((1, 1, C, hd), 1.0 / math.sqrt(C))},
        "key": {"kernel": normal((1, 1, C, hd), 1.0 / math.sqrt(C))},
        "value": {"kernel": normal((1, 1, C, hd), 1.0 / math.sqrt(C))},
        "RelativeLogits_0": {"rel_pos_emb_w": normal((2 * W - 1, head_ch), head_ch ** -0.5),
                             "rel_pos_emb_h": normal((2 * H - 1, head_ch), head_ch ** -0.5)},
    }


# ----------------------------------------------------------------------------- CvT
def same_pad_amounts(n, stride, k=3):
    """jax / flax padding='SAME': out = ceil(n / stride); total = max((out-1)*stride + k - n, 0);
    low = total // 2, high = total - low."""
    out = -(-n // stride)
    total = max((out - 1) * stride + k - n, 0)
    return total // 2, total - total // 2


def conv_projection(x, p, stride, eps=1e-5):
    """ConvProjectionBlock.__call__ (cvt_attention.py:21-40), training mode: depthwise 3x3 nn.Conv
    (feature_group_count = C, padding SAME, no bias; kernel [3,3,1,C]) -> nn.BatchNorm on batch
    statistics (biased variance over B,H,W; epsilon 1e-5) -> 1x1 nn.Conv without bias
    (kernel [1,1,C,out]).  x [B,H,W,C]."""
    B, H, W, C = x.shape
    pt, pb = same_pad_amounts(H, stride)
    pl, pr = same_pad_amounts(W, stride)
    xi = torch.nn.functional.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
    wk = p["Conv_0"]["kernel"].permute(3, 2, 0, 1)                               # [C,1,3,3]
    y = torch.nn.functional.conv2d(xi, wk, stride=stride, groups=C).permute(0, 2, 3, 1)
    mean = y.mean(dim=(0, 1, 2))
    var = (y * y).mean(dim=(0, 1, 2)) - mean * mean                              # flax: E[x^2]-E[x]^2
    z = (y - mean) * torch.rsqrt(var + eps) * p["BatchNorm_0"]["scale"] + p["BatchNorm_0"]["bias"]
    return z @ p["Conv_1"]["kernel"].reshape(C, -1)                              # 1x1 conv


def cvt_attention(x, p, num_heads, strides=(1, 2, 2)):
    """CvTAttentionBlock.__call__ (cvt_attention.py:58-113) with inputs_q = inputs_kv = x
    [B,H,W,C] (CvTSelfAttentionBlock, :116-120)."""
    B = x.shape[0]
    q, k, v = (conv_projection(x, p[f"ConvProjectionBlock_{i}"], s) for i, s in enumerate(strides))
    hd = q.shape[-1]
    d = hd // num_heads
    q, k, v = (t.reshape(B, -1, num_heads, d) for t in (q, k, v))                # :81-83
    q = q / math.sqrt(d)                                                         # :85
    w = softmax_last(torch.einsum("bqhd,bkhd->bhqk", q, k))                      # :87-94
    o = torch.einsum("bhqk,bkhd->bqhd", w, v)                                    # :103-104
    return torch.einsum("bqhd,hdc->bqc", o, p["DenseGeneral_0"]["kernel"])       # :106-109


def cvt_stage_block(tokens, p, num_heads):
    """StageBlock.__call__ (cvt.py:50-68) incl. zero_pad_and_reshape (cvt.py:10-16)."""
    B, L, C = tokens.shape
    S = int(math.ceil(math.sqrt(L)))
    padded = torch.cat([tokens, tokens.new_zeros(B, S * S - L, C)], dim=1)       # cvt.py:14
    grid = padded.reshape(B, S, S, C)                                            # cvt.py:15
    x = cvt_attention(grid, p["CvTSelfAttentionBlock_0"], num_heads) + padded    # cvt.py:53-61
    y = layer_norm(x, p["LayerNorm_0"]["scale"], p["LayerNorm_0"]["bias"])       # cvt.py:63
    return x + ff_block(y, p["FFBlock_0"])                                       # cvt.py:64-67


def synthetic_cvt_stage_params(C, num_heads, seed=42, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)

    def normal(shape, std):
        return (torch.randn(shape, generator=g, dtype=torch.float64) * std).to(dtype)
    d = C // num_heads
    hid = 4 * C

    def proj():
        return {"Conv_0": {"kernel": normal((3, 3, 1, C), 1.0 / 3.0)},
                "BatchNorm_0": {"scale": 1.0 + normal((C,), 0.1), "bias": normal((C,), 0.1)},
                "Conv_1": {"kernel": normal((1, 1, C, C), 1.0 / math.sqrt(C))}}
    return {
        "CvTSelfAttentionBlock_0": {"ConvProjectionBlock_0": proj(), "ConvProjectionBlock_1": proj(),
                                    "ConvProjectionBlock_2": proj(),
                                    "DenseGeneral_0": {"kernel": normal((num_heads, d, C), 1.0 / math.sqrt(C))}},
        "LayerNorm_0": {"scale": 1.0 + normal((C,), 0.02), "bias": normal((C,), 0.02)},
        "FFBlock_0": {"Dense_0": {"kernel": normal((C, hid), 1.0 / math.sqrt(C)), "bias": normal((hid,), 0.02)},
                      "Dense_1": {"kernel": normal((hid, C), 1.0 / math.sqrt(hid)), "bias": normal((C,), 0.02)}},
    }


def conv_projection_eval(x, p, stride, ra_mean, ra_var, eps=1e-5):
    """ConvProjectionBlock with use_running_average=True (is_training=False, cvt_attention.py:33):
    the BatchNorm normalises with the running averages."""
    B, H, W, C = x.shape
    pt, pb = same_pad_amounts(H, stride)
    pl, pr = same_pad_amounts(W, stride)
    xi = torch.nn.functional.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
    wk = p["Conv_0"]["kernel"].permute(3, 2, 0, 1)
    y = torch.nn.functional.conv2d(xi, wk, stride=stride, groups=C).permute(0, 2, 3, 1)
    z = (y - ra_mean) * torch.rsqrt(ra_var + eps) * p["BatchNorm_0"]["scale"] + p["BatchNorm_0"]["bias"]
    return z @ p["Conv_1"]["kernel"].reshape(C, -1)


def conv_projection_batch_stats(x, p, stride):
    """(mean, biased var) over B,H,W of the depthwise conv output -- what flax nn.BatchNorm folds into
    its running averages: ra <- momentum * ra + (1 - momentum) * stat."""
    B, H, W, C = x.shape
    pt, pb = same_pad_amounts(H, stride)
    pl, pr = same_pad_amounts(W, stride)
    xi = torch.nn.functional.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
    wk = p["Conv_0"]["kernel"].permute(3, 2, 0, 1)
    y = torch.nn.functional.conv2d(xi, wk, stride=stride, groups=C).permute(0, 2, 3, 1)
    mean = y.mean(dim=(0, 1, 2))
    return mean, (y * y).mean(dim=(0, 1, 2)) - mean * mean


def conv2d_same(x, kernel, bias, stride):
    """nn.Conv(features, (k,k), strides=(s,s), padding='SAME') on [B,H,W,Cin] with a Flax kernel
    [kh,kw,cin,cout] (cvt.py:27-31)."""
    k = kernel.shape[0]
    B, H, W, C = x.shape
    pt, pb = same_pad_amounts(H, stride, k)
    pl, pr = same_pad_amounts(W, stride, k)
    xi = torch.nn.functional.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
    y = torch.nn.functional.conv2d(xi, kernel.permute(3, 2, 0, 1), bias=bias, stride=stride)
    return y.permute(0, 2, 3, 1)


def conv_token_embed(x, p, stride):
    """ConvTokenEmbedBlock.__call__ (cvt.py:25-35): conv -> 'b H W c -> b (H W) c' -> LayerNorm."""
    y = conv2d_same(x, p["Conv_0"]["kernel"], p["Conv_0"]["bias"], stride)      # cvt.py:27-31
    y = y.reshape(y.shape[0], -1, y.shape[-1])                                   # cvt.py:32
    return layer_norm(y, p["LayerNorm_0"]["scale"], p["LayerNorm_0"]["bias"])    # cvt.py:34


def cvt_forward(params, images_nhwc, cfg):
    """CvT.__call__ (cvt.py:131-171) + Stage.__call__ (cvt.py:86-113), training mode."""
    x = images_nhwc
    n = len(cfg["stage_sizes"])
    for i in range(n):
        sp = params[f"Stage_{i}"]
        x = conv_token_embed(x, sp["ConvTokenEmbedBlock_0"], cfg["embed_strides"][i])   # cvt.py:88-91
        if i == n - 1:                                                           # insert_cls, cvt.py:93-98
            x = torch.cat([sp["cls"].expand(x.shape[0], -1, -1), x], dim=1)
        for j in range(cfg["stage_sizes"][i]):                                   # cvt.py:100-110
            x = cvt_stage_block(x, sp[f"StageBlock_{j}"], cfg["num_heads"][i])
        if i < n - 1:
            S = int(math.sqrt(x.shape[1]))                                       # cvt.py:147-149
            x = x.reshape(x.shape[0], S, S, x.shape[-1])
    cls_token = x[:, 0]                                                          # cvt.py:165
    return cls_token @ params["Dense_0"]["kernel"] + params["Dense_0"]["bias"]   # cvt.py:166-170


def cvt_config(name):
    cfgs = {
        "cvt-13": dict(stage_sizes=(1, 2, 10), num_heads=(1, 3, 6), embed_dim=(64, 192, 384)),
        "cvt-21": dict(stage_sizes=(1, 4, 16), num_heads=(1, 3, 6), embed_dim=(64, 192, 384)),
        "cvt-w24": dict(stage_sizes=(2, 2, 20), num_heads=(3, 12, 16), embed_dim=(192, 768, 1024)),
    }
    c = dict(cfgs[name])
    c.update(embed_kernel_size=(7, 3, 3), embed_strides=(4, 2, 2), num_classes=1000)
    return c


def synthetic_cvt_params(cfg, seed=42, dtype=torch.float32):
    """SURVEY 8d synthetic init with the Flax auto-names of cvt.py."""
    g = torch.Generator().manual_seed(seed)

    def normal(shape, std):
        return (torch.randn(shape, generator=g, dtype=torch.float64) * std).to(dtype)
    params = {}
    cin = 3
    n = len(cfg["stage_sizes"])
    for i in range(n):
        C, k = cfg["embed_dim"][i], cfg["embed_kernel_size"][i]
        sp = {"ConvTokenEmbedBlock_0": {
            "Conv_0": {"kernel": normal((k, k, cin, C), 1.0 / math.sqrt(k * k * cin)), "bias": normal((C,), 0.02)},
            "LayerNorm_0": {"scale": 1.0 + normal((C,), 0.02), "bias": normal((C,), 0.02)}}}
        if i == n - 1:
            sp["cls"] = normal((1, 1, C), 0.02)
        for j in range(cfg["stage_sizes"][i]):
            sp[f"StageBlock_{j}"] = synthetic_cvt_stage_params(C, cfg["num_heads"][i],
                                                               seed=seed + 100 * i + j + 1, dtype=dtype)
        params[f"Stage_{i}"] = sp
        cin = C
    nc = cfg["num_classes"]
    params["Dense_0"] = {"kernel": normal((cin, nc), 0.02), "bias": normal((nc,), 0.02)}
    return params


def cvt_loss_and_grads(params, images_hwcn, labels, cfg, label_smoothing=0.0):
    from .vit_oracle import hwcn_to_nhwc, loss_fn
    p = tree_map(lambda t: t.detach().clone().requires_grad_(True), params)
    logits = cvt_forward(p, hwcn_to_nhwc(images_hwcn), cfg)
    loss = loss_fn(logits, labels, cfg["num_classes"], label_smoothing)
    loss.backward()
    grads = tree_map(lambda t: t.grad if t.grad is not None else torch.zeros_like(t), p)
    return loss.detach(), logits.detach(), grads


# ----------------------------------------------------------------------------- BoTNet conv stages
class _RoundBf16(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x.bfloat16().float()

    @staticmethod
    def backward(ctx, g):
        return g.bfloat16().float()


# When set, the conv-stage functions below round their results (and the gradients flowing back
# through them) to bf16 at exactly the tensors a bf16 pipeline stores: conv operands and outputs,
# BatchNorm(+activation) outputs, block outputs.  This "ideal bf16 pipeline" is the noise FLOOR the
# CUDA path is compared with (BatchNorm'd conv stacks amplify operand rounding far more than the
# LayerNorm'd transformer: its worst gradient is ~0.1 off the fp32 result at the test shapes).
EMULATE_BF16 = [False]


def _rb(x):
    return _RoundBf16.apply(x) if EMULATE_BF16[0] else x


def swish(x):
    """flax.linen.swish: x * sigmoid(x) (botnet.py:31,216,264)."""
    return _rb(x * torch.sigmoid(x))


def batch_norm_train(y, p, eps=1e-5):
    """flax nn.BatchNorm(use_running_average=False) (botnet.py:286-290): statistics over all axes
    but the last, biased variance E[x^2] - E[x]^2."""
    ax = tuple(range(y.dim() - 1))
    mean = y.mean(dim=ax)
    var = (y * y).mean(dim=ax) - mean * mean
    return _rb((y - mean) * torch.rsqrt(var + eps) * p["scale"] + p["bias"])


def conv2d_nobias(x, kernel, stride=1, padding="SAME"):
    """nn.Conv(use_bias=False) on [B,H,W,Cin] with a Flax kernel [kh,kw,cin,cout]; padding 'SAME'
    or an explicit symmetric amount (botnet.py:294)."""
    x, kernel = _rb(x), _rb(kernel)
    if padding == "SAME":
        return _rb(conv2d_same(x, kernel, None, stride))
    xi = torch.nn.functional.pad(x.permute(0, 3, 1, 2), (padding, padding, padding, padding))
    return _rb(torch.nn.functional.conv2d(xi, kernel.permute(3, 2, 0, 1), stride=stride).permute(0, 2, 3, 1))


def max_pool_3x3_s2_same(x):
    """nn.max_pool(y, (3, 3), strides=(2, 2), padding='SAME') (botnet.py:297): -inf padding."""
    B, H, W, C = x.shape
    pt, pb = same_pad_amounts(H, 2, 3)
    pl, pr = same_pad_amounts(W, 2, 3)
    xi = torch.nn.functional.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb), value=float("-inf"))
    return torch.nn.functional.max_pool2d(xi, 3, 2).permute(0, 2, 3, 1)


def squeeze_excite(x, p, act=swish):
    """SqueezeExciteBlock.__call__ (squeeze_excite.py:21-38), intended reading of :32-33:
    x = mean(inputs, (1, 2), keepdims=True) -> Dense -> act -> Dense -> sigmoid(x) * inputs."""
    s = x.mean(dim=(1, 2), keepdim=True)
    s = act(s @ p["Dense_0"]["kernel"] + p["Dense_0"]["bias"])
    s = s @ p["Dense_1"]["kernel"] + p["Dense_1"]["bias"]
    return torch.sigmoid(s) * x


def bottleneck_block(x, p, strides, se=True):
    """BottleneckResNetBlock.__call__ (botnet.py:34-67)."""
    residual = x
    y = conv2d_nobias(x, p["Conv_0"]["kernel"])                                   # :43
    y = swish(batch_norm_train(y, p["BatchNorm_0"]))                              # :44-45
    y = conv2d_nobias(y, p["Conv_1"]["kernel"], strides)                          # :46
    y = swish(batch_norm_train(y, p["BatchNorm_1"]))                              # :47-48
    y = conv2d_nobias(y, p["Conv_2"]["kernel"])                                   # :49-50
    y = batch_norm_train(y, p["BatchNorm_2"])                                     # :51
    if se:
        y = squeeze_excite(y, p["SqueezeExciteBlock_0"])                          # :53-56
    if residual.shape != y.shape:                                                 # :58-63
        residual = conv2d_nobias(residual, p["Conv_3"]["kernel"], strides)
        residual = swish(batch_norm_train(residual, p["BatchNorm_3"]))
    return swish(residual + y)                                                    # :65


def bot_block(x, p, num_heads, head_dim):
    """BoTBlock.__call__ (botnet.py:218-252) with strides (1, 1) (stride_one=True, every
    create_model configuration)."""
    residual = x
    y = conv2d_nobias(x, p["Conv_0"]["kernel"])                                   # :229
    y = swish(batch_norm_train(y, p["BatchNorm_0"]))                              # :230-231
    y = bot_mhsa(y, p["BoTMHSA_0"], num_heads, head_dim)                          # :232
    y = swish(batch_norm_train(y, p["BatchNorm_1"]))                              # :238-239
    y = conv2d_nobias(y, p["Conv_1"]["kernel"])                                   # :240-241
    y = batch_norm_train(y, p["BatchNorm_2"])                                     # :242
    if residual.shape != y.shape:                                                 # :244-249
        residual = conv2d_nobias(residual, p["Conv_2"]["kernel"])
        residual = swish(batch_norm_train(residual, p["BatchNorm_3"]))
    return swish(residual + y)                                                    # :251


def botnet_forward(params, images_nhwc, cfg):
    """BoTNet.__call__ (botnet.py:278-326), train=True."""
    y = conv2d_nobias(images_nhwc, params["Conv_0"]["kernel"], 2, 3)              # :292-295
    y = swish(batch_norm_train(y, params["BatchNorm_0"]))                         # :296-297
    y = max_pool_3x3_s2_same(y)                                                   # :298
    n = 0
    sizes = cfg["stage_sizes"]
    for i, block_size in enumerate(sizes[:-1]):                                   # :299-311
        for j in range(block_size):
            strides = 2 if i > 0 and j == 0 else 1
            y = bottleneck_block(y, params[f"BottleneckResNetBlock_{n}"], strides)
            n += 1
    for j in range(sizes[-1]):                                                    # :312-320
        y = bot_block(y, params[f"BoTBlock_{j}"], cfg["num_heads"], cfg["head_dim"])
    y = y.mean(dim=(1, 2))                                                        # :321
    return y @ params["Dense_0"]["kernel"] + params["Dense_0"]["bias"]            # :322-325


def synthetic_botnet_params(cfg, img_size, seed=42, dtype=torch.float32):
    """SURVEY 8d synthetic init with the Flax auto-names of botnet.py (the zero-initialised last
    BatchNorm scale of each block, botnet.py:51,242, is given a non-degenerate value)."""
    g = torch.Generator().manual_seed(seed)

    def normal(shape, std):
        return (torch.randn(shape, generator=g, dtype=torch.float64) * std).to(dtype)

    def conv(k, cin, cout):
        return {"kernel": normal((k, k, cin, cout), 1.0 / math.sqrt(k * k * cin))}

    def bn(C):
        return {"scale": 1.0 + normal((C,), 0.1), "bias": normal((C,), 0.1)}

    f0, pf = cfg["initial_filters"], 4
    sizes = cfg["stage_sizes"]
    params = {"Conv_0": conv(7, 3, f0), "BatchNorm_0": bn(f0)}
    cin, n = f0, 0
    for i, block_size in enumerate(sizes[:-1]):
        f = f0 * 2 ** i
        for j in range(block_size):
            strides = 2 if i > 0 and j == 0 else 1
            hid = max(1, int(f * pf * 0.0625))
            p = {"Conv_0": conv(1, cin, f), "BatchNorm_0": bn(f), "Conv_1": conv(3, f, f), "BatchNorm_1": bn(f),
                 "Conv_2": conv(1, f, f * pf), "BatchNorm_2": bn(f * pf),
                 "SqueezeExciteBlock_0": {
                     "Dense_0": {"kernel": normal((f * pf, hid), 1.0 / math.sqrt(f * pf)), "bias": normal((hid,), 0.02)},
                     "Dense_1": {"kernel": normal((hid, f * pf), 1.0 / math.sqrt(hid)), "bias": normal((f * pf,), 0.02)}}}
            if cin != f * pf or strides != 1:
                p["Conv_3"] = conv(1, cin, f * pf)
                p["BatchNorm_3"] = bn(f * pf)
            params[f"BottleneckResNetBlock_{n}"] = p
            n += 1
            cin = f * pf
    f = f0 * 2 ** (len(sizes) - 1)
    hw = img_size // 4
    for i in range(1, len(sizes) - 1):
        hw = -(-hw // 2)
    h, d = cfg["num_heads"], cfg["head_dim"]
    for j in range(sizes[-1]):
        mh = synthetic_bot_params(f, h, d, hw, hw, seed=seed + 10 + j, dtype=dtype)
        p = {"Conv_0": conv(1, cin, f), "BatchNorm_0": bn(f), "BoTMHSA_0": mh, "BatchNorm_1": bn(h * d),
             "Conv_1": conv(1, h * d, f * pf), "BatchNorm_2": bn(f * pf)}
        if cin != f * pf:
            p["Conv_2"] = conv(1, cin, f * pf)
            p["BatchNorm_3"] = bn(f * pf)
        params[f"BoTBlock_{j}"] = p
        cin = f * pf
    nc = cfg["num_classes"]
    params["Dense_0"] = {"kernel": normal((cin, nc), 0.02), "bias": normal((nc,), 0.02)}
    return params


def botnet_loss_and_grads(params, images_hwcn, labels, cfg, label_smoothing=0.0, emulate_bf16=False):
    from .vit_oracle import hwcn_to_nhwc, loss_fn
    p = tree_map(lambda t: t.detach().clone().requires_grad_(True), params)
    EMULATE_BF16[0] = bool(emulate_bf16)
    try:
        logits = botnet_forward(p, hwcn_to_nhwc(images_hwcn), cfg)
    finally:
        EMULATE_BF16[0] = False
    loss = loss_fn(logits, labels, cfg["num_classes"], label_smoothing)
    loss.backward()
    grads = tree_map(lambda t: t.grad if t.grad is not None else torch.zeros_like(t), p)
    return loss.detach(), logits.detach(), grads


# ----------------------------------------------------------------------------- rotary helpers
def rotate_every_two(x):
    """position_embed.py:8-14 literally: x1 = x[:, :, ::2], x2 = x[:, :, 1::2],
    stack((-x2, x1), -1) -> '... d j -> ... (d j)'."""
    x1 = x[..., ::2]
    x2 = x[..., 1::2]
    return torch.stack((-x2, x1), dim=-1).reshape(x.shape)


def fixed_positional_embedding(n, dim, dtype=torch.float32):
    """FixedPositionalEmbedding.__call__ (position_embed.py:26-34) literally, in float32 (the
    RotaryPositionalEmbedding dtype, position_embed.py:38): intervals = arange(0, dim, 2);
    inv_freq = 1 / (10e4**intervals / dim)  -- the power overflows to inf from 10e4**8 on, so those
    frequencies are exactly 0; freqs = einsum('i,j->ij', arange(n), inv_freq);
    emb = concatenate([freqs, freqs], -1)."""
    intervals = torch.arange(0, dim, 2, dtype=dtype)
    inv_freq = 1.0 / (torch.pow(torch.tensor(10e4, dtype=dtype), intervals) / dim)
    t = torch.arange(n, dtype=dtype)
    freqs = torch.einsum("i,j->ij", t, inv_freq)
    return torch.cat([freqs, freqs], dim=-1)


def apply_rotary_pos_emb(x, n_axis=0):
    """apply_rotary_pos_emb (position_embed.py:17-20) under the intended reading (the reference
    unpacks (sin, cos) from the single tensor FixedPositionalEmbedding returns, which cannot run):
    sin, cos = sin / cos of the [n, dim/2] frequencies, each repeated 'b n -> b (n j)', j=2 and
    broadcast over the axes between the sequence axis and the last; x cos + rotate_every_two(x) sin."""
    n, dim = x.shape[n_axis], x.shape[-1]
    freqs = fixed_positional_embedding(n, dim)[:, :dim // 2].to(x.dtype)
    sin = torch.repeat_interleave(torch.sin(freqs), 2, dim=-1)
    cos = torch.repeat_interleave(torch.cos(freqs), 2, dim=-1)
    shape = [1] * x.dim()
    shape[n_axis], shape[-1] = n, dim
    return x * cos.view(shape) + rotate_every_two(x) * sin.view(shape)


# ----------------------------------------------------------------------------- MLP-Mixer
def mixer_block(x, p):
    """MixerBlock.__call__ (mlp_mixer.py:16-31)."""
    y = layer_norm(x, p["LayerNorm_0"]["scale"], p["LayerNorm_0"]["bias"])      # :18
    y = y.transpose(-1, -2)                                                     # :19 '... l d -> ... d l'
    y = ff_block(y, p["FFBlock_0"])                                             # :20-22 over the tokens
    y = y.transpose(-1, -2)                                                     # :23
    x = y + x                                                                   # :24
    y = layer_norm(x, p["LayerNorm_1"]["scale"], p["LayerNorm_1"]["bias"])      # :26
    y = ff_block(y, p["FFBlock_1"])                                             # :27-29 over the channels
    return x + y                                                                # :30


def mixer_forward(params, images_nhwc, cfg):
    """MLPMixer.__call__ (mlp_mixer.py:44-60): patch embedding with bias, blocks, LayerNorm, mean over
    the tokens, Dense head."""
    ph, pw = cfg["patch_shape"]
    pe = params["PatchEmbedBlock_0"]["Dense_0"]
    x = patch_embed(images_nhwc, pe["kernel"], ph, pw, bias=pe["bias"])         # :46-49
    for i in range(cfg["num_layers"]):
        x = mixer_block(x, params[f"MixerBlock_{i}"])                           # :51-55
    x = layer_norm(x, params["LayerNorm_0"]["scale"], params["LayerNorm_0"]["bias"])   # :57
    x = x.mean(dim=1)                                                           # :58
    return x @ params["Dense_0"]["kernel"] + params["Dense_0"]["bias"]          # :59


def mixer_config(name):
    """create_model.py:184-213 (the unreachable second 'mixer_s_patch32' entry is B/16)."""
    cfgs = {"mixer_s_patch32": (8, 512, 32), "mixer_s_patch16": (8, 512, 16),
            "mixer_b_patch32": (12, 768, 32), "mixer_b_patch16": (12, 768, 16),
            "mixer_l_patch32": (24, 1024, 32), "mixer_l_patch16": (32, 1024, 16)}
    layers, dim, p = cfgs[name]
    return dict(num_layers=layers, embed_dim=dim, patch_shape=(p, p), tokens_expand_ratio=0.5,
                channels_expand_ratio=4, num_classes=1000)


def synthetic_mixer_params(cfg, img_size=224, seed=42, dtype=torch.float32):
    """SURVEY 8d synthetic init with the Flax auto-names of mlp_mixer.py."""
    g = torch.Generator().manual_seed(seed)

    def normal(shape, std):
        return (torch.randn(shape, generator=g, dtype=torch.float64) * std).to(dtype)

    def dense(k, n):
        return {"kernel": normal((k, n), 1.0 / math.sqrt(k)), "bias": normal((n,), 0.02)}

    def ln(c):
        return {"scale": 1.0 + normal((c,), 0.02), "bias": normal((c,), 0.02)}
    ph, pw = cfg["patch_shape"]
    C = cfg["embed_dim"]
    L = (img_size // ph) * (img_size // pw)
    ht = max(1, int(cfg["tokens_expand_ratio"] * L))        # ff.py:22-23 on the token axis
    hc = max(1, int(cfg["channels_expand_ratio"] * C))
    params = {"PatchEmbedBlock_0": {"Dense_0": dense(ph * pw * 3, C)}}
    for i in range(cfg["num_layers"]):
        params[f"MixerBlock_{i}"] = {
            "LayerNorm_0": ln(C), "FFBlock_0": {"Dense_0": dense(L, ht), "Dense_1": dense(ht, L)},
            "LayerNorm_1": ln(C), "FFBlock_1": {"Dense_0": dense(C, hc), "Dense_1": dense(hc, C)}}
    params["LayerNorm_0"] = ln(C)
    nc = cfg["num_classes"]
    params["Dense_0"] = {"kernel": normal((C, nc), 0.02), "bias": normal((nc,), 0.02)}
    return params


def mixer_loss_and_grads(params, images_hwcn, labels, cfg, label_smoothing=0.0):
    p = tree_map(lambda t: t.detach().clone().requires_grad_(True), params)
    logits = mixer_forward(p, hwcn_to_nhwc(images_hwcn), cfg)
    loss = loss_fn(logits, labels, cfg["num_classes"], label_smoothing)
    loss.backward()
    grads = tree_map(lambda t: t.grad if t.grad is not None else torch.zeros_like(t), p)
    return loss.detach(), logits.detach(), grads


# ----------------------------------------------------------------------------- TNT
def pixel_embed(images_nhwc, p, patch_shape, sub_shape):
    """PixelEmbedBlock.__call__ (tnt.py:16-30): 'b (h p1) (w p2) c -> (b h w) p1 p2 c', then
    'n (p1 t1) (p2 t2) c -> n (p1 p2) (c t1 t2)', then Dense(+bias)."""
    b, H, W, c = images_nhwc.shape
    ph, pw = patch_shape
    t1, t2 = sub_shape
    gh, gw = H // ph, W // pw
    x = images_nhwc.reshape(b, gh, ph, gw, pw, c).permute(0, 1, 3, 2, 4, 5).reshape(b * gh * gw, ph, pw, c)
    q1, q2 = ph // t1, pw // t2
    x = x.reshape(-1, q1, t1, q2, t2, c).permute(0, 1, 3, 5, 2, 4).reshape(-1, q1 * q2, c * t1 * t2)
    return x @ p["Dense_0"]["kernel"] + p["Dense_0"]["bias"]


def tnt_encoder_block(patch_inputs, pixel_inputs, p, inner_heads, outer_heads):
    """tnt.py EncoderBlock.__call__ (tnt.py:59-87)."""
    from .vit_oracle import attention
    ln = lambda x, q: layer_norm(x, q["scale"], q["bias"])   # noqa: E731
    inner_x = ln(pixel_inputs, p["LayerNorm_0"])                                          # :61
    inner_x = attention(inner_x, inner_x, p["SelfAttentionBlock_0"], inner_heads)         # :62-66
    inner_x = inner_x + pixel_inputs                                                      # :67
    inner_y = ff_block(ln(inner_x, p["LayerNorm_1"]), p["FFBlock_0"])                     # :68-71
    inner_output = inner_x + inner_y                                                      # :72
    # Inner2OuterBlock (tnt.py:36-45)
    b = patch_inputs.shape[0]
    d2 = p["Inner2OuterBlock_0"]["Dense_0"]
    x = inner_output.reshape(inner_output.shape[0], -1)                                   # :40
    x = x @ d2["kernel"] + d2["bias"]                                                     # :41
    x = x.reshape(b, -1, x.shape[-1])                                                     # :42
    x = torch.nn.functional.pad(x, (0, 0, 1, 0))                                          # :43
    outer_x = x + patch_inputs                                                            # :44
    outer_x = ln(outer_x, p["LayerNorm_2"])                                               # :75
    outer_x = attention(outer_x, outer_x, p["SelfAttentionBlock_1"], outer_heads)         # :76-80
    outer_x = outer_x + patch_inputs                                                      # :81
    outer_y = ff_block(ln(outer_x, p["LayerNorm_3"]), p["FFBlock_1"])                     # :82-85
    return outer_x + outer_y, inner_output                                                # :86-87


def tnt_forward(params, images_nhwc, cfg):
    """TNT.__call__ (tnt.py:133-178)."""
    pixel = pixel_embed(images_nhwc, params["PixelEmbedBlock_0"], cfg["patch_shape"],
                        cfg["transformed_patch_shape"])                                   # :135-139
    pe = params["PatchEmbedBlock_0"]["Dense_0"]
    patch = patch_embed(images_nhwc, pe["kernel"], *cfg["patch_shape"], bias=pe["bias"])  # :141-144
    b = patch.shape[0]
    patch = torch.cat([params["cls"].expand(b, -1, -1), patch], dim=1)                    # :146-153
    pixel = pixel + params["AddAbsPosEmbed_0"]["pos_embed"]                               # :155
    patch = patch + params["AddAbsPosEmbed_1"]["pos_embed"]                               # :156
    for i in range(cfg["num_layers"]):
        patch, pixel = tnt_encoder_block(patch, pixel, params["Encoder_0"][f"EncoderBlock_{i}"],
                                         cfg["inner_num_heads"], cfg["outer_num_heads"])
    cls = patch[:, 0]                                                                     # :172
    return cls @ params["Dense_0"]["kernel"] + params["Dense_0"]["bias"]                  # :173-177


def tnt_config(name):
    """create_model.py:50-63."""
    cfgs = {"tnt_s_patch16": (4, 10, 40, 640), "tnt_b_patch16": (4, 6, 24, 384)}
    ih, oh, idim, odim = cfgs[name]
    return dict(num_layers=12, inner_num_heads=ih, outer_num_heads=oh, inner_embed_dim=idim,
                outer_embed_dim=odim, patch_shape=(16, 16), transformed_patch_shape=(4, 4),
                inner_expand_ratio=4, outer_expand_ratio=4, num_classes=1000)


def synthetic_tnt_params(cfg, img_size=224, seed=42, dtype=torch.float32):
    """SURVEY 8d synthetic init with the Flax auto-names of tnt.py."""
    g = torch.Generator().manual_seed(seed)

    def normal(shape, std):
        return (torch.randn(shape, generator=g, dtype=torch.float64) * std).to(dtype)

    def dense(k, n):
        return {"kernel": normal((k, n), 1.0 / math.sqrt(k)), "bias": normal((n,), 0.02)}

    def ln(c):
        return {"scale": 1.0 + normal((c,), 0.02), "bias": normal((c,), 0.02)}

    def attn(c, h):
        d = c // h
        return {"queries": {"kernel": normal((c, h, d), 1.0 / math.sqrt(c))},
                "keys": {"kernel": normal((c, h, d), 1.0 / math.sqrt(c))},
                "values": {"kernel": normal((c, h, d), 1.0 / math.sqrt(c))},
                "DenseGeneral_0": {"kernel": normal((h, d, c), 1.0 / math.sqrt(c))}}
    ph, pw = cfg["patch_shape"]
    t1, t2 = cfg["transformed_patch_shape"]
    ci, co = cfg["inner_embed_dim"], cfg["outer_embed_dim"]
    L = (img_size // ph) * (img_size // pw)
    npix = (ph // t1) * (pw // t2)
    enc = {}
    for i in range(cfg["num_layers"]):
        enc[f"EncoderBlock_{i}"] = {
            "LayerNorm_0": ln(ci), "SelfAttentionBlock_0": attn(ci, cfg["inner_num_heads"]),
            "LayerNorm_1": ln(ci),
            "FFBlock_0": {"Dense_0": dense(ci, int(cfg["inner_expand_ratio"] * ci)),
                          "Dense_1": dense(int(cfg["inner_expand_ratio"] * ci), ci)},
            "Inner2OuterBlock_0": {"Dense_0": dense(npix * ci, co)},
            "LayerNorm_2": ln(co), "SelfAttentionBlock_1": attn(co, cfg["outer_num_heads"]),
            "LayerNorm_3": ln(co),
            "FFBlock_1": {"Dense_0": dense(co, int(cfg["outer_expand_ratio"] * co)),
                          "Dense_1": dense(int(cfg["outer_expand_ratio"] * co), co)}}
    nc = cfg["num_classes"]
    return {"PixelEmbedBlock_0": {"Dense_0": dense(3 * t1 * t2, ci)},
            "PatchEmbedBlock_0": {"Dense_0": dense(ph * pw * 3, co)},
            "cls": normal((1, 1, co), 0.02),
            "AddAbsPosEmbed_0": {"pos_embed": normal((1, npix, ci), 0.02)},
            "AddAbsPosEmbed_1": {"pos_embed": normal((1, L + 1, co), 0.02)},
            "Encoder_0": enc,
            "Dense_0": {"kernel": normal((co, nc), 0.02), "bias": normal((nc,), 0.02)}}


def tnt_loss_and_grads(params, images_hwcn, labels, cfg, label_smoothing=0.0):
    p = tree_map(lambda t: t.detach().clone().requires_grad_(True), params)
    logits = tnt_forward(p, hwcn_to_nhwc(images_hwcn), cfg)
    loss = loss_fn(logits, labels, cfg["num_classes"], label_smoothing)
    loss.backward()
    grads = tree_map(lambda t: t.grad if t.grad is not None else torch.zeros_like(t), p)
    return loss.detach(), logits.detach(), grads


# ----------------------------------------------------------------------------- CeiT
def image_to_token(images_nhwc, p, cfg, eps=1e-5):
    """Image2TokenBlock.__call__ (image_to_token.py:22-48): Conv k x k / stride s with explicit
    padding = the patch shape, no bias -> BatchNorm (batch statistics) -> max_pool (flax default
    padding 'VALID') -> 'b (h ph) (w pw) c -> b (h w) (ph pw c)' -> Dense without bias."""
    ph, pw = cfg["patch_shape"]
    assert ph == pw
    y = conv2d_nobias(images_nhwc, p["Conv_0"]["kernel"], cfg["conv_stride"], ph)         # :24-29
    y = batch_norm_train(y, p["BatchNorm_0"], eps)                                        # :30-33
    y = torch.nn.functional.max_pool2d(y.permute(0, 3, 1, 2), cfg["pool_window_size"],
                                       cfg["pool_stride"]).permute(0, 2, 3, 1)            # :34-38
    return _rb(patch_embed(_rb(y), _rb(p["Dense_0"]["kernel"]), ph, pw))                  # :39-48


def leff_block(x, p, kernel_size, eps=1e-5):
    """LeFFBlock.__call__ (leff.py:20-63), expand_ratio given: the class token bypasses the block;
    the patch tokens go Dense -> BatchNorm -> gelu -> (on the sqrt(l) x sqrt(l) grid) Conv k x k
    'SAME' with bias -> BatchNorm -> gelu -> Dense -> BatchNorm -> gelu."""
    from .vit_oracle import gelu_tanh
    cls, tokens = x[:, 0], x[:, 1:]                                                       # :22
    b, l, c = tokens.shape
    act = lambda t: _rb(gelu_tanh(t))   # noqa: E731
    h = _rb(_rb(tokens) @ _rb(p["Dense_0"]["kernel"]) + p["Dense_0"]["bias"])             # :41
    h = act(batch_norm_train(h, p["BatchNorm_0"], eps))                                   # :42-43
    side = int(math.sqrt(l))                                                              # :45
    h = h.reshape(b, side, side, -1)                                                      # :46
    h = _rb(conv2d_same(_rb(h), _rb(p["Conv_0"]["kernel"]), p["Conv_0"]["bias"], 1))      # :48-53
    h = act(batch_norm_train(h, p["BatchNorm_1"], eps))                                   # :54-55
    h = h.reshape(b, l, -1)                                                               # :57
    h = _rb(_rb(h) @ _rb(p["Dense_1"]["kernel"]) + p["Dense_1"]["bias"])                  # :59
    h = act(batch_norm_train(h, p["BatchNorm_2"], eps))                                   # :60-61
    return torch.cat([cls[:, None], h], dim=1)                                            # :63-65


def _ceit_attention(xq, xkv, p, num_heads):
    """attention.py:21-67 through vit_oracle.attention; under EMULATE_BF16 the same arithmetic with
    round-to-bf16 at the tensors a bf16 pipeline stores (operand, weights, q / k / v, P, context)."""
    if not EMULATE_BF16[0]:
        return attention(xq, xkv, p, num_heads)
    wq, wk, wv = (_rb(p[n]["kernel"]) for n in ("queries", "keys", "values"))
    xq, xkv = _rb(xq), _rb(xkv)
    q = _rb(torch.einsum("bqc,chd->bqhd", xq, wq))
    k = _rb(torch.einsum("bkc,chd->bkhd", xkv, wk))
    v = _rb(torch.einsum("bkc,chd->bkhd", xkv, wv))
    w = softmax_last(torch.einsum("bqhd,bkhd->bhqk", q / math.sqrt(wq.shape[-1]), k))
    o = _rb(torch.einsum("bhqk,bkhd->bqhd", _rb(w), v))
    return torch.einsum("bqhd,hdc->bqc", o, _rb(p["DenseGeneral_0"]["kernel"]))


def ceit_encoder_block(x, p, num_heads, kernel_size):
    """ceit.py EncoderBlock.__call__ (ceit.py:28-44), post-LN."""
    ln = lambda t, q: layer_norm(t, q["scale"], q["bias"])   # noqa: E731
    y = _ceit_attention(x, x, p["SelfAttentionBlock_0"], num_heads)                       # :29-31
    y = y + x                                                                             # :32
    y = ln(y, p["LayerNorm_0"])                                                           # :33
    z = leff_block(y, p["LeFFBlock_0"], kernel_size)                                      # :35-40
    return ln(y + z, p["LayerNorm_1"])                                                    # :41-43


def ceit_forward(params, images_nhwc, cfg):
    """CeiT.__call__ (ceit.py:114-154) + Encoder.__call__ (ceit.py:57-72) + LCSelfAttentionBlock
    (ceit.py:11-16)."""
    x = image_to_token(images_nhwc, params["Image2TokenBlock_0"], cfg)                    # :117-126
    b = x.shape[0]
    x = torch.cat([params["cls"].expand(b, -1, -1), x], dim=1)                            # :128-133
    enc = params["Encoder_0"]
    cls_tokens = []
    for i in range(cfg["num_layers"]):                                                    # :66-70
        x = ceit_encoder_block(x, enc[f"EncoderBlock_{i}"], cfg["num_heads"], cfg["leff_kernel_size"])
        cls_tokens.append(x[:, :1])
    t = torch.cat(cls_tokens, dim=1)                                                      # :72
    t = _rb(t)
    t = _ceit_attention(t[:, -1:], t, params["LCSelfAttentionBlock_0"], cfg["num_heads"])  # :14-16,143-147
    cls = _rb(t[:, -1])                                                                   # :148
    return cls @ _rb(params["Dense_0"]["kernel"]) + params["Dense_0"]["bias"]             # :149-153


def ceit_config(name):
    """create_model.py:64-78 + the CeiT dataclass defaults (ceit.py:100-112)."""
    heads, dim = {"ceit_t": (3, 192), "ceit_s": (6, 384), "ceit_b": (12, 768)}[name]
    return dict(num_layers=12, num_heads=heads, embed_dim=dim, patch_shape=(4, 4), num_ch=32,
                conv_kernel_size=7, conv_stride=2, pool_window_size=3, pool_stride=2, expand_ratio=4,
                leff_kernel_size=3, num_classes=1000)


def synthetic_ceit_params(cfg, seed=42, dtype=torch.float32):
    """SURVEY 8d synthetic init with the Flax auto-names of ceit.py / leff.py / image_to_token.py."""
    g = torch.Generator().manual_seed(seed)

    def normal(shape, std):
        return (torch.randn(shape, generator=g, dtype=torch.float64) * std).to(dtype)

    def dense(k, n, bias=True):
        d = {"kernel": normal((k, n), 1.0 / math.sqrt(k))}
        if bias:
            d["bias"] = normal((n,), 0.02)
        return d

    def bn(c):
        return {"scale": 1.0 + normal((c,), 0.1), "bias": normal((c,), 0.1)}

    def ln(c):
        return {"scale": 1.0 + normal((c,), 0.02), "bias": normal((c,), 0.02)}

    def attn(c, h):
        d = c // h
        return {"queries": {"kernel": normal((c, h, d), 1.0 / math.sqrt(c))},
                "keys": {"kernel": normal((c, h, d), 1.0 / math.sqrt(c))},
                "values": {"kernel": normal((c, h, d), 1.0 / math.sqrt(c))},
                "DenseGeneral_0": {"kernel": normal((h, d, c), 1.0 / math.sqrt(c))}}
    C, H = cfg["embed_dim"], cfg["num_heads"]
    ph, pw = cfg["patch_shape"]
    k, nch = cfg["conv_kernel_size"], cfg["num_ch"]
    hid = max(1, int(cfg["expand_ratio"] * C))
    kl = cfg["leff_kernel_size"]
    enc = {}
    for i in range(cfg["num_layers"]):
        enc[f"EncoderBlock_{i}"] = {
            "SelfAttentionBlock_0": attn(C, H), "LayerNorm_0": ln(C),
            "LeFFBlock_0": {"Dense_0": dense(C, hid), "BatchNorm_0": bn(hid),
                            "Conv_0": {"kernel": normal((kl, kl, hid, hid), 1.0 / math.sqrt(kl * kl * hid)),
                                       "bias": normal((hid,), 0.02)},
                            "BatchNorm_1": bn(hid), "Dense_1": dense(hid, C), "BatchNorm_2": bn(C)},
            "LayerNorm_1": ln(C)}
    nc = cfg["num_classes"]
    return {"Image2TokenBlock_0": {"Conv_0": {"kernel": normal((k, k, 3, nch), 1.0 / math.sqrt(k * k * 3))},
                                   "BatchNorm_0": bn(nch), "Dense_0": dense(ph * pw * nch, C, bias=False)},
            "cls": normal((1, 1, C), 0.02),
            "Encoder_0": enc,
            "LCSelfAttentionBlock_0": attn(C, H),
            "Dense_0": {"kernel": normal((C, nc), 0.02), "bias": normal((nc,), 0.02)}}


def ceit_loss_and_grads(params, images_hwcn, labels, cfg, label_smoothing=0.0, emulate_bf16=False):
    p = tree_map(lambda t: t.detach().clone().requires_grad_(True), params)
    EMULATE_BF16[0] = bool(emulate_bf16)
    try:
        logits = ceit_forward(p, hwcn_to_nhwc(images_hwcn), cfg)
    finally:
        EMULATE_BF16[0] = False
    loss = loss_fn(logits, labels, cfg["num_classes"], label_smoothing)
    loss.backward()
    grads = tree_map(lambda t: t.grad if t.grad is not None else torch.zeros_like(t), p)
    return loss.detach(), logits.detach(), grads
