"""ORACLE (test infrastructure, NOT product code) -- fp32 CPU restatement of the reference hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  The product path (self-attention-experiments-vision_b200/) never does.

PARITY UNPINNED: the reference ships no golden vectors (its tests assert logits *shape* only,
models/vit_test.py:26) and its arithmetic lives in third-party packages that are not installed
here and cannot be (jax~=0.2.13 / jaxlib~=0.1.66, flax + optax at unpinned git HEAD,
requirements.txt:7-8,16-18).  This file restates the published semantics of those library calls at
the reference's own call sites; it is validated by (i) float64 finite differences of its own
backward, (ii) an independent torch.nn.functional cross-check, (iii) the reference's implied
known-answer test (zero-init head => logits exactly 0), (iv) parameter-tree names/shapes.

Every function cites the reference file:line (under /root/reference) it follows.  Arithmetic is
explicit torch CPU tensor math in fp32 (or fp64 when the inputs are); gradients come from torch
autograd over these explicit formulas.
"""
import math

import torch


# ----------------------------------------------------------------------------- library semantics
def gelu_tanh(x):
    """jax.nn.gelu default approximate=True, used as FFBlock.activation_fn (ff.py:12,25)."""
    return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * x ** 3)))


def layer_norm(x, scale, bias, eps=1e-6):
    """flax.linen.LayerNorm defaults (vit.py:19,26,57): eps 1e-6, var = E[x^2] - E[x]^2."""
    mean = x.mean(-1, keepdim=True)
    mean2 = (x * x).mean(-1, keepdim=True)
    var = torch.clamp(mean2 - mean * mean, min=0.0)
    return (x - mean) * torch.rsqrt(var + eps) * scale + bias


def softmax_last(x):
    """flax.linen.softmax over the last (key) axis (attention.py:48): exp(x - max) / sum."""
    e = torch.exp(x - x.max(-1, keepdim=True).values)
    return e / e.sum(-1, keepdim=True)


# ----------------------------------------------------------------------------- layers
def patch_embed(images_nhwc, kernel, ph, pw, bias=None):
    """PatchEmbedBlock.__call__ (patch_embed.py:16-26): 'b (h ph) (w pw) c -> b (h w) (ph pw c)'
    then Dense(embed_dim, use_bias=False)."""
    b, H, W, c = images_nhwc.shape
    gh, gw = H // ph, W // pw
    x = images_nhwc.reshape(b, gh, ph, gw, pw, c).permute(0, 1, 3, 2, 4, 5)
    x = x.reshape(b, gh * gw, ph * pw * c)
    y = x @ kernel
    return y + bias if bias is not None else y


def attention(inputs_q, inputs_kv, p, num_heads, talking=None):
    """AttentionBlock.__call__ (attention.py:21-67).  p: {'queries','keys','values':{'kernel'
    [C,H,d]}, 'DenseGeneral_0': {'kernel' [H,d,C]}}; use_bias=False (default, attention.py:17);
    dropout rates 0 => identity (attention.py:54-55,65-66)."""
    wq, wk, wv = p["queries"]["kernel"], p["keys"]["kernel"], p["values"]["kernel"]
    head_ch = wq.shape[-1]
    q = torch.einsum("bqc,chd->bqhd", inputs_q, wq)            # attention.py:35
    k = torch.einsum("bkc,chd->bkhd", inputs_kv, wk)           # attention.py:36
    v = torch.einsum("bkc,chd->bkhd", inputs_kv, wv)           # attention.py:37
    q = q / math.sqrt(head_ch)                                  # attention.py:39
    w = torch.einsum("bqhd,bkhd->bhqk", q, k)                  # attention.py:41-42
    if talking is not None:                                     # attention.py:44-46
        w = torch.einsum("hi,bhqk->biqk", talking[0], w)       # talking_heads.py:13
    w = softmax_last(w)                                         # attention.py:48
    if talking is not None:                                     # attention.py:50-52
        w = torch.einsum("hi,bhqk->biqk", talking[1], w)
    o = torch.einsum("bhqk,bkhd->bqhd", w, v)                  # attention.py:57-58
    return torch.einsum("bqhd,hdc->bqc", o, p["DenseGeneral_0"]["kernel"])  # attention.py:60-63


def ff_block(x, p):
    """FFBlock.__call__ (ff.py:15-34): Dense_0(+bias) -> gelu -> Dense_1(+bias)."""
    h = x @ p["Dense_0"]["kernel"] + p["Dense_0"]["bias"]      # ff.py:24
    h = gelu_tanh(h)                                            # ff.py:25
    return h @ p["Dense_1"]["kernel"] + p["Dense_1"]["bias"]   # ff.py:27


def encoder_block(x, p, num_heads):
    """vit.py EncoderBlock.__call__ (vit.py:17-32), pre-LN."""
    y = layer_norm(x, p["LayerNorm_0"]["scale"], p["LayerNorm_0"]["bias"])     # vit.py:19
    y = attention(y, y, p["SelfAttentionBlock_0"], num_heads)                  # vit.py:20-23
    x = y + x                                                                  # vit.py:24
    y = layer_norm(x, p["LayerNorm_1"]["scale"], p["LayerNorm_1"]["bias"])     # vit.py:26
    y = ff_block(y, p["FFBlock_0"])                                            # vit.py:27-30
    return x + y                                                               # vit.py:31


def vit_forward(params, images_nhwc, num_layers, num_heads, patch_shape):
    """ViT.__call__ (vit.py:73-99) + Encoder.__call__ (vit.py:44-58)."""
    ph, pw = patch_shape
    x = patch_embed(images_nhwc, params["PatchEmbedBlock_0"]["Dense_0"]["kernel"], ph, pw)  # :77
    b = x.shape[0]
    cls = params["cls"].expand(b, -1, -1)                                      # vit.py:81-84
    x = torch.cat([cls, x], dim=1)                                             # vit.py:85
    enc = params["Encoder_0"]
    x = x + enc["AddAbsPosEmbed_0"]["pos_embed"]                               # position_embed.py:54-56
    for i in range(num_layers):                                                # vit.py:49-55
        x = encoder_block(x, enc[f"EncoderBlock_{i}"], num_heads)
    x = layer_norm(x, enc["LayerNorm_0"]["scale"], enc["LayerNorm_0"]["bias"])  # vit.py:57
    cls_tok = x[:, 0]                                                          # vit.py:95
    return cls_tok @ params["Dense_0"]["kernel"] + params["Dense_0"]["bias"]   # vit.py:96-98


# ----------------------------------------------------------------------------- train step pieces
def loss_fn(logits, labels, num_classes, label_smoothing=0.0, device_count=1, mix_labels=None, ratio=None):
    """train.py:83-91: one_hot -> (mixup: ratio*y + (1-ratio)*y1, train.py:84-87) ->
    optax.smooth_labels -> optax.softmax_cross_entropy -> mean (-> / device_count, kept as a
    switch; SURVEY D7 uses the plain mean)."""
    y = torch.nn.functional.one_hot(labels.long(), num_classes).to(logits.dtype)  # train.py:83
    if mix_labels is not None:                                                 # train.py:84-87
        y1 = torch.nn.functional.one_hot(mix_labels.long(), num_classes).to(logits.dtype)
        y = ratio[:, None] * y + (1.0 - ratio[:, None]) * y1
    y = (1.0 - label_smoothing) * y + label_smoothing / num_classes            # train.py:88
    logp = logits - torch.logsumexp(logits, dim=-1, keepdim=True)
    loss = -(y * logp).sum(-1).mean()                                          # train.py:89-90
    return loss / device_count                                                 # train.py:91


def warmup_cosine_decay_schedule(init_value, peak_value, warmup_steps, decay_steps, end_value=0.0):
    """optax.warmup_cosine_decay_schedule as called in train.py:215-220, restated from optax's
    published source: join_schedules([linear_schedule(init, peak, warmup),
    cosine_decay_schedule(peak, decay - warmup, alpha=end/peak)], [warmup])."""
    def linear(count):
        count = min(max(count, 0), warmup_steps)
        frac = 1.0 - count / warmup_steps
        return (init_value - peak_value) * frac + peak_value

    def cosine(count):
        D = decay_steps - warmup_steps
        count = min(count, D)
        cosine_decay = 0.5 * (1.0 + math.cos(math.pi * count / D))
        alpha = end_value / peak_value
        return peak_value * ((1.0 - alpha) * cosine_decay + alpha)

    def schedule(count):
        return linear(count) if count < warmup_steps else cosine(count - warmup_steps)
    return schedule


def hwcn_to_nhwc(images_hwcn):
    """train.py:80: rearrange(batch['images'], 'H W C N -> N H W C')."""
    return images_hwcn.permute(3, 0, 1, 2).contiguous()


def vit_loss_and_grads(params, images_hwcn, labels, cfg, label_smoothing=0.0):
    """train.py:79-95: value_and_grad of the loss w.r.t. every parameter leaf."""
    leaves = []

    def req(t):
        t = t.detach().clone().requires_grad_(True)
        leaves.append(t)
        return t

    p = tree_map(req, params)
    logits = vit_forward(p, hwcn_to_nhwc(images_hwcn), cfg["num_layers"], cfg["num_heads"],
                         cfg["patch_shape"])
    loss = loss_fn(logits, labels, cfg["num_classes"], label_smoothing)
    loss.backward()
    grads = tree_map(lambda t: t.grad if t.grad is not None else torch.zeros_like(t), p)
    return loss.detach(), logits.detach(), grads


def adam_chain_step(p, g, m, v, t, lr, max_norm=None, b1=0.9, b2=0.999, eps=1e-8, wd=0.0):
    """train.py:25-27 / simple_train.py:25-27 on flat tensors: clip_by_global_norm ->
    scale_by_adam -> additive_weight_decay -> scale(-lr) (descent) -> apply_updates."""
    if max_norm is not None and max_norm > 0:
        gn = torch.sqrt((g * g).sum())
        if gn >= max_norm:
            g = g * (max_norm / gn)
    # optax.scale_by_adam arithmetic is float32 throughout: (1 - decay) is a Python double rounded to
    # fp32 by weak typing, the bias correction 1 - decay**count is evaluated in fp32
    f32 = torch.float32
    omb1, omb2 = torch.tensor(1 - b1, dtype=f32), torch.tensor(1 - b2, dtype=f32)
    b1t, b2t = torch.tensor(b1, dtype=f32), torch.tensor(b2, dtype=f32)
    m = omb1 * g + b1t * m
    v = omb2 * (g * g) + b2t * v
    tt = torch.tensor(float(t), dtype=f32)
    mhat = m / (1 - torch.pow(b1t, tt))
    vhat = v / (1 - torch.pow(b2t, tt))
    upd = mhat / (torch.sqrt(vhat) + eps) + wd * p
    return p - lr * upd, m, v


# ----------------------------------------------------------------------------- params
def tree_map(fn, tree):
    if isinstance(tree, dict):
        return {k: tree_map(fn, v) for k, v in tree.items()}
    return fn(tree)


def tree_leaves(tree, prefix=""):
    if isinstance(tree, dict):
        for k in tree:
            yield from tree_leaves(tree[k], f"{prefix}/{k}" if prefix else k)
    else:
        yield prefix, tree


def vit_config(name):
    cfgs = {
        # DeiT-Ti / DeiT-S are ViT(num_layers=12, heads 3/6, embed 192/384, patch 16)
        "deit_ti_patch16": dict(num_layers=12, num_heads=3, embed_dim=192, patch_shape=(16, 16)),
        "deit_s_patch16": dict(num_layers=12, num_heads=6, embed_dim=384, patch_shape=(16, 16)),
        "vit_b_patch16": dict(num_layers=12, num_heads=12, embed_dim=768, patch_shape=(16, 16)),
        "vit_b_patch32": dict(num_layers=12, num_heads=12, embed_dim=768, patch_shape=(32, 32)),
    }
    c = dict(cfgs[name])
    c["num_classes"] = 1000
    c["expand_ratio"] = 4
    return c


def synthetic_vit_params(cfg, img_size=224, seed=42, default_init=False, dtype=torch.float32):
    """Parameter tree with the Flax auto-names of vit.py.  Non-default synthetic init
    (SURVEY 8d): kernels N(0,1/fan_in), biases N(0,0.02), LN scale 1+N(0,0.02), cls/pos N(0,0.02),
    head kernel N(0,0.02).  default_init=True reproduces the reference initialisers' *zeros*
    (cls zeros vit.py:82, head kernel zeros vit.py:98, biases zeros) for the zero-logits KAT."""
    g = torch.Generator().manual_seed(seed)
    C, H = cfg["embed_dim"], cfg["num_heads"]
    d = C // H
    ph, pw = cfg["patch_shape"]
    n_tok = (img_size // ph) * (img_size // pw)
    hid = int(cfg["expand_ratio"] * C)

    def normal(shape, std):
        return (torch.randn(shape, generator=g, dtype=torch.float64) * std).to(dtype)

    def kern(shape, fan_in):
        return normal(shape, 1.0 / math.sqrt(fan_in))

    def bias(n):
        return torch.zeros(n, dtype=dtype) if default_init else normal((n,), 0.02)

    def ln():
        if default_init:
            return {"scale": torch.ones(C, dtype=dtype), "bias": torch.zeros(C, dtype=dtype)}
        return {"scale": 1.0 + normal((C,), 0.02), "bias": normal((C,), 0.02)}

    enc = {"AddAbsPosEmbed_0": {"pos_embed": normal((1, n_tok + 1, C), 0.02)}}
    for i in range(cfg["num_layers"]):
        enc[f"EncoderBlock_{i}"] = {
            "LayerNorm_0": ln(),
            "SelfAttentionBlock_0": {
                "queries": {"kernel": kern((C, H, d), C)},
                "keys": {"kernel": kern((C, H, d), C)},
                "values": {"kernel": kern((C, H, d), C)},
                "DenseGeneral_0": {"kernel": kern((H, d, C), C)},
            },
            "LayerNorm_1": ln(),
            "FFBlock_0": {
                "Dense_0": {"kernel": kern((C, hid), C), "bias": bias(hid)},
                "Dense_1": {"kernel": kern((hid, C), hid), "bias": bias(C)},
            },
        }
    enc["LayerNorm_0"] = ln()
    nc = cfg["num_classes"]
    return {
        "PatchEmbedBlock_0": {"Dense_0": {"kernel": kern((ph * pw * 3, C), ph * pw * 3)}},
        "cls": torch.zeros(1, 1, C, dtype=dtype) if default_init else normal((1, 1, C), 0.02),
        "Encoder_0": enc,
        "Dense_0": {"kernel": torch.zeros(C, nc, dtype=dtype) if default_init else normal((C, nc), 0.02),
                    "bias": bias(nc)},
    }


def synthetic_batch(batch, img_size=224, num_classes=1000, seed=0, dtype=torch.float32):
    """images N(0,1) in the reference layout [H,W,3,B] (train.py:80), labels uniform."""
    g = torch.Generator().manual_seed(seed)
    images = torch.randn((img_size, img_size, 3, batch), generator=g, dtype=torch.float32).to(dtype)
    g1 = torch.Generator().manual_seed(seed + 1)
    labels = torch.randint(0, num_classes, (batch,), generator=g1, dtype=torch.int32)
    return images, labels
