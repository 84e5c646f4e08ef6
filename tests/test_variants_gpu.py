"""GPU parity of the attention variants and CaiT against the fp32 CPU oracle
(oracle/variants_oracle.py) on identical synthetic inputs, through the module surface -> C ABI.

  * materialised-logits attention core (sav_bgemm + sav_thsoftmax_*): talking heads (d = 48),
    class attention (Nq = 1), ragged key counts;
  * BoTMHSA with 2-D relative position logits (14 x 14 grid, 4 heads x 128): output, input
    gradient, all parameter gradients;
  * CaiT (talking heads + LayerScale + drop-path + class attention): logits and EVERY parameter
    gradient, with layerscale = 0.1 (so that the branches matter, SURVEY 7 'Tolerance') and with
    the reference's 1e-5 init.

Tolerance: per tensor max|delta| / max|ref| <= 2e-2 (BASELINE.json north_star, bf16 path).
The [H,H] talking-heads transforms are held to TH_TOL = 3e-2 at model level: their gradient is a
16..64-element reduction over every (image, query, key) with heavy cancellation; the row kernel
itself reproduces it to < 1e-4 when fed the same bf16 operands (op-level cases below, measured
on B200), the model-level 1.3e-2 .. 2.3e-2 is the bf16 rounding of q / k / v / dO upstream of it.
"""
import json
import math
import os

import pytest
import torch

from oracle import variants_oracle as V
from oracle import vit_oracle as O

TOL = 2e-2
TH_TOL = 3e-2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rel(a, b):
    a = a.detach().float().cpu()
    b = b.detach().float().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-20))


def _dump(tag, errs):
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, f"parity_{tag}.json"), "w") as f:
        json.dump({"worst": sorted(errs.items(), key=lambda kv: -kv[1])[:12], "n_tensors": len(errs)}, f,
                  indent=1)


def _check(tag, errs, tol=TOL):
    _dump(tag, errs)
    print(tag, "worst:", sorted(errs.items(), key=lambda kv: -kv[1])[:6])
    bad = {k: v for k, v in errs.items() if not v <= tol}
    assert not bad, f"{tag}: {len(bad)} tensors over {tol}: {sorted(bad.items(), key=lambda kv: -kv[1])[:10]}"


# ------------------------------------------------------------------------------ op level
def _attn_ref(q, k, v, scale, t1=None, t2=None):
    s = torch.einsum("bqhd,bkhd->bhqk", q, k) * scale
    if t1 is not None:
        s = torch.einsum("hi,bhqk->biqk", t1, s)
    p = torch.softmax(s, dim=-1)
    if t2 is not None:
        p = torch.einsum("hi,bhqk->biqk", t2, p)
    return torch.einsum("bhqk,bkhd->bqhd", p, v)


@pytest.mark.gpu
@pytest.mark.parametrize("B,H,Nq,Nk,D,talking", [
    (3, 4, 196, 196, 48, True),      # CaiT-XXS self-attention layer
    (2, 8, 196, 196, 48, True),      # CaiT-S
    (2, 6, 50, 50, 48, True),        # ragged (not a tile multiple), 6 heads
    (2, 16, 70, 70, 48, True),       # CaiT-M head count (sliced backward), ragged
    (1, 12, 197, 197, 64, True),     # 12 heads, 197 keys
    (5, 4, 1, 197, 48, False),       # class attention: one query row
    (2, 3, 70, 33, 32, False),       # Nq != Nk, small head dim
    (2, 4, 196, 196, 128, False),    # BoTNet head dim without the relative term
])
def test_materialised_attention_core(B, H, Nq, Nk, D, talking):
    from sav import ops
    g = torch.Generator().manual_seed(B * 1000 + H * 100 + Nq)
    q = torch.randn((B, Nq, H, D), generator=g)
    k = torch.randn((B, Nk, H, D), generator=g)
    v = torch.randn((B, Nk, H, D), generator=g)
    d_o = torch.randn((B, Nq, H, D), generator=g)
    t1 = t2 = None
    if talking:
        t1 = torch.linalg.qr(torch.randn((H, H), generator=g))[0] + 0.1 * torch.randn((H, H), generator=g)
        t2 = torch.linalg.qr(torch.randn((H, H), generator=g))[0] + 0.1 * torch.randn((H, H), generator=g)
    scale = 1.0 / math.sqrt(D)
    # reference on the bf16-rounded operands, fp32 math
    qr, kr, vr = (t.bfloat16().float().requires_grad_(True) for t in (q, k, v))
    tr = [t.clone().requires_grad_(True) for t in (t1, t2)] if talking else [None, None]
    o_ref = _attn_ref(qr, kr, vr, scale, tr[0], tr[1])
    (o_ref * d_o.bfloat16().float()).sum().backward()

    qd, kd, vd, dod = (t.cuda().bfloat16() for t in (q, k, v, d_o))
    m1 = t1.cuda().contiguous() if talking else None
    m2 = t2.cuda().contiguous() if talking else None
    o, S, P = ops.attn_mat_fwd(qd, kd, vd, scale, m1, m2)
    dq, dk, dv = torch.empty_like(qd), torch.empty_like(kd), torch.empty_like(vd)
    dm1 = torch.zeros_like(m1) if talking else None
    dm2 = torch.zeros_like(m2) if talking else None
    ops.attn_mat_bwd(qd, kd, vd, S, P, dod, dq, dk, dv, scale, m1, m2, dm1, dm2)
    torch.cuda.synchronize()
    errs = {"o": _rel(o, o_ref), "dq": _rel(dq, qr.grad), "dk": _rel(dk, kr.grad), "dv": _rel(dv, vr.grad)}
    if talking:
        errs["dmix_pre"] = _rel(dm1, tr[0].grad)
        errs["dmix_post"] = _rel(dm2, tr[1].grad)
    _check(f"attn_mat_B{B}H{H}q{Nq}k{Nk}d{D}", errs, tol=1.5e-2)


# ------------------------------------------------------------------------------ BoTNet MHSA
class _Cotangent:
    """Stand-in loss for value_and_grad: seeds the output gradient with a fixed cotangent."""

    def __init__(self, out, cot):
        self.out, self.cot = out, cot

    def seed(self):
        self.out.g = self.cot


@pytest.mark.gpu
def test_bot_mhsa_relative_logits():
    import sav
    from models import BoTMHSA
    from sav.tape import Var
    B, Hh, W, C, h, d = 4, 14, 14, 512, 4, 128
    g = torch.Generator().manual_seed(7)
    x = torch.randn((B, Hh, W, C), generator=g).bfloat16().float()
    cot = torch.randn((B, Hh, W, h * d), generator=g).bfloat16().float()
    params = V.synthetic_bot_params(C, h, d, Hh, W, seed=42)
    pr = O.tree_map(lambda t: t.detach().clone().requires_grad_(True), params)
    xr = x.clone().requires_grad_(True)
    out_ref = V.bot_mhsa(xr, pr, h, d)
    (out_ref * cot).sum().backward()

    model = BoTMHSA(num_heads=h, head_ch=d)
    xin = Var(x.cuda().bfloat16())
    variables = sav.load_params(model, params, xin)
    holder = {}

    def fn(p, xv):
        out = model.apply({"params": p}, xv)
        holder["out"] = out
        return _Cotangent(out, cot.cuda().bfloat16())
    xv = Var(x.cuda().bfloat16())
    _, grads = sav.value_and_grad(fn)(variables["params"], xv)
    torch.cuda.synchronize()
    errs = {"out": _rel(holder["out"].v, out_ref), "dx": _rel(xv.g, xr.grad)}
    ours = dict(O.tree_leaves(grads))
    for path, t in O.tree_leaves(pr):
        errs[path] = _rel(ours[path], t.grad)
    _check("bot_mhsa", errs)


# ------------------------------------------------------------------------------ CaiT
def _cait_case(tag, num_layers, heads, dim, batch, img, layerscale, drop_rate, seed=0):
    import sav
    from models import CaiT
    from sav.train import HWCN, make_loss_fn
    cfg = dict(num_layers=num_layers, num_heads=heads, embed_dim=dim, layerscale_eps=1e-5,
               num_layers_token_only=2, patch_shape=(16, 16), num_classes=1000, expand_ratio=4)
    params = V.synthetic_cait_params(cfg, img_size=img, seed=42, layerscale=layerscale)
    images, labels = O.synthetic_batch(batch, img_size=img, seed=seed)
    factors = None
    sd_seed = 1234
    if drop_rate > 0:
        # the module draws its per-sample factors from the device Philox stream keyed by
        # (seed, step, call index) once per StochasticDepthBlock call, in execution order
        # (stochastic_depth.py:17-21); the oracle is fed the same factors
        from sav.rng import drop_path_factors
        factors = [f.cpu() for f in drop_path_factors(sd_seed, 0, 2 * (num_layers + 2), batch,
                                                      1.0 - drop_rate, "cuda")]
        assert any(float(f.min()) == 0.0 for f in factors) and all(float(f.max()) > 1.0 for f in factors)
    loss_ref, logits_ref, grads_ref = V.cait_loss_and_grads(params, images, labels, cfg, 0.1, factors)

    model = CaiT(num_classes=1000, num_layers=num_layers, num_layers_token_only=2, num_heads=heads,
                 embed_dim=dim, patch_shape=(16, 16), stoch_depth_rate=drop_rate, layerscale_eps=1e-5)
    dev_images = images.cuda()
    variables = sav.load_params(model, params, HWCN(dev_images), is_training=False)

    def loss_fn(p, im, lb):
        from sav import functional as F
        logits = model.apply({"params": p}, im, is_training=True, rngs={"stochastic_depth": sd_seed})
        return F.softmax_cross_entropy(logits, lb, 0.1), logits
    (loss, logits), grads = sav.value_and_grad(loss_fn, has_aux=True)(
        variables["params"], HWCN(dev_images), labels.cuda())
    torch.cuda.synchronize()
    assert abs(loss.item() - float(loss_ref)) <= 2e-2 * abs(float(loss_ref)), (loss.item(), float(loss_ref))
    errs = {"logits": _rel(logits.v, logits_ref)}
    ours = dict(O.tree_leaves(grads))
    for path, g_ref in O.tree_leaves(grads_ref):
        errs[path] = _rel(ours[path], g_ref)
    th = {k: v for k, v in errs.items() if k.endswith("talking_heads_transform")}
    rest = {k: v for k, v in errs.items() if k not in th}
    _check(tag, rest)
    _check(tag + "_th", th, tol=TH_TOL)


@pytest.mark.gpu
def test_cait_xxs_layerscale_0p1_all_grads():
    # CaiT-XXS width (4 heads x 48), 4 SA + 2 CA layers, 224^2 (196 tokens), layerscale 0.1
    _cait_case("cait_xxs4_ls0p1", num_layers=4, heads=4, dim=192, batch=8, img=224, layerscale=0.1,
               drop_rate=0.0)


@pytest.mark.gpu
def test_cait_s_width_drop_path_all_grads():
    # CaiT-S width (8 heads x 48), drop-path 0.1 with shared masks, layerscale 0.1
    _cait_case("cait_s3_ls0p1_dp", num_layers=3, heads=8, dim=384, batch=8, img=224, layerscale=0.1,
               drop_rate=0.1)


@pytest.mark.gpu
def test_cait_m_width_16_heads_all_grads():
    # CaiT-M width (16 heads x 48 = 768): the many-head talking-heads path (sliced backward)
    _cait_case("cait_m2_ls0p1", num_layers=2, heads=16, dim=768, batch=4, img=224, layerscale=0.1,
               drop_rate=0.0)


@pytest.mark.gpu
def test_cait_reference_layerscale_init():
    # the reference's LayerScale init 1e-5 (cait_test.py:18): branches nearly vanish in the stream
    _cait_case("cait_xxs4_ls1e-5", num_layers=4, heads=4, dim=192, batch=8, img=224, layerscale=None,
               drop_rate=0.0)


# ------------------------------------------------------------------------------ CvT stage block
@pytest.mark.gpu
@pytest.mark.parametrize("B,L,C,heads", [
    (4, 197, 384, 6),     # CvT stage 3: 197 tokens -> 15 x 15 map, keys on 8 x 8 (odd SAME padding)
    (2, 784, 192, 3),     # stage 2: 28 x 28 queries, 14 x 14 keys
    (1, 3136, 64, 1),     # stage 1: 56 x 56 queries, 28 x 28 keys, one head
])
def test_cvt_stage_block(B, L, C, heads):
    import sav
    from models import StageBlock
    from sav.tape import Var
    g = torch.Generator().manual_seed(L)
    x = torch.randn((B, L, C), generator=g)
    params = V.synthetic_cvt_stage_params(C, heads, seed=42)
    pr = O.tree_map(lambda t: t.detach().clone().requires_grad_(True), params)
    xr = x.clone().requires_grad_(True)
    out_ref = V.cvt_stage_block(xr, pr, heads)
    cot = torch.randn(out_ref.shape, generator=g)
    (out_ref * cot).sum().backward()

    model = StageBlock(num_heads=heads, embed_dim=C)
    variables = sav.load_params(model, params, Var(x.cuda()), is_training=True)
    holder = {}

    class Cot:
        def __init__(self, out):
            self.out = out

        def seed(self):
            self.out.g = cot.cuda()

    def fn(p, xv):
        holder["out"] = model.apply({"params": p}, xv, is_training=True)
        return Cot(holder["out"])
    xv = Var(x.cuda())
    _, grads = sav.value_and_grad(fn)(variables["params"], xv)
    torch.cuda.synchronize()
    errs = {"out": _rel(holder["out"].v, out_ref), "dx": _rel(xv.g, xr.grad)}
    ours = dict(O.tree_leaves(grads))
    for path, t in O.tree_leaves(pr):
        errs[path] = _rel(ours[path], t.grad)
    # a constant added to every key shifts all logits of a query equally: the softmax -- hence the
    # loss -- does not depend on the key projection's BatchNorm bias, its true gradient is 0
    kb = "CvTSelfAttentionBlock_0/ConvProjectionBlock_1/BatchNorm_0/bias"
    ref_kb = dict(O.tree_leaves(pr))[kb].grad
    scale_ref = dict(O.tree_leaves(pr))["CvTSelfAttentionBlock_0/ConvProjectionBlock_2/BatchNorm_0/bias"].grad
    assert float(ref_kb.abs().max()) < 1e-4 * float(scale_ref.abs().max())
    assert float(ours[kb].float().abs().max().cpu()) < 2e-2 * float(scale_ref.abs().max())
    del errs[kb]
    _check(f"cvt_stage_B{B}L{L}C{C}", errs)


# ------------------------------------------------------------------------------ rotary helpers
@pytest.mark.gpu
@pytest.mark.parametrize("shape,axis,dtype", [
    ((197, 6, 64), 0, torch.bfloat16),        # [seq, heads, dim], seq_dim = 0 (the reference default)
    ((3, 50, 4, 48), 1, torch.bfloat16),      # batched [B, N, H, D]
    ((33, 2, 128), 0, torch.float32),
])
def test_rotary_matches_oracle_forward_and_reverse(shape, axis, dtype):
    """position_embed.py:8-45 (intended reading, parity unpinned): forward against the restated
    helpers, reverse mode against autograd of the restatement, FixedPositionalEmbedding table."""
    import sav
    from models.layers import FixedPositionalEmbedding, RotaryPositionalEmbedding
    from sav.tape import Var
    g = torch.Generator().manual_seed(len(shape) * 10 + axis)
    x = torch.randn(shape, generator=g).to(dtype)
    cot = torch.randn(shape, generator=g).to(dtype)
    xr = x.float().clone().requires_grad_(True)
    y_ref = V.apply_rotary_pos_emb(xr, axis)
    (y_ref * cot.float()).sum().backward()

    model = RotaryPositionalEmbedding()
    xv = Var(x.cuda())
    variables = model.init({"params": 0}, xv, seq_dim=axis)
    holder = {}

    def fn(p, v):
        holder["y"] = model.apply({"params": p}, v, seq_dim=axis)
        return _Cotangent(holder["y"], cot.cuda())
    sav.value_and_grad(fn)(variables["params"], xv)
    torch.cuda.synchronize()
    tol = 1e-2 if dtype == torch.bfloat16 else 2e-5     # one bf16 rounding of the output
    assert _rel(holder["y"].v, y_ref) <= tol
    assert _rel(xv.g, xr.grad) <= tol
    emb = FixedPositionalEmbedding().apply({"params": variables["params"]}, xv, seq_dim=axis)
    ref = V.fixed_positional_embedding(shape[axis], shape[-1])
    assert torch.equal(emb.v.cpu(), ref)


# ------------------------------------------------------------------------------ full CvT model
def _small_cvt_cfg():
    return dict(stage_sizes=(1, 1, 2), num_heads=(1, 3, 6), embed_dim=(64, 192, 384),
                embed_kernel_size=(7, 3, 3), embed_strides=(4, 2, 2), num_classes=1000)


@pytest.mark.gpu
def test_cvt_model_logits_and_all_grads():
    """CvT (cvt.py:116-171) end to end: conv token embedding (7x7 stride 4 / 3x3 stride 2 as im2col
    + tcgen05 GEMM) -> LayerNorm into the stream -> StageBlocks -> class token in the last stage ->
    head on the stream's class row.  CvT-13's widths and head counts at reduced depth, 96x96 images
    (576 / 144 / 36+1 tokens: the last stage pads 37 tokens to a 7x7 map, keys on 4x4)."""
    import sav
    from models import CvT
    from sav.train import HWCN, make_loss_fn
    cfg = _small_cvt_cfg()
    params = V.synthetic_cvt_params(cfg, seed=42)
    images, labels = O.synthetic_batch(16, img_size=96, seed=0)
    loss_ref, logits_ref, grads_ref = V.cvt_loss_and_grads(params, images, labels, cfg, 0.1)
    model = CvT(num_classes=1000, stage_sizes=cfg["stage_sizes"], num_heads=cfg["num_heads"],
                embed_dim=cfg["embed_dim"])
    x = HWCN(images.cuda())
    variables = sav.load_params(model, params, x, is_training=True)
    (loss, logits), grads = sav.value_and_grad(make_loss_fn(model, 0.1), has_aux=True)(
        variables["params"], x, labels.cuda())
    torch.cuda.synchronize()
    assert abs(loss.item() - float(loss_ref)) <= 2e-2 * abs(float(loss_ref))
    errs = {"logits": _rel(logits.v, logits_ref)}
    l2 = {}
    ours = dict(O.tree_leaves(grads))
    zero_ref = []
    for path, g_ref in O.tree_leaves(grads_ref):
        # the key projection's BatchNorm bias has an analytically zero gradient (softmax shift
        # invariance, see test_cvt_stage_block): compare it on the scale of its sibling
        if path.endswith("ConvProjectionBlock_1/BatchNorm_0/bias"):
            zero_ref.append(path)
            continue
        errs[path] = _rel(ours[path], g_ref)
        l2[path] = float((ours[path].float().cpu() - g_ref).norm() / g_ref.norm())
    ref = dict(O.tree_leaves(grads_ref))
    zero_ratio = {}
    for path in zero_ref:
        sib = ref[path.replace("ConvProjectionBlock_1", "ConvProjectionBlock_2")]
        zero_ratio[path] = float(ours[path].float().abs().max().cpu()) / float(sib.abs().max())
    print("cvt zero-gradient biases / sibling:", sorted(zero_ratio.items(), key=lambda kv: -kv[1])[:2])
    print("cvt worst max-norm:", sorted(errs.items(), key=lambda kv: -kv[1])[:3])
    print("cvt worst relative L2:", sorted(l2.items(), key=lambda kv: -kv[1])[:4])
    assert max(zero_ratio.values()) < 2e-2, zero_ratio
    # Few tokens per weight gradient (36 .. 576 per image x 16 images) and three more bf16 operand
    # sites per attention than ViT (depthwise-conv output, BatchNorm output, per projection): the
    # edge-shape tolerance of tests/test_vit_parity_gpu.py applies to the max-norm, every tensor must
    # be within 3e-2 in relative L2, the logits within the north-star 2e-2.
    # Run-to-run spread (12 runs of this test on one B200, profiles/r02bg_cvt_parity_spread.txt): worst
    # max-norm 2.44e-2 .. 3.33e-2 (always a BatchNorm / depthwise-conv parameter of the 576-token stage),
    # worst relative L2 2.04e-2 .. 2.69e-2.  The fp32 reductions that arrive in a different order from
    # run to run (split-K dW, BatchNorm sums, the attention backward's dK / dV groups) differ by 1e-7,
    # but a flipped bf16 rounding behind a batch-16 BatchNorm moves a whole channel; the max-norm bound
    # therefore leaves 2 sigma above the observed maximum, the L2 bound stays.
    _check("cvt_small", errs, tol=4e-2)
    assert max(l2.values()) <= 3e-2, sorted(l2.items(), key=lambda kv: -kv[1])[:5]
    assert errs["logits"] <= TOL


@pytest.mark.gpu
def test_cvt_batchnorm_running_averages_and_eval_mode():
    """cvt_attention.py:33-36: nn.BatchNorm(use_running_average=not is_training, momentum=0.9).
    init() creates batch_stats (mean 0, var 1); apply(mutable=['batch_stats']) in training mode
    folds the batch statistics in; eval mode normalises with them."""
    import sav
    from models import StageBlock
    from sav.tape import Var
    B, L, C, heads = 4, 197, 384, 6
    g = torch.Generator().manual_seed(3)
    x = torch.randn((B, L, C), generator=g)
    params = V.synthetic_cvt_stage_params(C, heads, seed=42)
    model = StageBlock(num_heads=heads, embed_dim=C)
    variables = model.init({"params": 0}, Var(x.cuda()), is_training=True)
    variables = {"params": variables["params"].arena.load_tree(params), "batch_stats": variables["batch_stats"]}
    bs = dict(O.tree_leaves(variables["batch_stats"]))
    assert len(bs) == 6 and all(float(v.abs().max()) == 0 for k, v in bs.items() if k.endswith("mean"))
    # training step with mutable batch_stats: ra = 0.9 * ra + 0.1 * batch stat
    out, upd = model.apply(variables, Var(x.cuda()), is_training=True, mutable=["batch_stats"])
    torch.cuda.synchronize()
    S = 15
    padded = torch.cat([x, x.new_zeros(B, S * S - L, C)], dim=1).reshape(B, S, S, C)
    pa = params["CvTSelfAttentionBlock_0"]
    new = dict(O.tree_leaves(upd["batch_stats"]))
    ras = []
    for i, stride in enumerate((1, 2, 2)):
        mean, var = V.conv_projection_batch_stats(padded, pa[f"ConvProjectionBlock_{i}"], stride)
        pre = f"CvTSelfAttentionBlock_0/ConvProjectionBlock_{i}/BatchNorm_0/"
        ra_m, ra_v = 0.1 * mean, 0.9 + 0.1 * var
        # the device statistics are taken over the bf16-rounded conv output's fp32 values
        assert _rel(new[pre + "mean"], ra_m) <= 5e-3 and _rel(new[pre + "var"], ra_v) <= 5e-3
        ras.append((new[pre + "mean"].cpu(), new[pre + "var"].cpu()))
    # eval mode uses exactly those running averages
    out_eval = model.apply({"params": variables["params"], "batch_stats": upd["batch_stats"]},
                           Var(x.cuda()), is_training=False)
    torch.cuda.synchronize()
    q, k, v = (V.conv_projection_eval(padded, pa[f"ConvProjectionBlock_{i}"], s, *ras[i])
               for i, s in enumerate((1, 2, 2)))
    d = C // heads
    q, k, v = (t.reshape(B, -1, heads, d) for t in (q, k, v))
    w = V.softmax_last(torch.einsum("bqhd,bkhd->bhqk", q / math.sqrt(d), k))
    a = torch.einsum("bqhd,hdc->bqc", torch.einsum("bhqk,bkhd->bqhd", w, v), pa["DenseGeneral_0"]["kernel"])
    xr = a + padded.reshape(B, S * S, C)
    ref = xr + V.ff_block(V.layer_norm(xr, params["LayerNorm_0"]["scale"], params["LayerNorm_0"]["bias"]),
                          params["FFBlock_0"])
    assert _rel(out_eval.v, ref) <= TOL
    # without mutable the running averages are left alone (train.py's call shape)
    before = {k: v.clone() for k, v in new.items()}
    model.apply({"params": variables["params"], "batch_stats": upd["batch_stats"]}, Var(x.cuda()),
                is_training=True)
    torch.cuda.synchronize()
    assert all(torch.equal(before[k], v) for k, v in O.tree_leaves(upd["batch_stats"]))


# ------------------------------------------------------------------------------ full BoTNet model
def _small_botnet_cfg():
    return dict(stage_sizes=(1, 1, 1, 1), num_heads=4, head_dim=128, initial_filters=64, num_classes=1000)


@pytest.mark.gpu
def test_botnet_model_logits_and_all_grads():
    """BoTNet (botnet.py:255-331) end to end at the reference widths (64..2048 channels, 4 heads x 128
    in the BoT block) and reduced depth, 64x64 images: 7x7 stride-2 stem (im2col + GEMM), BatchNorm +
    swish, 3x3 max-pool, three bottleneck blocks with squeeze-excite and projection skips (two of
    them strided), one bottleneck-transformer block (MHSA with 2-D relative logits on a 4x4 map),
    global mean, head.  Logits and every parameter gradient against the fp32 oracle."""
    import sav
    from models import BoTNet
    from sav.train import HWCN, make_loss_fn
    cfg = _small_botnet_cfg()
    B, img = 16, 64
    params = V.synthetic_botnet_params(cfg, img, seed=42)
    images, labels = O.synthetic_batch(B, img_size=img, seed=0)
    loss_ref, logits_ref, grads_ref = V.botnet_loss_and_grads(params, images, labels, cfg, 0.1)
    model = BoTNet(num_classes=1000, stage_sizes=cfg["stage_sizes"])
    x = HWCN(images.cuda())
    variables = sav.load_params(model, params, x, is_training=True)
    (loss, logits), grads = sav.value_and_grad(make_loss_fn(model, 0.1), has_aux=True)(
        variables["params"], x, labels.cuda())
    torch.cuda.synchronize()
    assert abs(loss.item() - float(loss_ref)) <= 2e-2 * abs(float(loss_ref)), (loss.item(), float(loss_ref))
    errs = {"logits": _rel(logits.v, logits_ref)}
    l2 = {}
    ours = dict(O.tree_leaves(grads))
    for path, g_ref in O.tree_leaves(grads_ref):
        errs[path] = _rel(ours[path], g_ref)
        l2[path] = float((ours[path].float().cpu() - g_ref).norm() / g_ref.norm().clamp_min(1e-30))
    print("botnet worst relative L2:", sorted(l2.items(), key=lambda kv: -kv[1])[:4])
    _dump("botnet_small", errs)
    # The floor: the oracle with round-to-bf16 at exactly the tensors a bf16 pipeline stores (conv
    # operands / outputs, BatchNorm + swish outputs, block outputs) -- an IDEAL bf16 implementation.
    # BatchNorm'd conv stacks amplify operand rounding far more than the LayerNorm'd transformer: at
    # these shapes the ideal pipeline is already ~0.12 off the fp32 result on the stem kernel and
    # ~4e-2 in the median.  The CUDA path must sit on that floor, not above it.
    _, logits_emu, grads_emu = V.botnet_loss_and_grads(params, images, labels, cfg, 0.1, emulate_bf16=True)
    emu = {p: _rel(g, dict(O.tree_leaves(grads_ref))[p]) for p, g in O.tree_leaves(grads_emu)}
    e_gpu, e_emu = sorted(errs[p] for p in emu), sorted(emu.values())
    print("botnet max-norm errors  gpu: worst %.3f median %.3f | ideal bf16: worst %.3f median %.3f | logits %.3f vs %.3f"
          % (e_gpu[-1], e_gpu[len(e_gpu) // 2], e_emu[-1], e_emu[len(e_emu) // 2], errs["logits"],
             _rel(logits_emu, logits_ref)))
    assert e_gpu[-1] <= 1.3 * e_emu[-1] and e_gpu[len(e_gpu) // 2] <= 1.3 * e_emu[len(e_emu) // 2]
    assert errs["logits"] <= max(TOL, 1.5 * _rel(logits_emu, logits_ref))


@pytest.mark.gpu
def test_maxpool_and_conv2d_ops_match_torch():
    """Op-level: 3x3 / stride-2 SAME max-pool (values and gradient routing to the arg-max) and the
    im2col + GEMM convolution (7x7 stride 2 pad 3; 3x3 stride 2 SAME on an odd map) with reverse mode."""
    from sav import functional as F
    from sav.tape import Tape, Var
    g = torch.Generator().manual_seed(9)
    x = torch.randn((3, 13, 13, 64), generator=g).bfloat16()
    xr = x.float().requires_grad_(True)
    ref = V.max_pool_3x3_s2_same(xr)
    cot = torch.randn(ref.shape, generator=g).bfloat16()
    (ref * cot.float()).sum().backward()
    with Tape() as tape:
        xv = Var(x.cuda())
        out = F.maxpool3x3s2(xv)
        out.g = cot.cuda()
        tape.backward()
    torch.cuda.synchronize()
    assert torch.equal(out.v.float().cpu(), ref.detach())
    assert _rel(xv.g, xr.grad) <= 1e-2

    class P:      # minimal parameter stand-in for functional.conv2d
        def __init__(self, w):
            self.master = w.cuda()
            self.shadow = self.master.bfloat16()
            self.grad = torch.zeros_like(self.master)

        def touch_grad(self):
            pass
    for (k, s, pad, H, cin, cout) in ((7, 2, 3, 20, 8, 64), (3, 2, "SAME", 13, 64, 96), (1, 2, "SAME", 8, 64, 128)):
        xin = torch.randn((2, H, H, cin), generator=g).bfloat16()
        w = torch.randn((k, k, cin, cout), generator=g) / (k * k * cin) ** 0.5
        xr = xin.float().requires_grad_(True)
        wr = w.bfloat16().float().requires_grad_(True)
        ref = V.conv2d_nobias(xr, wr, s, pad)
        cot = torch.randn(ref.shape, generator=g).bfloat16()
        (ref * cot.float()).sum().backward()
        pw = P(w)
        with Tape() as tape:
            xv = Var(xin.cuda())
            out = F.conv2d(xv, pw, k, s, pad)
            out.g = cot.cuda()
            tape.backward()
        torch.cuda.synchronize()
        assert tuple(out.v.shape) == tuple(ref.shape)
        assert _rel(out.v, ref) <= 1e-2 and _rel(xv.g, xr.grad) <= 1e-2 and _rel(pw.grad, wr.grad) <= 1e-2, (k, s)
