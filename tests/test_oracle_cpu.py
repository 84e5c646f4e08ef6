"""CPU checks of the oracle (oracle/vit_oracle.py), the checker every GPU parity test trusts.

The reference has no golden vectors (its tests assert logits shape only, models/vit_test.py:26) and
cannot be imported here, so the oracle is validated by: the reference's implied
known-answer test (zero-init head => logits exactly 0), the reference's shape tests, the Flax
parameter-tree names, an independent torch.nn.functional cross-check of every formula, float64
finite differences of the backward pass, and the committed regression fixture tests/golden.
"""
import importlib.util
import math
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as TF

from oracle import vit_oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
MICRO = dict(num_layers=2, num_heads=2, embed_dim=16, patch_shape=(4, 4), num_classes=7, expand_ratio=4)


def test_gelu_is_jax_tanh_form():
    x = torch.linspace(-6, 6, 1001, dtype=torch.float64)
    assert torch.allclose(O.gelu_tanh(x), TF.gelu(x, approximate="tanh"), atol=1e-12)
    # and NOT the erf form (jax.nn.gelu default approximate=True, ff.py:12)
    assert (O.gelu_tanh(x) - TF.gelu(x)).abs().max() > 1e-4


def test_layernorm_matches_flax_semantics():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(5, 7, 32, generator=g, dtype=torch.float64) * 3 + 1
    s = torch.randn(32, generator=g, dtype=torch.float64)
    b = torch.randn(32, generator=g, dtype=torch.float64)
    ref = TF.layer_norm(x, (32,), s, b, eps=1e-6)   # flax epsilon default 1e-6 (not torch's 1e-5)
    assert torch.allclose(O.layer_norm(x, s, b), ref, atol=1e-10)
    assert not torch.allclose(O.layer_norm(x, s, b), TF.layer_norm(x, (32,), s, b, eps=1e-3), atol=1e-6)


def test_attention_matches_sdpa():
    g = torch.Generator().manual_seed(1)
    B, N, C, H = 2, 9, 32, 4
    d = C // H
    x = torch.randn(B, N, C, generator=g, dtype=torch.float64)
    p = {k: {"kernel": torch.randn(C, H, d, generator=g, dtype=torch.float64) / math.sqrt(C)}
         for k in ("queries", "keys", "values")}
    p["DenseGeneral_0"] = {"kernel": torch.randn(H, d, C, generator=g, dtype=torch.float64) / math.sqrt(C)}
    ours = O.attention(x, x, p, H)
    q = torch.einsum("bnc,chd->bhnd", x, p["queries"]["kernel"])
    k = torch.einsum("bnc,chd->bhnd", x, p["keys"]["kernel"])
    v = torch.einsum("bnc,chd->bhnd", x, p["values"]["kernel"])
    ctx = TF.scaled_dot_product_attention(q, k, v)            # scale = 1/sqrt(d), softmax over keys
    ref = torch.einsum("bhnd,hdc->bnc", ctx, p["DenseGeneral_0"]["kernel"])
    assert torch.allclose(ours, ref, atol=1e-10)


def test_patch_feature_order_is_ph_pw_c():
    """einops 'b (h ph) (w pw) c -> b (h w) (ph pw c)' (patch_embed.py:19-22): c fastest."""
    img = torch.arange(2 * 8 * 8 * 3, dtype=torch.float64).reshape(2, 8, 8, 3)
    eye = torch.eye(4 * 4 * 3, dtype=torch.float64)
    tok = O.patch_embed(img, eye, 4, 4)
    # token (gy=1, gx=0) of image 0, feature (py=2, px=3, c=1)
    assert tok[0, 2, (2 * 4 + 3) * 3 + 1] == img[0, 4 + 2, 0 + 3, 1]


def test_zero_head_known_answer():
    """vit.py:96-98: the head kernel is zeros-initialised => default-init logits are exactly 0 and
    every non-head gradient is exactly 0."""
    cfg = dict(MICRO)
    params = O.synthetic_vit_params(cfg, 8, seed=0, default_init=True)
    images, labels = O.synthetic_batch(3, 8, num_classes=cfg["num_classes"], seed=0)
    loss, logits, grads = O.vit_loss_and_grads(params, images, labels, cfg, 0.1)
    assert torch.all(logits == 0)
    assert abs(float(loss) - math.log(cfg["num_classes"])) < 1e-6
    for path, g in O.tree_leaves(grads):
        if not path.startswith("Dense_0"):
            assert torch.all(g == 0), path


@pytest.mark.parametrize("name,patch", [("vit_b_patch32", 32), ("vit_b_patch16", 16)])
def test_reference_shape_tests(name, patch):
    """models/vit_test.py:13-26: logits of a [2,224,224,3] batch have shape (2, 1000)."""
    cfg = O.vit_config(name)
    cfg["num_layers"] = 1  # depth does not change shapes; keeps the CPU suite fast
    params = O.synthetic_vit_params(cfg, 224, seed=0)
    logits = O.vit_forward(params, torch.ones(2, 224, 224, 3), cfg["num_layers"], cfg["num_heads"],
                           cfg["patch_shape"])
    assert tuple(logits.shape) == (2, 1000)


def test_param_tree_names_follow_flax_autonaming():
    cfg = O.vit_config("deit_ti_patch16")
    names = {p for p, _ in O.tree_leaves(O.synthetic_vit_params(cfg, 224))}
    for must in ("PatchEmbedBlock_0/Dense_0/kernel", "cls", "Encoder_0/AddAbsPosEmbed_0/pos_embed",
                 "Encoder_0/EncoderBlock_0/LayerNorm_0/scale",
                 "Encoder_0/EncoderBlock_11/SelfAttentionBlock_0/queries/kernel",
                 "Encoder_0/EncoderBlock_3/SelfAttentionBlock_0/DenseGeneral_0/kernel",
                 "Encoder_0/EncoderBlock_3/FFBlock_0/Dense_1/bias", "Encoder_0/LayerNorm_0/bias",
                 "Dense_0/kernel", "Dense_0/bias"):
        assert must in names, must
    n_params = sum(t.numel() for _, t in O.tree_leaves(O.synthetic_vit_params(cfg, 224)))
    assert n_params == 5_708_008  # DeiT-Ti without q/k/v/proj/patch biases (attention.py:17, patch_embed.py:12)


def test_backward_against_float64_finite_differences():
    cfg = dict(MICRO)
    params = O.synthetic_vit_params(cfg, 8, seed=5, dtype=torch.float64)
    images, labels = O.synthetic_batch(2, 8, num_classes=cfg["num_classes"], seed=1, dtype=torch.float64)
    _, _, grads = O.vit_loss_and_grads(params, images, labels, cfg, 0.1)
    flat = dict(O.tree_leaves(params))
    gflat = dict(O.tree_leaves(grads))

    def loss_of():
        logits = O.vit_forward(params, O.hwcn_to_nhwc(images), cfg["num_layers"], cfg["num_heads"],
                               cfg["patch_shape"])
        return float(O.loss_fn(logits, labels, cfg["num_classes"], 0.1))

    rng = np.random.default_rng(0)
    eps = 1e-6
    for path in ("Encoder_0/EncoderBlock_0/SelfAttentionBlock_0/queries/kernel",
                 "Encoder_0/EncoderBlock_1/FFBlock_0/Dense_0/kernel",
                 "Encoder_0/EncoderBlock_0/LayerNorm_1/scale", "cls",
                 "Encoder_0/AddAbsPosEmbed_0/pos_embed", "PatchEmbedBlock_0/Dense_0/kernel",
                 "Encoder_0/EncoderBlock_1/SelfAttentionBlock_0/DenseGeneral_0/kernel"):
        t = flat[path]
        for _ in range(3):
            idx = tuple(int(rng.integers(0, s)) for s in t.shape)
            old = float(t[idx])
            t[idx] = old + eps
            lp = loss_of()
            t[idx] = old - eps
            lm = loss_of()
            t[idx] = old
            fd = (lp - lm) / (2 * eps)
            assert abs(fd - float(gflat[path][idx])) <= 1e-6 * max(1.0, abs(fd)), (path, idx)


def test_adam_chain_matches_torch_adamw_form():
    """train.py:25-27: scale_by_adam -> additive_weight_decay -> scale(-lr): decoupled decay."""
    g = torch.Generator().manual_seed(0)
    p = torch.randn(50, generator=g, dtype=torch.float64)
    grad = torch.randn(50, generator=g, dtype=torch.float64)
    m = torch.zeros_like(p); v = torch.zeros_like(p)
    p1, m1, v1 = O.adam_chain_step(p, grad, m, v, 1, lr=1e-2, wd=0.1)
    mhat = (0.1 * grad) / (1 - 0.9)
    vhat = (0.001 * grad * grad) / (1 - 0.999)
    assert torch.allclose(p1, p - 1e-2 * (mhat / (vhat.sqrt() + 1e-8) + 0.1 * p), atol=1e-12)
    # clip_by_global_norm scales only when ||g|| >= max_norm
    p2, _, _ = O.adam_chain_step(p, grad, m, v, 1, lr=1e-2, max_norm=1e9)
    assert torch.allclose(p2, O.adam_chain_step(p, grad, m, v, 1, lr=1e-2)[0])


def test_golden_fixture_regression():
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    now = mod.compute()
    gold = np.load(os.path.join(HERE, "golden", "vit_micro.npz"))
    assert set(gold.files) == set(now)
    for k in gold.files:
        np.testing.assert_allclose(now[k], gold[k], rtol=2e-4, atol=1e-6, err_msg=k)
