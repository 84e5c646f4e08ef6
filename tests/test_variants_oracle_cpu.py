"""CPU checks of the variants oracle (oracle/variants_oracle.py) and of the CaiT / BoTMHSA module
surface (shape propagation on the 'meta' device, like the reference's cait_test.py / botnet_test.py
which assert logits shape only).

The oracle is pinned by: float64 finite differences of its own gradients, the literal
pad/reshape restatement of `_to_absolute_logits` (botnet.py:78-93) against the direct gather, the
index identity abs[i, j] = rel[i, j - i + l - 1], a torch.nn.functional cross-check of plain
attention, and the zero-head known-answer test (cait.py:179-182 zero-init head => logits = bias).
"""
import math

import pytest
import torch

from oracle import variants_oracle as V
from oracle import vit_oracle as O


def test_to_absolute_logits_index_identity():
    b, h, l = 2, 3, 5
    rel = torch.randn(b, h, l, 2 * l - 1, dtype=torch.float64)
    out = V.to_absolute_logits(rel)
    assert out.shape == (b, h, l, l)
    for i in range(l):
        for j in range(l):
            assert torch.equal(out[:, :, i, j], rel[:, :, i, j - i + l - 1])


@pytest.mark.parametrize("H,W", [(3, 5), (4, 4), (14, 14)])
def test_bot_mhsa_literal_restatement_equals_direct_gather(H, W):
    torch.manual_seed(0)
    C, h, d = 32, 2, 16
    p = V.synthetic_bot_params(C, h, d, H, W, dtype=torch.float64)
    x = torch.randn(2, H, W, C, dtype=torch.float64)
    a, b = V.bot_mhsa(x, p, h, d), V.bot_mhsa_direct(x, p, h, d)
    assert (a - b).abs().max() < 1e-12


def test_bot_mhsa_without_tables_is_plain_attention():
    torch.manual_seed(1)
    C, h, d, H, W = 24, 3, 8, 4, 5
    p = V.synthetic_bot_params(C, h, d, H, W, dtype=torch.float64)
    p["RelativeLogits_0"]["rel_pos_emb_w"].zero_()
    p["RelativeLogits_0"]["rel_pos_emb_h"].zero_()
    x = torch.randn(2, H, W, C, dtype=torch.float64)
    out = V.bot_mhsa(x, p, h, d)
    xf = x.reshape(2, H * W, C)
    q, k, v = ((xf @ p[n]["kernel"].reshape(C, h * d)).reshape(2, H * W, h, d).transpose(1, 2)
               for n in ("query", "key", "value"))
    ref = torch.nn.functional.scaled_dot_product_attention(q, k, v)
    assert (out - ref.transpose(1, 2).reshape(2, H, W, h * d)).abs().max() < 1e-12


def _fd_check(fn, params, n_probe=6, eps=1e-6, seed=0):
    """Central finite differences of scalar fn(params) against autograd at a few coordinates of
    every leaf (float64)."""
    p = O.tree_map(lambda t: t.detach().clone().requires_grad_(True), params)
    fn(p).backward()
    g = torch.Generator().manual_seed(seed)
    worst = 0.0
    for (path, leaf), (_, base) in zip(O.tree_leaves(p), O.tree_leaves(params)):
        assert base.is_contiguous(), path
        flat = base.view(-1)
        for _ in range(min(n_probe, flat.numel())):
            i = int(torch.randint(0, flat.numel(), (1,), generator=g))
            old = float(flat[i])
            flat[i] = old + eps
            fp = float(fn(params))
            flat[i] = old - eps
            fm = float(fn(params))
            flat[i] = old
            fd = (fp - fm) / (2 * eps)
            an = float(leaf.grad.reshape(-1)[i])
            worst = max(worst, abs(fd - an) / max(1e-4, abs(fd), abs(an)))
    return worst


def test_cait_oracle_gradients_match_finite_differences():
    cfg = dict(num_layers=2, num_heads=2, embed_dim=16, layerscale_eps=1e-5, num_layers_token_only=2,
               patch_shape=(4, 4), num_classes=7, expand_ratio=4)
    params = V.synthetic_cait_params(cfg, img_size=12, seed=3, layerscale=0.3, dtype=torch.float64)
    images, labels = O.synthetic_batch(3, img_size=12, num_classes=7, seed=0, dtype=torch.float64)
    gen = torch.Generator().manual_seed(5)
    factors = [torch.floor(0.8 + torch.rand((3,), generator=gen)).double() / 0.8 for _ in range(8)]
    nhwc = O.hwcn_to_nhwc(images)

    def fn(p):
        return O.loss_fn(V.cait_forward(p, nhwc, cfg, factors), labels, 7, 0.1)
    assert _fd_check(fn, params) < 1e-5


def test_bot_oracle_gradients_match_finite_differences():
    C, h, d, H, W = 12, 2, 8, 3, 4
    params = V.synthetic_bot_params(C, h, d, H, W, dtype=torch.float64)
    x = torch.randn(2, H, W, C, dtype=torch.float64, generator=torch.Generator().manual_seed(0))
    cot = torch.randn(2, H, W, h * d, dtype=torch.float64, generator=torch.Generator().manual_seed(1))

    def fn(p):
        return (V.bot_mhsa(x, p, h, d) * cot).sum()
    assert _fd_check(fn, params) < 1e-6


def test_cait_zero_head_known_answer():
    """cait.py:179-182: head kernel zeros => logits equal the head bias for any input."""
    cfg = dict(num_layers=1, num_heads=2, embed_dim=16, layerscale_eps=1e-5, num_layers_token_only=2,
               patch_shape=(4, 4), num_classes=5, expand_ratio=4)
    params = V.synthetic_cait_params(cfg, img_size=8, seed=0)
    params["Dense_0"]["kernel"].zero_()
    images, _ = O.synthetic_batch(2, img_size=8, num_classes=5)
    logits = V.cait_forward(params, O.hwcn_to_nhwc(images), cfg)
    assert torch.equal(logits, params["Dense_0"]["bias"].expand(2, -1))


def test_class_attention_uses_only_row_zero_as_query():
    """cait.py:10-15: changing the queries of patch tokens must not change the CA block output."""
    cfg = dict(num_layers=0, num_heads=2, embed_dim=16, layerscale_eps=1e-5, num_layers_token_only=1,
               patch_shape=(4, 4), num_classes=5, expand_ratio=4)
    p = V.synthetic_cait_params(cfg, img_size=8, seed=0, layerscale=0.5)["CAEncoderBlock_0"]
    x = torch.randn(2, 4, 16)
    cls = torch.randn(2, 1, 16)
    out = V.cait_ca_block(x, cls, p, 2)
    assert out.shape == (2, 1, 16)


def test_cvt_same_padding_matches_jax_convention():
    # even input, stride 2: the single padded row / column is on the HIGH side; odd input: both
    assert V.same_pad_amounts(56, 2) == (0, 1)
    assert V.same_pad_amounts(15, 2) == (1, 1)
    assert V.same_pad_amounts(14, 1) == (1, 1)


def test_cvt_stage_block_oracle_shapes_and_finite_differences():
    C, h = 16, 2
    params = V.synthetic_cvt_stage_params(C, h, dtype=torch.float64)
    # 10 tokens -> padded to 16 (4x4 map); keys/values on the 2x2 strided map
    x = torch.randn(3, 10, C, dtype=torch.float64, generator=torch.Generator().manual_seed(0))
    out = V.cvt_stage_block(x, params, h)
    assert out.shape == (3, 16, C)
    cot = torch.randn(3, 16, C, dtype=torch.float64, generator=torch.Generator().manual_seed(1))

    def fn(p):
        return (V.cvt_stage_block(x, p, h) * cot).sum()
    assert _fd_check(fn, params, n_probe=4, eps=1e-5) < 1e-4   # key-BN bias: exact-zero gradient vs FD noise


def test_cvt_stage_block_shape_and_param_names():
    """models/cvt_test.py only asserts logits shape; here: the stage block pads 197 tokens to a
    15 x 15 map (cvt.py:10-16) and the parameter tree matches the oracle's."""
    from models import StageBlock
    from sav.tape import Var
    m = StageBlock(num_heads=6, embed_dim=384)
    x = Var(torch.empty((2, 197, 384), device="meta", dtype=torch.float32))
    out, variables = m.init_with_output({"params": 0}, x, is_training=True)
    assert tuple(out.shape) == (2, 225, 384)
    ours = {k: tuple(v.shape) for k, v in O.tree_leaves(variables["params"])}
    ref = {k: tuple(v.shape) for k, v in O.tree_leaves(V.synthetic_cvt_stage_params(384, 6))}
    assert ours == ref


# ------------------------------------------------------------------------- module surface (meta)
@pytest.mark.parametrize("name,heads,dim", [("cait_xxs_24", 4, 192), ("cait_xs_24", 6, 288),
                                            ("cait_s_24", 8, 384)])
def test_cait_logits_shape_and_param_tree_like_reference(name, heads, dim):
    """models/cait_test.py:13-40: init_with_output on ones[2,224,224,3] -> logits (2, 1000); the
    parameter tree has the oracle's (= Flax auto-naming) names and shapes."""
    from models import create_model
    model = create_model(name, num_classes=1000)
    x = torch.empty((2, 224, 224, 3), device="meta")
    logits, variables = model.init_with_output({"params": 0, "stochastic_depth": 1}, x, is_training=True)
    assert tuple(logits.shape) == (2, 1000)
    ours = {k: tuple(v.shape) for k, v in O.tree_leaves(variables["params"])}
    cfg = V.cait_config(name)
    ref = {k: tuple(v.shape) for k, v in O.tree_leaves(V.synthetic_cait_params(cfg))}
    assert ours == ref


def test_cait_training_needs_the_stochastic_depth_rng():
    """cait_test.py:37-39 supplies no 'stochastic_depth' rng although drop-path > 0;
    like Flax's make_rng, that is a loud error here."""
    from models import create_model
    model = create_model("cait_xxs_24")
    x = torch.empty((2, 224, 224, 3), device="meta")
    with pytest.raises(KeyError, match="stochastic_depth"):
        model.init_with_output({"params": 0}, x, is_training=True)


def test_bot_mhsa_shape_and_param_names():
    from models import BoTMHSA
    from sav.tape import Var
    m = BoTMHSA(num_heads=4, head_ch=128)
    x = Var(torch.empty((2, 14, 14, 512), device="meta", dtype=torch.bfloat16))
    out, variables = m.init_with_output({"params": 0}, x)
    assert tuple(out.shape) == (2, 14, 14, 512)
    ours = {k: tuple(v.shape) for k, v in O.tree_leaves(variables["params"])}
    ref = {k: tuple(v.shape) for k, v in O.tree_leaves(V.synthetic_bot_params(512, 4, 128, 14, 14))}
    assert ours == ref


def test_talking_heads_block_is_part_of_the_surface():
    from models.layers import TalkingHeadsBlock
    from models.layers.attentions import TalkingHeadsBlock as T2
    assert TalkingHeadsBlock is T2


def test_rotary_helpers_are_exported_and_shape_only_on_meta():
    """models/layers/__init__.py:4 exports RotaryPositionalEmbedding / FixedPositionalEmbedding."""
    from models.layers import FixedPositionalEmbedding, RotaryPositionalEmbedding
    from sav.tape import Var
    x = Var(torch.empty((197, 6, 64), device="meta", dtype=torch.bfloat16))
    out, variables = RotaryPositionalEmbedding().init_with_output({"params": 0}, x)
    assert tuple(out.shape) == (197, 6, 64) and not list(O.tree_leaves(variables["params"]))
    emb, _ = FixedPositionalEmbedding().init_with_output({"params": 0}, x)
    assert tuple(emb.shape) == (197, 64)


def test_rotary_oracle_properties():
    """The restated helpers: rotate_every_two is the literal (-x1, x0, -x3, x2, ...) pattern, the
    rotation is orthogonal (norm preserving, inverse = rotation by -theta), position 0 is the
    identity, and the literal float32 frequency table is zero from the fifth frequency on."""
    x = torch.arange(8.0).reshape(1, 1, 8)
    assert V.rotate_every_two(x).flatten().tolist() == [-1, 0, -3, 2, -5, 4, -7, 6]
    emb = V.fixed_positional_embedding(5, 64)
    assert emb.shape == (5, 64) and torch.equal(emb[:, :32], emb[:, 32:])
    assert float(emb[1, 0]) == 64.0 and float(emb[1, 1]) == pytest.approx(64.0 / 1e10)
    assert torch.all(emb[:, 4:32] == 0)       # 10e4**8 overflows float32 -> 1/inf
    g = torch.Generator().manual_seed(0)
    x = torch.randn((7, 3, 64), generator=g)
    y = V.apply_rotary_pos_emb(x, 0)
    assert torch.allclose(y[0], x[0]) and torch.allclose(y.norm(dim=-1), x.norm(dim=-1), rtol=1e-5)
    pairs = lambda t: t.reshape(7, 3, 32, 2)
    assert torch.allclose(pairs(y).norm(dim=-1), pairs(x).norm(dim=-1), rtol=1e-5)
