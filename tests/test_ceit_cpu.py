"""CeiT host logic without a GPU: constructors, parameter tree / logits shape on the meta device (the
reference's own test, ceit_test.py, asserts the logits shape only), and the oracle restatement against
float64 finite differences."""
import pytest
import torch

from oracle import variants_oracle as V
from oracle import vit_oracle as O


@pytest.mark.parametrize("name", ["ceit_t", "ceit_s", "ceit_b"])
def test_ceit_constructors_param_tree_and_logits_shape(name):
    import sav  # noqa: F401
    from models import create_model
    cfg = V.ceit_config(name)
    model = create_model(name, num_classes=1000)
    assert (model.num_layers, model.num_heads, model.embed_dim) == (12, cfg["num_heads"], cfg["embed_dim"])
    x = torch.ones((2, 224, 224, 3), device="meta")
    logits, variables = model.init_with_output({"params": 0}, x, is_training=True)
    assert tuple(logits.shape) == (2, 1000)
    got = {k: tuple(v.shape) for k, v in O.tree_leaves(variables["params"])}
    want = {k: tuple(v.shape) for k, v in O.tree_leaves(V.synthetic_ceit_params(cfg))}
    assert got == want
    # running averages of every BatchNorm (image_to_token.py:30, leff.py:35-39)
    stats = {k: tuple(v.shape) for k, v in O.tree_leaves(variables["batch_stats"])}
    assert stats["Image2TokenBlock_0/BatchNorm_0/mean"] == (32,)
    hid = 4 * cfg["embed_dim"]
    assert stats["Encoder_0/EncoderBlock_11/LeFFBlock_0/BatchNorm_1/var"] == (hid,)
    assert len(stats) == 2 * (1 + 3 * 12)


def test_image_to_token_geometry():
    """224 -> conv 7x7 / 2, padding 4 -> 113 -> max_pool 3x3 / 2 'VALID' -> 56 -> 4x4 patches -> 196
    tokens of 512 features (image_to_token.py:24-44)."""
    cfg = V.ceit_config("ceit_t")
    p = V.synthetic_ceit_params(cfg)["Image2TokenBlock_0"]
    x = torch.randn((1, 224, 224, 3), generator=torch.Generator().manual_seed(0))
    assert tuple(V.image_to_token(x, p, cfg).shape) == (1, 196, 192)
    assert tuple(p["Dense_0"]["kernel"].shape) == (512, 192)


def test_ceit_oracle_matches_finite_differences():
    cfg = dict(num_layers=2, num_heads=2, embed_dim=8, patch_shape=(2, 2), num_ch=4, conv_kernel_size=3,
               conv_stride=2, pool_window_size=3, pool_stride=2, expand_ratio=2, leff_kernel_size=3,
               num_classes=5)
    # 18 -> conv 3x3 / 2, padding 2 -> (18 + 4 - 3) // 2 + 1 = 10 -> pool (10 - 3) // 2 + 1 = 4 -> 2 x 2 tokens
    params = V.synthetic_ceit_params(cfg, seed=3, dtype=torch.float64)
    images, labels = O.synthetic_batch(3, img_size=18, num_classes=5, seed=1, dtype=torch.float64)
    loss, logits, grads = V.ceit_loss_and_grads(params, images, labels, cfg, 0.1)
    assert tuple(logits.shape) == (3, 5)
    flat, gflat = dict(O.tree_leaves(params)), dict(O.tree_leaves(grads))
    gen = torch.Generator().manual_seed(0)
    for path in ("Image2TokenBlock_0/Conv_0/kernel", "Image2TokenBlock_0/BatchNorm_0/scale",
                 "Image2TokenBlock_0/Dense_0/kernel", "cls",
                 "Encoder_0/EncoderBlock_0/SelfAttentionBlock_0/keys/kernel",
                 "Encoder_0/EncoderBlock_0/LeFFBlock_0/Conv_0/kernel",
                 "Encoder_0/EncoderBlock_0/LeFFBlock_0/Conv_0/bias",
                 "Encoder_0/EncoderBlock_0/LeFFBlock_0/BatchNorm_1/bias",
                 "Encoder_0/EncoderBlock_0/LayerNorm_1/scale",
                 "Encoder_0/EncoderBlock_1/LeFFBlock_0/Dense_1/kernel",
                 "Encoder_0/EncoderBlock_1/LayerNorm_0/bias",
                 "LCSelfAttentionBlock_0/queries/kernel", "LCSelfAttentionBlock_0/DenseGeneral_0/kernel"):
        t = flat[path]
        idx = tuple(int(torch.randint(0, s, (1,), generator=gen)) for s in t.shape)
        old = float(t[idx])
        eps = 1e-6
        vals = []
        for d in (eps, -eps):
            t[idx] = old + d
            vals.append(float(V.ceit_loss_and_grads(params, images, labels, cfg, 0.1)[0]))
        t[idx] = old
        fd = (vals[0] - vals[1]) / (2 * eps)
        assert abs(fd - float(gflat[path][idx])) <= 1e-6 * max(1.0, abs(fd)), (path, fd, float(gflat[path][idx]))
