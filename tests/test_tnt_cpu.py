"""TNT host logic without a GPU: constructors, parameter tree / logits shape on the meta device (the
reference's own test, tnt_test.py, asserts the logits shape only), and the oracle restatement against
float64 finite differences."""
import pytest
import torch

from oracle import variants_oracle as V
from oracle import vit_oracle as O


@pytest.mark.parametrize("name", ["tnt_s_patch16", "tnt_b_patch16"])
def test_tnt_constructors_param_tree_and_logits_shape(name):
    import sav  # noqa: F401
    from models import create_model
    cfg = V.tnt_config(name)
    model = create_model(name, num_classes=1000)
    assert (model.num_layers, model.inner_num_heads, model.outer_num_heads, model.inner_embed_dim,
            model.outer_embed_dim) == (12, cfg["inner_num_heads"], cfg["outer_num_heads"],
                                       cfg["inner_embed_dim"], cfg["outer_embed_dim"])
    x = torch.ones((2, 224, 224, 3), device="meta")
    logits, variables = model.init_with_output({"params": 0}, x, is_training=True)
    assert tuple(logits.shape) == (2, 1000)
    got = {k: tuple(v.shape) for k, v in O.tree_leaves(variables["params"])}
    want = {k: tuple(v.shape) for k, v in O.tree_leaves(V.synthetic_tnt_params(cfg, img_size=224))}
    assert got == want


def test_tnt_oracle_matches_finite_differences():
    cfg = dict(num_layers=2, inner_num_heads=2, outer_num_heads=2, inner_embed_dim=4, outer_embed_dim=8,
               patch_shape=(4, 4), transformed_patch_shape=(2, 2), inner_expand_ratio=2,
               outer_expand_ratio=2, num_classes=5)
    params = V.synthetic_tnt_params(cfg, img_size=8, seed=3, dtype=torch.float64)
    images, labels = O.synthetic_batch(3, img_size=8, num_classes=5, seed=1, dtype=torch.float64)
    loss, logits, grads = V.tnt_loss_and_grads(params, images, labels, cfg, 0.1)
    assert tuple(logits.shape) == (3, 5)
    flat, gflat = dict(O.tree_leaves(params)), dict(O.tree_leaves(grads))
    gen = torch.Generator().manual_seed(0)
    for path in ("PixelEmbedBlock_0/Dense_0/kernel", "AddAbsPosEmbed_0/pos_embed", "cls",
                 "Encoder_0/EncoderBlock_0/SelfAttentionBlock_0/keys/kernel",
                 "Encoder_0/EncoderBlock_0/Inner2OuterBlock_0/Dense_0/kernel",
                 "Encoder_0/EncoderBlock_1/FFBlock_0/Dense_1/bias",
                 "Encoder_0/EncoderBlock_1/LayerNorm_2/scale"):
        t = flat[path]
        idx = tuple(int(torch.randint(0, s, (1,), generator=gen)) for s in t.shape)
        old = float(t[idx])
        eps = 1e-6
        vals = []
        for d in (eps, -eps):
            t[idx] = old + d
            vals.append(float(V.tnt_loss_and_grads(params, images, labels, cfg, 0.1)[0]))
        t[idx] = old
        fd = (vals[0] - vals[1]) / (2 * eps)
        assert abs(fd - float(gflat[path][idx])) <= 1e-6 * max(1.0, abs(fd)), (path, fd, float(gflat[path][idx]))
