"""MLP-Mixer host logic without a GPU: constructor table, parameter tree / logits shape on the meta
device (the reference's own test, mlp_mixer_test.py:13-38, asserts the logits shape only), and the
oracle restatement against float64 finite differences."""
import pytest
import torch

from oracle import variants_oracle as V
from oracle import vit_oracle as O


@pytest.mark.parametrize("name,layers,dim,patch", [
    ("mixer_s_patch32", 8, 512, 32), ("mixer_s_patch16", 8, 512, 16), ("mixer_b_patch32", 12, 768, 32),
    ("mixer_b_patch16", 12, 768, 16), ("mixer_l_patch32", 24, 1024, 32), ("mixer_l_patch16", 32, 1024, 16)])
def test_mixer_constructors_and_logits_shape(name, layers, dim, patch):
    import sav  # noqa: F401
    from models import create_model
    model = create_model(name, num_classes=1000)
    assert (model.num_layers, model.embed_dim, tuple(model.patch_shape)) == (layers, dim, (patch, patch))
    x = torch.ones((2, 224, 224, 3), device="meta")
    logits, variables = model.init_with_output({"params": 0}, x, is_training=True)
    assert tuple(logits.shape) == (2, 1000)                       # mlp_mixer_test.py:37
    got = {k: tuple(v.shape) for k, v in O.tree_leaves(variables["params"])}
    want = {k: tuple(v.shape) for k, v in O.tree_leaves(
        V.synthetic_mixer_params(V.mixer_config(name), img_size=224)) if k.startswith("MixerBlock_0/")
        or not k.startswith("MixerBlock_")}
    for k, shape in want.items():
        assert got[k] == shape, (k, got.get(k), shape)
    assert len(got) == 4 + 12 * layers + 2                        # embed, head, final LN + 12 leaves per block


def test_mixer_oracle_matches_finite_differences():
    cfg = dict(num_layers=2, embed_dim=8, patch_shape=(4, 4), tokens_expand_ratio=0.5,
               channels_expand_ratio=4, num_classes=5)
    params = V.synthetic_mixer_params(cfg, img_size=12, seed=3, dtype=torch.float64)
    images, labels = O.synthetic_batch(3, img_size=12, num_classes=5, seed=1, dtype=torch.float64)
    loss, logits, grads = V.mixer_loss_and_grads(params, images, labels, cfg, 0.1)
    assert tuple(logits.shape) == (3, 5)
    flat = dict(O.tree_leaves(params))
    gflat = dict(O.tree_leaves(grads))
    gen = torch.Generator().manual_seed(0)
    for path in ("MixerBlock_0/FFBlock_0/Dense_0/kernel", "MixerBlock_1/FFBlock_0/Dense_1/bias",
                 "MixerBlock_0/LayerNorm_0/scale", "PatchEmbedBlock_0/Dense_0/bias",
                 "MixerBlock_1/FFBlock_1/Dense_0/kernel"):
        t = flat[path]
        idx = tuple(int(torch.randint(0, s, (1,), generator=gen)) for s in t.shape)
        old = float(t[idx])
        eps = 1e-6
        vals = []
        for d in (eps, -eps):
            t[idx] = old + d
            vals.append(float(V.mixer_loss_and_grads(params, images, labels, cfg, 0.1)[0]))
        t[idx] = old
        fd = (vals[0] - vals[1]) / (2 * eps)
        assert abs(fd - float(gflat[path][idx])) <= 1e-6 * max(1.0, abs(fd)), (path, fd, float(gflat[path][idx]))
