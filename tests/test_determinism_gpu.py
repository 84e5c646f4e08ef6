"""No kernel may consume uninitialised memory: the same step run again after the caching allocator
has been filled with NaN bit patterns (so every `torch.empty` the host code hands to a kernel comes
back full of NaNs instead of the zeros a fresh process sees) must reproduce the first run.  The
weight-gradient GEMMs combine split-K partials with fp32 reduce-adds in arrival order, so the
comparison is to fp32 rounding, not bitwise."""
import pytest
import torch

from oracle import vit_oracle as O


def _poison(nbytes=6 << 30):
    """Allocate, NaN-fill and free a large block (and a spread of small ones) so the allocator's
    free lists hand out NaN-filled memory from now on."""
    big = torch.full((nbytes // 4,), float("nan"), device="cuda")
    smalls = [torch.full((n,), float("nan"), device="cuda") for n in (257, 4099, 65537, 1 << 20, 5 << 20) for _ in range(8)]
    torch.cuda.synchronize()
    del big, smalls


@pytest.mark.gpu
@pytest.mark.parametrize("name,batch,img", [("deit_ti_patch16", 8, 224), ("deit_s_patch16", 32, 224)])
def test_gradients_do_not_depend_on_stale_memory(name, batch, img):
    import sav
    from models import create_model
    from sav.train import HWCN, make_loss_fn
    cfg = O.vit_config(name)
    params = O.synthetic_vit_params(cfg, img_size=img, seed=42)
    images, labels = O.synthetic_batch(batch, img_size=img, seed=0)
    model = create_model(name)
    x = HWCN(images.cuda())
    y = labels.cuda()
    variables = sav.load_params(model, params, x, is_training=True)
    fn = sav.value_and_grad(make_loss_fn(model, 0.1), has_aux=True)

    def run():
        (loss, logits), grads = fn(variables["params"], x, y)
        torch.cuda.synchronize()
        return loss.item(), logits.v.clone(), {k: v.clone() for k, v in O.tree_leaves(grads)}
    l0, lg0, g0 = run()
    _poison()
    l1, lg1, g1 = run()
    _poison(2 << 30)
    l2, lg2, g2 = run()
    # the loss is an atomicAdd over samples: equal to fp32 rounding, the logits bit for bit
    assert abs(l0 - l1) <= 1e-6 * abs(l0) and abs(l0 - l2) <= 1e-6 * abs(l0)
    assert torch.equal(lg0, lg1) and torch.equal(lg0, lg2)
    worst = 0.0
    for k in g0:
        assert torch.isfinite(g1[k]).all() and torch.isfinite(g2[k]).all(), k
        for other in (g1, g2):
            d = float((g0[k] - other[k]).abs().max() / g0[k].abs().max().clamp_min(1e-30))
            worst = max(worst, d)
            assert d <= 1e-4, (k, d)
    print("worst run-to-run relative difference", worst)
