"""Parity of the CUDA path (through the module surface -> C ABI) against the fp32 CPU oracle on
identical synthetic inputs and parameters: logits and EVERY parameter gradient.

Tolerance (BASELINE.json north_star): bf16 path, per tensor max|delta| / max|ref| <= 2e-2.
"""
import pytest
import torch

from oracle import vit_oracle as O

TOL = 2e-2


def _rel(a, b):
    a = a.detach().float().cpu()
    b = b.detach().float().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-20))


def _run_case(name, batch, img, label_smoothing=0.1):
    import sav
    from models import create_model
    from sav.train import HWCN, make_loss_fn
    cfg = O.vit_config(name)
    params = O.synthetic_vit_params(cfg, img_size=img, seed=42)
    images, labels = O.synthetic_batch(batch, img_size=img, seed=0)
    loss_ref, logits_ref, grads_ref = O.vit_loss_and_grads(params, images, labels, cfg, label_smoothing)

    model = create_model(name)
    dev_images = images.cuda()
    variables = sav.load_params(model, params, HWCN(dev_images), is_training=True)
    loss_fn = make_loss_fn(model, label_smoothing)
    (loss, logits), grads = sav.value_and_grad(loss_fn, has_aux=True)(
        variables["params"], HWCN(dev_images), labels.cuda())
    torch.cuda.synchronize()
    errs = {"logits": _rel(logits.v, logits_ref)}
    assert abs(loss.item() - float(loss_ref)) <= 2e-2 * abs(float(loss_ref)), (loss.item(), float(loss_ref))
    ours = dict(O.tree_leaves(grads))
    for path, g_ref in O.tree_leaves(grads_ref):
        errs[path] = _rel(ours[path], g_ref)
    return errs


@pytest.mark.gpu
def test_deit_ti_b8_logits_and_all_grads():
    errs = _run_case("deit_ti_patch16", batch=8, img=224)
    worst = sorted(errs.items(), key=lambda kv: -kv[1])[:8]
    print("worst:", worst)
    bad = {k: v for k, v in errs.items() if not v <= TOL}
    assert not bad, f"{len(bad)} tensors over tolerance {TOL}: {sorted(bad.items(), key=lambda kv: -kv[1])[:10]}"


@pytest.mark.gpu
def test_deit_s_small_image_ragged_batch():
    # 96x96 image -> 36+1 tokens (single partial attention tile), odd batch
    errs = _run_case("deit_s_patch16", batch=3, img=96, label_smoothing=0.0)
    bad = {k: v for k, v in errs.items() if not v <= TOL}
    assert not bad, f"over tolerance: {sorted(bad.items(), key=lambda kv: -kv[1])[:10]}"
