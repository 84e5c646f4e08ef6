"""MLP-Mixer (models/mlp_mixer.py:10-60) on the B200 path against the fp32 oracle
(oracle/variants_oracle.py: mixer_forward): logits and every parameter gradient."""
import pytest
import torch

from oracle import variants_oracle as V
from oracle import vit_oracle as O

TOL = 2e-2   # max|d| / max|ref| per tensor, bf16 path vs fp32 oracle


def _rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.gpu
@pytest.mark.parametrize("B,R,C,pad,src_f32,out_f32,res", [
    (3, 196, 512, 200, False, False, False),     # tokens -> channels-major rows, padded
    (3, 512, 196, 512, False, True, True),       # back, fused with the residual add (pitch 200 source)
    (2, 49, 768, 56, True, False, False),        # fp32 stream gradient in
    (1, 98, 196, 104, True, False, False),       # a weight: fp32 master -> bf16 W^T
    (2, 33, 70, 40, False, False, False),        # ragged tiles
])
def test_transpose_pad(B, R, C, pad, src_f32, out_f32, res):
    from sav import ops
    g = torch.Generator().manual_seed(R * C)
    pitch = C + (8 - C % 8) % 8 + 8
    src = torch.randn((B, R, pitch), generator=g).to(torch.float32 if src_f32 else torch.bfloat16).cuda()
    out = torch.full((B, C, pad + 8), 7.0, dtype=torch.float32 if out_f32 else torch.bfloat16, device="cuda")
    r = torch.randn((B, C, R), generator=g).cuda() if res else None
    ops.transpose_pad(src, R, C, out, pad, residual=r)
    torch.cuda.synchronize()
    want = src[:, :, :C].float().transpose(1, 2)
    if res:
        want = want + r
    want = want.to(out.dtype).float()
    assert torch.equal(out[:, :, :R].float(), want)
    assert torch.all(out[:, :, R:pad] == 0) and torch.all(out[:, :, pad:] == 7.0)   # pad zeroed, rest untouched


def _small_cfg(patch):
    cfg = V.mixer_config(f"mixer_s_patch{patch}")
    cfg["num_layers"] = 3
    return cfg


@pytest.mark.gpu
@pytest.mark.parametrize("patch,batch", [(16, 8), (32, 16)])
def test_mixer_logits_and_all_grads(patch, batch):
    """Mixer-S widths at reduced depth: /16 has 196 tokens -> token-mixing hidden 98 (neither a
    multiple of 8: both padded inside TokenMix), /32 has 49 tokens -> hidden 24."""
    import sav
    from models import MLPMixer
    from sav.train import HWCN, make_loss_fn
    cfg = _small_cfg(patch)
    params = V.synthetic_mixer_params(cfg, img_size=224, seed=42)
    images, labels = O.synthetic_batch(batch, img_size=224, seed=0)
    loss_ref, logits_ref, grads_ref = V.mixer_loss_and_grads(params, images, labels, cfg, 0.1)
    model = MLPMixer(num_classes=1000, num_layers=cfg["num_layers"], embed_dim=cfg["embed_dim"],
                     patch_shape=cfg["patch_shape"])
    x = HWCN(images.cuda())
    variables = sav.load_params(model, params, x, is_training=True)
    (loss, logits), grads = sav.value_and_grad(make_loss_fn(model, 0.1), has_aux=True)(
        variables["params"], x, labels.cuda())
    torch.cuda.synchronize()
    assert abs(loss.item() - float(loss_ref)) <= 2e-2 * abs(float(loss_ref))
    errs = {"logits": _rel(logits.v, logits_ref)}
    l2 = {}
    ours = dict(O.tree_leaves(grads))
    ref = dict(O.tree_leaves(grads_ref))
    for path, g_ref in O.tree_leaves(grads_ref):
        if path.endswith("FFBlock_0/Dense_1/bias"):
            # The token-mixing output bias adds one constant to every channel of a token, and every
            # reader of the stream starts with a LayerNorm over the channels: its gradient is
            # analytically zero (the oracle returns fp32 round-off, ~1e-9 of its sibling).  Compare
            # it on the scale of the first Dense's bias gradient instead.
            sib = ref[path.replace("Dense_1", "Dense_0")]
            assert float(g_ref.abs().max()) < 1e-5 * float(sib.abs().max())
            assert float(ours[path].float().abs().max().cpu()) < TOL * float(sib.abs().max()), path
            continue
        errs[path] = _rel(ours[path], g_ref)
        l2[path] = float((ours[path].float().cpu() - g_ref).norm() / g_ref.norm())
    worst = sorted(errs.items(), key=lambda kv: -kv[1])[:4]
    print("mixer worst max-norm:", worst, "worst L2:", sorted(l2.items(), key=lambda kv: -kv[1])[:3])
    assert errs["logits"] <= TOL
    assert max(errs.values()) <= TOL, worst
    assert max(l2.values()) <= TOL
