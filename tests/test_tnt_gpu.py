"""TNT (models/tnt.py:10-182) on the B200 path against the fp32 oracle (oracle/variants_oracle.py:
tnt_forward): the small-head attention kernel and the pixel rearrangement alone, then logits and
every parameter gradient of the whole model."""
import math

import pytest
import torch

from oracle import variants_oracle as V
from oracle import vit_oracle as O

TOL = 2e-2   # max|d| / max|ref| per tensor, bf16 path vs fp32 oracle


def _rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.gpu
@pytest.mark.parametrize("B,N,H,D", [(37, 16, 4, 6), (5, 16, 4, 10), (3, 32, 2, 16), (2, 7, 3, 5)])
def test_small_head_attention_fwd_bwd(B, N, H, D):
    """sav_attn_small_fwd / _bwd against attention.py:39-58 in fp32 torch on the same bf16 inputs."""
    from sav import ops
    g = torch.Generator().manual_seed(B * N + D)
    qkv = torch.randn((B, N, 3, H, D), generator=g).bfloat16()
    d_o = torch.randn((B, N, H, D), generator=g).bfloat16()
    scale = 1.0 / math.sqrt(D)
    ref_in = qkv.float().requires_grad_(True)
    q, k, v = ref_in[:, :, 0], ref_in[:, :, 1], ref_in[:, :, 2]
    w = torch.softmax(torch.einsum("bqhd,bkhd->bhqk", q * scale, k), dim=-1)
    o_ref = torch.einsum("bhqk,bkhd->bqhd", w, v)
    o_ref.backward(d_o.float())
    o = ops.attn_small_fwd(qkv.cuda(), scale)
    dqkv = ops.attn_small_bwd(qkv.cuda(), d_o.cuda(), scale)
    torch.cuda.synchronize()
    assert _rel(o, o_ref) <= 1e-2                      # bf16 output rounding only
    for i, name in enumerate("qkv"):
        assert _rel(dqkv[:, :, i], ref_in.grad[:, :, i]) <= 1e-2, name


@pytest.mark.gpu
@pytest.mark.parametrize("hwcn", [False, True])
def test_pixel_patchify_matches_the_two_rearranges(hwcn):
    from sav import ops
    x = torch.randn((3, 32, 48, 3), generator=torch.Generator().manual_seed(1))
    eye = {"Dense_0": {"kernel": torch.eye(48), "bias": torch.zeros(48)}}
    want = V.pixel_embed(x, eye, (16, 16), (4, 4))                 # tnt.py:19-26 restated
    src = x.permute(1, 2, 3, 0).contiguous() if hwcn else x
    got = ops.pixel_patchify(src.cuda(), 16, 16, 4, 4, hwcn)
    torch.cuda.synchronize()
    assert tuple(got.shape) == (3 * 2 * 3, 16, 48)
    assert torch.equal(got.float().cpu(), want.bfloat16().float())


@pytest.mark.gpu
@pytest.mark.parametrize("name,batch", [("tnt_b_patch16", 4), ("tnt_s_patch16", 2)])
def test_tnt_logits_and_all_grads(name, batch):
    """Both create_model configurations at reduced depth (3 blocks, so the pixel stream's two-consumer
    gradient fan-in -- Inner2OuterBlock of block i and block i+1 -- is exercised twice), 224^2."""
    import sav
    from models import TNT
    from sav.train import HWCN, make_loss_fn
    cfg = V.tnt_config(name)
    cfg["num_layers"] = 3
    params = V.synthetic_tnt_params(cfg, img_size=224, seed=42)
    images, labels = O.synthetic_batch(batch, img_size=224, seed=0)
    loss_ref, logits_ref, grads_ref = V.tnt_loss_and_grads(params, images, labels, cfg, 0.1)
    model = TNT(num_classes=1000, num_layers=3, inner_num_heads=cfg["inner_num_heads"],
                outer_num_heads=cfg["outer_num_heads"], inner_embed_dim=cfg["inner_embed_dim"],
                outer_embed_dim=cfg["outer_embed_dim"])
    x = HWCN(images.cuda())
    variables = sav.load_params(model, params, x, is_training=True)
    (loss, logits), grads = sav.value_and_grad(make_loss_fn(model, 0.1), has_aux=True)(
        variables["params"], x, labels.cuda())
    torch.cuda.synchronize()
    assert abs(loss.item() - float(loss_ref)) <= 2e-2 * abs(float(loss_ref))
    errs = {"logits": _rel(logits.v, logits_ref)}
    l2 = {}
    ours = dict(O.tree_leaves(grads))
    for path, g_ref in O.tree_leaves(grads_ref):
        errs[path] = _rel(ours[path], g_ref)
        l2[path] = float((ours[path].float().cpu() - g_ref).norm() / g_ref.norm().clamp_min(1e-30))
    worst = sorted(errs.items(), key=lambda kv: -kv[1])[:4]
    print("tnt worst max-norm:", worst, "worst L2:", sorted(l2.items(), key=lambda kv: -kv[1])[:3])
    # On the bf16 floor like the ViT cases (tests/test_baseline_configs_gpu.py):
    # logits and every tensor's relative L2 within 2e-2; the max-norm of the worst of ~100 tensors is a
    # tail statistic (the inner transformer's 24 x 24 / 40 x 40 kernels have few elements to average
    # over) -- at most two tensors may exceed 2e-2, and those by no more than 1.3 x.
    assert errs["logits"] <= TOL
    assert max(l2.values()) <= TOL, sorted(l2.items(), key=lambda kv: -kv[1])[:4]
    over = {k: v for k, v in errs.items() if v > TOL}
    assert len(over) <= 2 and all(v <= 1.3 * TOL for v in over.values()), over
