"""CeiT (models/ceit.py:10-156, leff.py, image_to_token.py) on the B200 path against the fp32 oracle
(oracle/variants_oracle.py: ceit_forward): the new operators alone ('VALID' max-pool, activation
patchify, BatchNorm + tanh-GELU, biased convolution, the 1-query / 12-key layer-wise class attention),
then logits and every parameter gradient of the whole model.

Tolerances.  The Dense / Conv biases that sit directly in front of a BatchNorm (leff.py:41-42,48-54,
59-60) have an exactly-zero gradient -- BatchNorm subtracts the batch mean -- so a relative error
is meaningless there: they are checked to be ~0 on the scale of the same layer's kernel gradient.
Every other tensor is compared with the fp32 oracle AND with the oracle's ideal-bf16 emulation
(round-to-bf16 at the tensors a bf16 pipeline stores; BatchNorm'd stacks amplify operand rounding,
like BoTNet's, tests/test_variants_gpu.py): the CUDA path must sit on that floor.
"""
import json
import math
import os

import pytest
import torch

from oracle import variants_oracle as V
from oracle import vit_oracle as O

TOL = 2e-2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _l2(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


class _P:      # minimal parameter stand-in for the functional ops
    def __init__(self, w):
        self.master = w.float().cuda()
        self.shadow = self.master.bfloat16()
        self.grad = torch.zeros_like(self.master)

    def touch_grad(self):
        pass


@pytest.mark.gpu
def test_valid_maxpool_and_activation_patchify():
    """image_to_token.py:34-44: nn.max_pool 3x3 / 2 with flax's default 'VALID' padding on the odd
    113 x 113 map of a 224^2 image (bit-exact values, gradient routed to the arg-max), then
    'b (h ph) (w pw) c -> b (h w) (ph pw c)' and its inverse (bit-exact permutations)."""
    from sav import functional as F
    from sav.tape import Tape, Var
    g = torch.Generator().manual_seed(5)
    for (B, H, C) in ((2, 113, 32), (3, 9, 8), (1, 12, 64)):
        x = torch.randn((B, H, H, C), generator=g).bfloat16()
        xr = x.float().requires_grad_(True)
        ref = torch.nn.functional.max_pool2d(xr.permute(0, 3, 1, 2), 3, 2).permute(0, 2, 3, 1)
        cot = torch.randn(ref.shape, generator=g).bfloat16()
        (ref * cot.float()).sum().backward()
        with Tape() as tape:
            xv = Var(x.cuda())
            out = F.maxpool3x3s2(xv, valid=True)
            out.g = cot.cuda()
            tape.backward()
        torch.cuda.synchronize()
        assert tuple(out.v.shape) == (B, (H - 3) // 2 + 1, (H - 3) // 2 + 1, C)
        assert torch.equal(out.v.float().cpu(), ref.detach())
        assert _rel(xv.g, xr.grad) <= 1e-2
    x = torch.randn((3, 8, 12, 32), generator=g).bfloat16()
    want = O.patch_embed(x.float(), torch.eye(4 * 4 * 32), 4, 4)
    cot = torch.randn(want.shape, generator=g).bfloat16()
    with Tape() as tape:
        xv = Var(x.cuda())
        out = F.patchify_act(xv, 4, 4)
        out.g = cot.cuda()
        tape.backward()
    torch.cuda.synchronize()
    assert tuple(out.v.shape) == (3, 6, 512)
    assert torch.equal(out.v.float().cpu(), want)
    # the gradient of a permutation is the inverse permutation of the cotangent
    xr = x.float().requires_grad_(True)
    (O.patch_embed(xr, torch.eye(4 * 4 * 32), 4, 4) * cot.float()).sum().backward()
    assert torch.equal(xv.g.float().cpu(), xr.grad)


@pytest.mark.gpu
def test_batchnorm_gelu_and_biased_conv():
    """leff.py:41-55: BatchNorm (batch statistics over every row) + tanh-GELU in one kernel pair on a
    token-shaped [B, L, C] activation, and the 3x3 'SAME' convolution with bias (im2col + GEMM, bias in
    the epilogue, bias gradient = column sums of dY)."""
    from sav import functional as F
    from sav.tape import Tape, Var
    g = torch.Generator().manual_seed(11)
    B, L, C = 4, 196, 96
    y = (torch.randn((B, L, C), generator=g) * 1.5 + 0.3).bfloat16()
    gamma, beta = _P(1.0 + 0.1 * torch.randn(C, generator=g)), _P(0.1 * torch.randn(C, generator=g))
    yr = y.float().requires_grad_(True)
    gr, br = gamma.master.cpu().requires_grad_(True), beta.master.cpu().requires_grad_(True)
    ref = O.gelu_tanh(V.batch_norm_train(yr, {"scale": gr, "bias": br}))
    cot = torch.randn(ref.shape, generator=g).bfloat16()
    (ref * cot.float()).sum().backward()
    with Tape() as tape:
        yv = Var(y.cuda())
        out = F.batchnorm_act(yv, gamma, beta, 1e-5, act="gelu")
        out.g = cot.cuda()
        tape.backward()
    torch.cuda.synchronize()
    assert _rel(out.v, ref) <= 1e-2 and _rel(yv.g, yr.grad) <= 1.5e-2
    assert _rel(gamma.grad, gr.grad) <= 1e-2 and _rel(beta.grad, br.grad) <= 1e-2

    k, cin, cout, H = 3, 64, 96, 14
    xin = torch.randn((2, H, H, cin), generator=g).bfloat16()
    w = torch.randn((k, k, cin, cout), generator=g) / (k * k * cin) ** 0.5
    b = 0.1 * torch.randn(cout, generator=g)
    xr = xin.float().requires_grad_(True)
    wr = w.bfloat16().float().requires_grad_(True)
    b_r = b.clone().requires_grad_(True)
    ref = V.conv2d_same(xr, wr, b_r, 1)
    cot = torch.randn(ref.shape, generator=g).bfloat16()
    (ref * cot.float()).sum().backward()
    pw, pb = _P(w), _P(b)
    with Tape() as tape:
        xv = Var(xin.cuda())
        out = F.conv2d(xv, pw, k, 1, "SAME", bias=pb)
        out.g = cot.cuda()
        tape.backward()
    torch.cuda.synchronize()
    assert tuple(out.v.shape) == tuple(ref.shape)
    assert _rel(out.v, ref) <= 1e-2 and _rel(xv.g, xr.grad) <= 1e-2
    assert _rel(pw.grad, wr.grad) <= 1e-2 and _rel(pb.grad, b_r.grad) <= 1e-2


@pytest.mark.gpu
def test_layerwise_class_attention_shape():
    """ceit.py:11-16: one query (the last class token) over 12 keys, head dim 64 -- the
    materialised-logits path with Nq = 1, Nk = 12 (not a multiple of 8)."""
    from sav import ops
    g = torch.Generator().manual_seed(3)
    B, Nq, Nk, H, D = 4, 1, 12, 3, 64
    q = torch.randn((B, Nq, H, D), generator=g).bfloat16()
    k = torch.randn((B, Nk, H, D), generator=g).bfloat16()
    v = torch.randn((B, Nk, H, D), generator=g).bfloat16()
    d_o = torch.randn((B, Nq, H, D), generator=g).bfloat16()
    scale = 1.0 / math.sqrt(D)
    qr, kr, vr = (t.float().requires_grad_(True) for t in (q, k, v))
    w = torch.softmax(torch.einsum("bqhd,bkhd->bhqk", qr * scale, kr), dim=-1)
    ref = torch.einsum("bhqk,bkhd->bqhd", w, vr)
    ref.backward(d_o.float())
    qc, kc, vc = q.cuda(), k.cuda(), v.cuda()
    o, S, P = ops.attn_mat_fwd(qc, kc, vc, scale)
    dq, dk, dv = (torch.empty_like(t) for t in (qc, kc, vc))
    ops.attn_mat_bwd(qc, kc, vc, S, P, d_o.cuda(), dq, dk, dv, scale)
    torch.cuda.synchronize()
    assert _rel(o, ref) <= 1e-2
    assert _rel(dq, qr.grad) <= 1.5e-2 and _rel(dk, kr.grad) <= 1.5e-2 and _rel(dv, vr.grad) <= 1.5e-2


def _zero_grad_biases(cfg):
    """Biases in front of a BatchNorm: d loss / d bias == 0 exactly (the batch mean is subtracted)."""
    return {f"Encoder_0/EncoderBlock_{i}/LeFFBlock_0/{layer}/bias": f"Encoder_0/EncoderBlock_{i}/LeFFBlock_0/{layer}/kernel"
            for i in range(cfg["num_layers"]) for layer in ("Dense_0", "Conv_0", "Dense_1")}


@pytest.mark.gpu
@pytest.mark.parametrize("name,batch,layers", [("ceit_t", 8, 3), ("ceit_s", 4, 2)])
def test_ceit_logits_and_all_grads(name, batch, layers):
    """create_model's CeiT-T / CeiT-S widths at reduced depth, 224^2 (196 tokens + class token): the
    conv stem, post-LN blocks whose LayerNorm outputs are the residual stream, LeFF on the 14 x 14
    grid, the class tokens of every block (a three-consumer gradient fan-in on each block output:
    next block's attention operand, its skip input, the class-token list) and the layer-wise class
    attention.  Logits and every parameter gradient."""
    import sav
    from models import CeiT
    from sav.train import HWCN, make_loss_fn
    cfg = V.ceit_config(name)
    cfg["num_layers"] = layers
    params = V.synthetic_ceit_params(cfg, seed=42)
    images, labels = O.synthetic_batch(batch, img_size=224, seed=0)
    loss_ref, logits_ref, grads_ref = V.ceit_loss_and_grads(params, images, labels, cfg, 0.1)
    model = CeiT(num_classes=1000, num_layers=layers, num_heads=cfg["num_heads"], embed_dim=cfg["embed_dim"])
    x = HWCN(images.cuda())
    variables = sav.load_params(model, params, x, is_training=True)
    (loss, logits), grads = sav.value_and_grad(make_loss_fn(model, 0.1), has_aux=True)(
        variables["params"], x, labels.cuda())
    torch.cuda.synchronize()
    ours = dict(O.tree_leaves(grads))
    ref = dict(O.tree_leaves(grads_ref))
    assert set(ours) == set(ref)
    zero = _zero_grad_biases(cfg)
    errs, l2 = {"logits": _rel(logits.v, logits_ref)}, {}
    for path, g_ref in ref.items():
        if path in zero:
            continue
        errs[path], l2[path] = _rel(ours[path], g_ref), _l2(ours[path], g_ref)
    # ideal bf16 pipeline = the floor
    _, logits_emu, grads_emu = V.ceit_loss_and_grads(params, images, labels, cfg, 0.1, emulate_bf16=True)
    emu = {p: _rel(g, ref[p]) for p, g in O.tree_leaves(grads_emu) if p not in zero}
    emu_l2 = {p: _l2(g, ref[p]) for p, g in O.tree_leaves(grads_emu) if p not in zero}
    e_gpu, e_emu = sorted(errs[p] for p in emu), sorted(emu.values())
    l_gpu, l_emu = sorted(l2.values()), sorted(emu_l2.values())
    zero_ratio = {p: float(ours[p].float().abs().max().cpu() / ref[k].abs().max()) for p, k in zero.items()}
    summary = {"loss": loss.item(), "loss_ref": float(loss_ref), "logits": errs["logits"],
               "logits_emu": _rel(logits_emu, logits_ref),
               "max_norm_gpu_worst_median": [e_gpu[-1], e_gpu[len(e_gpu) // 2]],
               "max_norm_emu_worst_median": [e_emu[-1], e_emu[len(e_emu) // 2]],
               "l2_gpu_worst_median": [l_gpu[-1], l_gpu[len(l_gpu) // 2]],
               "l2_emu_worst_median": [l_emu[-1], l_emu[len(l_emu) // 2]],
               "worst": sorted(errs.items(), key=lambda kv: -kv[1])[:10],
               "worst_l2": sorted(l2.items(), key=lambda kv: -kv[1])[:6],
               "worst_emu": sorted(emu.items(), key=lambda kv: -kv[1])[:6],
               "zero_grad_bias_over_kernel_grad": max(zero_ratio.values())}
    print(name, json.dumps(summary, indent=1))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"parity_{name}.json"), "w") as f:
        json.dump(summary, f, indent=1)
    assert abs(loss.item() - float(loss_ref)) <= 2e-2 * abs(float(loss_ref)), (loss.item(), float(loss_ref))
    assert errs["logits"] <= max(TOL, 1.5 * summary["logits_emu"])
    # mathematically-zero gradients stay ~0 on the scale of the layer's kernel gradient
    assert max(zero_ratio.values()) <= 2e-2, sorted(zero_ratio.items(), key=lambda kv: -kv[1])[:3]
    # on the floor of an ideal bf16 pipeline: worst and median tensor, max-norm and relative L2
    assert e_gpu[-1] <= 1.3 * e_emu[-1] and e_gpu[len(e_gpu) // 2] <= max(TOL, 1.5 * e_emu[len(e_emu) // 2])
    assert l_gpu[-1] <= 1.3 * l_emu[-1] and l_gpu[len(l_gpu) // 2] <= max(TOL, 1.5 * l_emu[len(l_emu) // 2])
