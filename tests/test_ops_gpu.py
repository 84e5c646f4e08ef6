"""Operator-level cases through the C ABI (the case table of tests/gpu_diag.py as pytest cases): every
operand-layout combination of the tcgen05 GEMM, its epilogues (bias + GELU with the saved
pre-activation, the fp32 residual-stream epilogue with LayerScale / drop-path scales, dGELU,
accumulate / split-K / fp32 out), N and K tails, CTA-pair out-of-bounds tiles; LayerNorm on strided
rows; the patchify / embed stem; the depthwise 3x3 convolution of CvT; and the fused attention forward
and backward at 64 / 197 / 577 tokens, cross shapes (Nq != Nk, CvT's 784 x 196 and 3136 x 784) and a
ragged 300 x 100 -- each against fp32 torch on the same bf16 inputs, per tensor max|d| / max|ref| <= 2e-2."""
import importlib.util
import os

import pytest

_spec = importlib.util.spec_from_file_location("sav_gpu_diag", os.path.join(os.path.dirname(__file__), "gpu_diag.py"))
diag = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(diag)

TOL = 2e-2


@pytest.mark.gpu
@pytest.mark.timeout(180)
@pytest.mark.parametrize("name", sorted(diag.CASES))
def test_operator_case(name):
    fn, kw = diag.CASES[name]
    errs = fn(**kw)
    bad = {k: v for k, v in errs.items() if not v <= TOL}
    assert not bad, (name, errs)


@pytest.mark.gpu
def test_unfused_gelu_and_add_fall_back_to_kernels():
    """Compositions the epilogue fusion does not recognise (a gelu or an add whose input is not a
    pending Dense output) run as standalone kernels instead of raising: gelu(x) + y on bf16
    activations and a + b on fp32 stream tensors, forward and reverse mode against fp32 torch."""
    import torch
    from sav import functional as F
    from sav.tape import Tape, Var
    g = torch.Generator().manual_seed(2)
    x = torch.randn((3, 50, 64), generator=g).bfloat16()
    y = torch.randn((3, 50, 64), generator=g).bfloat16()
    cot = torch.randn((3, 50, 64), generator=g).bfloat16()
    xr, yr = x.float().requires_grad_(True), y.float().requires_grad_(True)
    ref = torch.nn.functional.gelu(xr, approximate="tanh") + yr
    (ref * cot.float()).sum().backward()
    with Tape() as tape:
        xv, yv = Var(x.cuda()), Var(y.cuda())
        out = F.gelu(xv) + yv
        out.g = cot.cuda()
        tape.backward()
    torch.cuda.synchronize()
    rel = lambda a, b: float((a.float().cpu() - b).abs().max() / b.abs().max())   # noqa: E731
    assert rel(out.v, ref.detach()) <= 1e-2
    assert rel(xv.g, xr.grad) <= 1e-2 and rel(yv.g, yr.grad) <= 1e-2
    a = torch.randn((2, 7, 32), generator=g)
    b = torch.randn((2, 7, 32), generator=g)
    cot32 = torch.randn((2, 7, 32), generator=g)
    with Tape() as tape:
        av, bv = Var(a.cuda()), Var(b.cuda())
        s = av + bv
        s.g = cot32.cuda()
        tape.backward()
    torch.cuda.synchronize()
    assert torch.equal(s.v.cpu(), a + b)
    assert torch.equal(av.g.cpu(), cot32) and torch.equal(bv.g.cpu(), cot32)
    assert av.g.data_ptr() != bv.g.data_ptr()     # fp32 gradients are accumulated in place downstream
