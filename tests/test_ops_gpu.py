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
