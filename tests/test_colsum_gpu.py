"""Bias-gradient column sums (reverse mode of the bias adds in ff.py:24,27 and attention.py:29-37,
60-63) through the C ABI: the two-pass entry point `sav_colsum_bf16_ws` and the workspace-free
`sav_colsum_bf16`, against an fp64 CPU sum of the same bf16 values.

Tolerance: the kernels add bf16 values in fp32; with M rows the rounding error of a column sum is
bounded by M * 2^-24 * sum|x|, so the test asserts |d| <= 1e-5 * sum|x| + 1e-6 per column."""
import ctypes

import pytest
import torch

SHAPES = [(1, 8), (7, 24), (256, 1000), (1576, 192), (50432, 384), (1576, 3072), (333, 2056),
          (12608, 768), (5000, 1152), (4097, 1536)]


def _ref(x):
    xd = x.detach().cpu().double()
    return xd.sum(0), xd.abs().sum(0)


@pytest.mark.gpu
@pytest.mark.parametrize("M,N", SHAPES)
def test_colsum_two_pass_matches_fp64_sum(M, N):
    from sav import ops
    g = torch.Generator().manual_seed(M * 131 + N)
    x = torch.randn((M, N), generator=g).bfloat16().cuda()
    ref, mag = _ref(x)
    out = ops.colsum(x)
    assert out.dtype == torch.float32 and out.shape == (N,)
    err = (out.cpu().double() - ref).abs()
    assert bool((err <= 1e-5 * mag + 1e-6).all()), float((err / (mag + 1e-6)).max())
    # accumulate=True adds to what is there (gradient arena semantics)
    base = torch.randn((N,), generator=g).cuda()
    out2 = ops.colsum(x, out=base.clone(), accumulate=True)
    err2 = (out2.cpu().double() - (ref + base.cpu().double())).abs()
    assert bool((err2 <= 1e-5 * mag + 1e-5).all())


@pytest.mark.gpu
def test_colsum_strided_view_and_legacy_entry_point():
    from sav import _abi, ops
    g = torch.Generator().manual_seed(5)
    wide = torch.randn((3000, 1152 + 16), generator=g).bfloat16().cuda()
    x = wide[:, 8:8 + 1152]                       # row pitch 1168, 16-byte aligned start
    ref, mag = _ref(x)
    out = ops.colsum(x)
    assert bool(((out.cpu().double() - ref).abs() <= 1e-5 * mag + 1e-6).all())
    # the workspace-free entry point must agree
    lib = _abi.lib()
    legacy = torch.zeros(1152, device="cuda")
    rc = lib.sav_colsum_bf16(ctypes.c_void_p(x.data_ptr()), 3000, 1152, x.stride(0),
                             ctypes.c_void_p(legacy.data_ptr()), 0,
                             ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    assert bool(((legacy.cpu().double() - ref).abs() <= 1e-5 * mag + 1e-6).all())


@pytest.mark.gpu
def test_colsum_rejects_small_workspace():
    from sav import _abi
    lib = _abi.lib()
    x = torch.zeros((64, 64), dtype=torch.bfloat16, device="cuda")
    out = torch.zeros(64, device="cuda")
    ws = torch.zeros(16, dtype=torch.uint8, device="cuda")
    need = lib.sav_colsum_workspace_bytes(64, 64)
    assert need > 16
    rc = lib.sav_colsum_bf16_ws(ctypes.c_void_p(x.data_ptr()), 64, 64, 64, ctypes.c_void_p(out.data_ptr()), 0,
                                ctypes.c_void_p(ws.data_ptr()), 16,
                                ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K", [(1576, 768, 192), (50432, 1536, 384), (300, 1000, 64), (129, 136, 128)])
def test_dgelu_gemm_epilogue_accumulates_the_bias_gradient(M, N, K):
    """sav_gemm aux_in mode with dgelu_colsum: D = bf16(acc * gelu'(pre)) AND dgelu_colsum[n] +=
    sum_m acc * gelu'(pre) -- the FC1 bias gradient accumulated by the FC2-dX GEMM's epilogue
    (reverse mode of ff.py:24-25), against an fp64 restatement on the same bf16 operands; partial
    M / N tiles and 1-CTA / CTA-pair tile shapes are covered by the shapes."""
    from sav import ops
    g = torch.Generator().manual_seed(M + N + K)
    dy = (torch.randn((M, K), generator=g) * 0.5).bfloat16()
    w = (torch.randn((N, K), generator=g) / K ** 0.5).bfloat16()       # b_mn=False: [N, K]
    pre = torch.randn((M, N), generator=g).bfloat16()
    acc = dy.double() @ w.double().t()
    x = pre.double()
    u = 0.7978845608028654 * (x + 0.044715 * x ** 3)
    t = torch.tanh(u)
    dg = 0.5 * (1 + t) + 0.5 * x * (1 - t * t) * 0.7978845608028654 * (1 + 3 * 0.044715 * x * x)
    ref = acc * dg
    base = torch.randn((N,), generator=g)
    colsum = base.clone().cuda()
    out = ops.gemm(dy.cuda(), w.cuda(), b_mn=False, aux_in=pre.cuda(), dgelu_colsum=colsum)
    plain = ops.gemm(dy.cuda(), w.cuda(), b_mn=False, aux_in=pre.cuda())
    torch.cuda.synchronize()
    assert torch.equal(out, plain)                       # the output itself is untouched
    assert float((out.double().cpu() - ref).abs().max() / ref.abs().max()) <= 1e-2
    want = ref.sum(0) + base.double()
    mag = ref.abs().sum(0)
    err = (colsum.double().cpu() - want).abs()
    # fp32 sums of fp32 products (tanh.approx inside gelu': 2^-11 relative per term)
    assert bool((err <= 2e-3 * mag / (M ** 0.5) + 1e-3 * mag + 1e-5).all()), float((err / (mag + 1e-6)).max())


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K", [(1576, 384, 384), (50432, 384, 1536), (300, 192, 192), (129, 256, 64),
                                   (5000, 512, 128), (788, 128, 384)])
def test_gemm_epilogue_with_fused_layernorm(M, N, K):
    """sav_gemm fp32-stream epilogue (bias + residual, vit.py:24,31) with the LayerNorm that follows
    fused in (vit.py:19,26): D is unchanged, ln_out = bf16(LayerNorm(D)) with flax statistics
    (var = E[x^2] - E[x]^2, eps 1e-6), mean / rstd saved -- N = 1 or 2 tiles, 1-CTA and CTA-pair
    tile shapes, partial M tiles."""
    from sav import ops
    assert ops.gemm_ln_tiles(N) in (1, 2, 3, 4) and ops.gemm_ln_tiles(768) == 0
    g = torch.Generator().manual_seed(M + N + K)
    a = torch.randn((M, K), generator=g).bfloat16()
    w = (torch.randn((K, N), generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn((N,), generator=g)
    res = torch.randn((M, N), generator=g)
    gamma = 1.0 + 0.1 * torch.randn((N,), generator=g)
    beta = 0.1 * torch.randn((N,), generator=g)
    plain = ops.gemm(a.cuda(), w.cuda(), b_mn=True, out_fp32=True, bias=bias.cuda(), residual=res.cuda())
    out, (y, mean, rstd) = ops.gemm(a.cuda(), w.cuda(), b_mn=True, out_fp32=True, bias=bias.cuda(),
                                    residual=res.cuda(), ln=(gamma.cuda(), beta.cuda(), 1e-6))
    torch.cuda.synchronize()
    assert torch.equal(out, plain)                        # the stream itself is bit-identical
    for _ in range(3):                                    # fixed-order statistics: reproducible bit for bit
        out2, (y2, mean2, rstd2) = ops.gemm(a.cuda(), w.cuda(), b_mn=True, out_fp32=True, bias=bias.cuda(),
                                            residual=res.cuda(), ln=(gamma.cuda(), beta.cuda(), 1e-6))
        assert torch.equal(y, y2) and torch.equal(mean, mean2) and torch.equal(rstd, rstd2)
    x = out.double().cpu()
    mu = x.mean(-1, keepdim=True)
    var = (x * x).mean(-1, keepdim=True) - mu * mu
    ref = (x - mu) * torch.rsqrt(var + 1e-6) * gamma.double() + beta.double()
    assert float((mean.double().cpu() - mu[:, 0]).abs().max()) <= 1e-5
    assert float((rstd.double().cpu() * torch.sqrt(var[:, 0] + 1e-6) - 1).abs().max()) <= 1e-4
    assert float((y.double().cpu() - ref).abs().max()) <= 2 ** -8 * float(ref.abs().max()) + 1e-3
