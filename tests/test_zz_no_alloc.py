"""sav_colsum_bf16_static: the column-sum entry point that allocates nothing and takes no workspace
(scratch in a file-scope `static __device__` array inside libsav_b200.so; no counterpart in the
reference -- include/sav_abi.h).  Called through the
C-ABI and compared with the existing workspace entry point sav_colsum_bf16_ws on two calls in a row.

Collected last on purpose; passes alone in a fresh process and
after the other tests: it depends on no state but the library's own scratch array."""
import ctypes as C

import pytest
import torch


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ws_colsum(lib, x, out, accumulate):
    M, N = x.shape
    need = int(lib.sav_colsum_workspace_bytes(M, N))
    ws = torch.empty(max(need, 16), dtype=torch.uint8, device="cuda")
    rc = lib.sav_colsum_bf16_ws(_ptr(x), M, N, x.stride(0), _ptr(out), int(accumulate), _ptr(ws), need,
                                _stream())
    assert rc == 0, lib.sav_last_error().decode()


@pytest.mark.gpu
def test_static_scratch_colsum_matches_workspace_colsum_twice_in_a_row():
    from sav import _abi
    lib = _abi.lib()
    cap = int(lib.sav_colsum_static_max_bytes())
    assert cap == 4 << 20
    g = torch.Generator().manual_seed(7)
    # call 1: the FC1 bias-gradient shape of the benchmarked step (50 432 x 1536, 1.8 MiB of partials);
    # call 2: a smaller, ragged shape read through a wider pitch, accumulating into a non-zero output
    # -- whatever call 1 left in the scratch array must not leak into it
    x1 = torch.randn((50432, 1536), generator=g).bfloat16().cuda()
    wide = torch.randn((1571, 448), generator=g).bfloat16().cuda()
    x2 = wide[:, :384]
    seed2 = torch.randn((384,), generator=g).cuda()
    want1 = torch.empty(1536, device="cuda")
    want2 = seed2.clone()
    _ws_colsum(lib, x1, want1, False)
    _ws_colsum(lib, x2, want2, True)
    got1 = torch.full((1536,), float("nan"), device="cuda")
    got2 = seed2.clone()
    torch.cuda.synchronize()
    assert int(lib.sav_colsum_workspace_bytes(*x1.shape)) <= cap
    before = torch.cuda.memory_allocated()
    for x, out, acc in ((x1, got1, False), (x2, got2, True)):
        rc = lib.sav_colsum_bf16_static(_ptr(x), x.shape[0], x.shape[1], x.stride(0), _ptr(out), int(acc),
                                        _stream())
        assert rc == 0, lib.sav_last_error().decode()
    assert torch.cuda.memory_allocated() == before          # nothing came from torch's allocator
    torch.cuda.synchronize()
    # same kernels, same summation order: the tolerance of tests/test_colsum_gpu.py is generous here
    for got, want, x in ((got1, want1, x1), (got2, want2, x2)):
        assert torch.isfinite(got).all()
        scale = float(x.float().abs().sum(0).max())
        assert float((got - want).abs().max()) <= 1e-6 * scale
    assert float((got1.double().cpu() - x1.double().sum(0).cpu()).abs().max()) <= 1e-5 * float(
        x1.float().abs().sum(0).max())
    assert float((got2.double().cpu() - (seed2.double().cpu() + x2.double().sum(0).cpu())).abs().max()) <= \
        1e-5 * float(x2.float().abs().sum(0).max())


@pytest.mark.gpu
def test_static_scratch_colsum_rejects_shapes_beyond_its_array():
    from sav import _abi
    lib = _abi.lib()
    M = N = 8192            # 296 row splits x 8192 columns x 4 B = 9.7 MB of partials > 4 MiB
    assert int(lib.sav_colsum_workspace_bytes(M, N)) > int(lib.sav_colsum_static_max_bytes())
    x = torch.zeros((M, N), dtype=torch.bfloat16, device="cuda")
    out = torch.zeros(N, device="cuda")
    rc = lib.sav_colsum_bf16_static(_ptr(x), M, N, N, _ptr(out), 0, _stream())
    assert rc == -1                      # SAV_ERR_SHAPE (include/sav_abi.h)
    msg = lib.sav_last_error().decode()
    assert "sav_colsum_bf16_static" in msg and "sav_colsum_bf16_ws" in msg
    torch.cuda.synchronize()
