"""GPU bring-up diagnostics: each kernel family against a torch-on-GPU fp32 restatement, one
subprocess per case (a hang or fault in one case cannot take the others down).

    python tests/gpu_diag.py            # run all cases, print a table, exit 1 on any failure
    python tests/gpu_diag.py --case gemm_kk_128
This is a development tool; the parity tests proper are tests/test_*_gpu.py (vs the oracle).
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "self-attention-experiments-vision_b200"))


def rel_err(a, b):
    import torch
    a = a.float(); b = b.float()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-12))


def case_gemm(a_mn, b_mn, N, epi="none", M=296, K=192, split_k=1):
    import torch
    from sav import ops
    torch.manual_seed(0)
    dev = "cuda"
    A = torch.randn(M, K, device=dev).bfloat16()
    B = torch.randn(N, K, device=dev).bfloat16()
    ref = A.float() @ B.float().t()
    Ain = A.t().contiguous() if a_mn else A
    Bin = B.t().contiguous() if b_mn else B
    kw = {}
    if epi == "bias_gelu":
        bias = torch.randn(N, device=dev)
        aux = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        kw = dict(bias=bias, act=1, aux_out=aux)
        pre = ref + bias
        ref = torch.nn.functional.gelu(pre, approximate="tanh")
    elif epi == "full":   # fp32 residual-stream epilogue: bias, LayerScale, drop-path, residual
        bias = torch.randn(N, device=dev)
        cs = torch.rand(N, device=dev) + 0.5
        rs = torch.rand((M + 49) // 50, device=dev) + 0.5
        res = torch.randn(M, N, device=dev)
        kw = dict(bias=bias, col_scale=cs, row_scale=rs, rows_per_scale=50, residual=res,
                  out_fp32=True)
        ref = (ref + bias) * cs * rs.repeat_interleave(50)[:M, None] + res
    elif epi == "bias_act":   # single-output GELU (inference form)
        bias = torch.randn(N, device=dev)
        kw = dict(bias=bias, act=1)
        ref = torch.nn.functional.gelu(ref + bias, approximate="tanh")
    elif epi == "accum":
        d0 = torch.randn(M, N, device=dev)
        kw = dict(out=d0.clone(), accumulate=True, split_k=split_k)
        ref = ref + d0
    elif epi == "dgelu":
        h = torch.randn(M, N, device=dev).bfloat16()
        kw = dict(aux_in=h)
        hf = h.float().requires_grad_(True)
        g, = torch.autograd.grad(torch.nn.functional.gelu(hf, approximate="tanh").sum(), hf)
        ref = ref * g
    elif epi == "fp32":
        kw = dict(out_fp32=True, split_k=split_k)
    out = ops.gemm(Ain, Bin, a_mn=a_mn, b_mn=b_mn, **kw)
    torch.cuda.synchronize()
    errs = {"out": rel_err(out, ref)}
    if epi == "bias_gelu":
        errs["aux"] = rel_err(kw["aux_out"], pre)
    return errs


def case_ln(strided=False):
    import torch
    from sav import ops
    torch.manual_seed(0)
    rows, C = 1000, 384
    pitch = 3 * C if strided else C
    xbuf = torch.randn(rows, pitch, device="cuda") * 2 + 0.5
    x = xbuf[:, :C]
    g = torch.randn(C, device="cuda"); b = torch.randn(C, device="cuda")
    y, mean, rstd = ops.layernorm_fwd(x, g, b)
    xf = x.detach().clone().requires_grad_(True)
    gf = g.clone().requires_grad_(True); bf = b.clone().requires_grad_(True)
    ref = torch.nn.functional.layer_norm(xf, (C,), gf, bf, eps=1e-6)
    dy = torch.randn(rows, C, device="cuda").bfloat16()
    add = torch.randn(rows, pitch, device="cuda")[:, :C]
    ref.backward(dy.float())
    ds = torch.zeros(C, device="cuda"); db = torch.zeros(C, device="cuda")
    dxb = torch.zeros(rows, pitch, device="cuda")
    dx16b = torch.zeros(rows, pitch, device="cuda", dtype=torch.bfloat16)
    dx, dx16 = ops.layernorm_bwd(dy, x, g, mean, rstd, ds, db, dx_add=add, dx=dxb[:, :C],
                                 dx16=dx16b[:, :C])
    torch.cuda.synchronize()
    return {"y": rel_err(y, ref), "dx": rel_err(dx, xf.grad + add), "dx16": rel_err(dx16, xf.grad + add),
            "dscale": rel_err(ds, gf.grad), "dbias": rel_err(db, bf.grad),
            "untouched": float(dxb[:, C:].abs().max()) if strided else 0.0}


def case_gemm_time(M=50432, N=1536, K=384, mode="gelu2", a_mn=False, b_mn=True, reps=20):
    """Times one GEMM shape of the DeiT-S block; returns TFLOP/s and effective GB/s (not an error
    metric: reported through the same table with tol ignored)."""
    import torch
    from sav import ops
    dev = "cuda"
    A = torch.randn((K, M) if a_mn else (M, K), device=dev).bfloat16()
    B = torch.randn((K, N) if b_mn else (N, K), device=dev).bfloat16()
    kw = {}
    byt = (M * K + N * K) * 2
    if mode == "gelu2":
        kw = dict(bias=torch.randn(N, device=dev), act=1,
                  aux_out=torch.empty(M, N, device=dev, dtype=torch.bfloat16)); byt += 4 * M * N
    elif mode == "plain":
        byt += 2 * M * N
    elif mode in ("dgelu", "dgelu_cs"):
        kw = dict(aux_in=torch.randn(M, N, device=dev).bfloat16()); byt += 4 * M * N
        if mode == "dgelu_cs":     # + the bias column sums accumulated by the epilogue
            kw["dgelu_colsum"] = torch.zeros(N, device=dev)
    elif mode in ("resid", "resid_ln"):
        kw = dict(bias=torch.randn(N, device=dev), residual=torch.randn(M, N, device=dev),
                  out_fp32=True); byt += 8 * M * N
        if mode == "resid_ln":     # + the LayerNorm of the result (bf16 operand, mean, rstd)
            kw["ln"] = (torch.ones(N, device=dev), torch.zeros(N, device=dev), 1e-6); byt += 2 * M * N
    elif mode == "dw":
        kw = dict(out=torch.zeros(M, N, device=dev), accumulate=True, split_k=0); byt += 4 * M * N
    out = kw.pop("out", None)
    if out is None:
        out = torch.empty(M, N, device=dev, dtype=torch.float32 if kw.get("out_fp32") else torch.bfloat16)
    kw.pop("out_fp32", None)
    for _ in range(3):
        ops.gemm(A, B, a_mn=a_mn, b_mn=b_mn, out=out, **kw)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        ops.gemm(A, B, a_mn=a_mn, b_mn=b_mn, out=out, **kw)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    return {"us": ms * 1e3, "tflops": 2.0 * M * N * K / ms / 1e9, "gbs": byt / ms / 1e6}


def case_attn(B=2, H=3, Nq=197, Nk=197, bwd=True):
    import torch
    from sav import ops
    torch.manual_seed(0)
    D = 64
    C = H * D
    qkv = torch.randn(B, max(Nq, Nk), 3, H, D, device="cuda").bfloat16()
    q, k, v = qkv[:, :Nq, 0], qkv[:, :Nk, 1], qkv[:, :Nk, 2]
    o, lse = ops.attn_fwd(q, k, v)
    qf, kf, vf = (t.float().permute(0, 2, 1, 3).detach().requires_grad_(True) for t in (q, k, v))
    s = (qf / 8.0) @ kf.transpose(-1, -2)
    pr = torch.softmax(s, -1)
    ref = (pr @ vf).permute(0, 2, 1, 3)
    o2 = o
    o = o2[0]
    errs = {"o": rel_err(o, ref), "o_lo": rel_err(o2[0].float() + o2[1].float(), ref),
            "lse": rel_err(lse, torch.logsumexp(s, -1))}
    if bwd:
        d_o = torch.randn(B, Nq, H, D, device="cuda").bfloat16()
        ref.backward(d_o.float())
        dqkv = torch.zeros_like(qkv)
        ops.attn_bwd(q, k, v, o2, d_o, lse, dqkv[:, :Nq, 0], dqkv[:, :Nk, 1], dqkv[:, :Nk, 2])
        torch.cuda.synchronize()
        errs["dq"] = rel_err(dqkv[:, :Nq, 0], qf.grad.permute(0, 2, 1, 3))
        errs["dk"] = rel_err(dqkv[:, :Nk, 1], kf.grad.permute(0, 2, 1, 3))
        errs["dv"] = rel_err(dqkv[:, :Nk, 2], vf.grad.permute(0, 2, 1, 3))
    torch.cuda.synchronize()
    return errs


def case_stem():
    import torch
    from sav import ops
    torch.manual_seed(0)
    N, H, W, Cc, p = 70, 32, 48, 3, 16
    img = torch.randn(H, W, Cc, N, device="cuda")
    out = ops.patchify(img, p, p, hwcn=True)
    nhwc = img.permute(3, 0, 1, 2).contiguous()
    ref = nhwc.reshape(N, H // p, p, W // p, p, Cc).permute(0, 1, 3, 2, 4, 5).reshape(-1, p * p * Cc)
    out2 = ops.patchify(nhwc, p, p, hwcn=False)
    torch.cuda.synchronize()
    return {"hwcn": rel_err(out, ref.bfloat16()), "nhwc": rel_err(out2, ref.bfloat16())}


CASES = {}
for am in (0, 1):
    for bm in (0, 1):
        for n in (384, 256):
            tag = f"gemm_{'m' if am else 'k'}{'m' if bm else 'k'}_{128 if n % 256 else 256}"
            CASES[tag] = (case_gemm, dict(a_mn=bool(am), b_mn=bool(bm), N=n))
CASES["gemm_bias_gelu"] = (case_gemm, dict(a_mn=False, b_mn=True, N=384, epi="bias_gelu"))
CASES["gemm_full_epi"] = (case_gemm, dict(a_mn=False, b_mn=True, N=384, epi="full"))
CASES["gemm_dgelu"] = (case_gemm, dict(a_mn=False, b_mn=False, N=384, epi="dgelu"))
CASES["gemm_bias_act"] = (case_gemm, dict(a_mn=False, b_mn=True, N=384, epi="bias_act"))
CASES["gemm_ntail"] = (case_gemm, dict(a_mn=False, b_mn=True, N=1000, M=256, K=384, epi="bias_act"))
CASES["gemm_ktail"] = (case_gemm, dict(a_mn=False, b_mn=False, N=384, M=256, K=1000))
CASES["gemm_accum"] = (case_gemm, dict(a_mn=True, b_mn=True, N=384, epi="accum", M=192, K=3000, split_k=0))
CASES["gemm_fp32"] = (case_gemm, dict(a_mn=True, b_mn=True, N=384, epi="fp32", M=192, K=3000))
CASES["gemm_splitk"] = (case_gemm, dict(a_mn=True, b_mn=True, N=384, epi="fp32", M=192, K=3000, split_k=5))
CASES["gemm_big"] = (case_gemm, dict(a_mn=False, b_mn=True, N=1152, M=4000, K=384))
CASES["gemm_pair_oob"] = (case_gemm, dict(a_mn=False, b_mn=True, N=192, M=1568, K=192))
CASES["gemm_pair_oob2"] = (case_gemm, dict(a_mn=False, b_mn=False, N=192, M=1568, K=192))


def case_conv(B=2, H=28, W=28, C=192, stride=2):
    import torch
    from sav import ops
    torch.manual_seed(0)
    x = torch.randn(B, H, W, C, device="cuda")
    w = torch.randn(3, 3, C, device="cuda") / 3
    y, stats = ops.dwconv3x3_fwd(x, w.reshape(9, C), stride)
    Ho, Wo = y.shape[1], y.shape[2]
    pt = max((Ho - 1) * stride + 3 - H, 0)
    pl = max((Wo - 1) * stride + 3 - W, 0)
    xi = torch.nn.functional.pad(x.permute(0, 3, 1, 2), (pl // 2, pl - pl // 2, pt // 2, pt - pt // 2))
    xi.requires_grad_(True)
    wk = w.permute(2, 0, 1)[:, None].contiguous().requires_grad_(True)
    ref = torch.nn.functional.conv2d(xi, wk, stride=stride, groups=C).permute(0, 2, 3, 1)
    errs = {"y": rel_err(y, ref), "sum": rel_err(stats[0], ref.sum((0, 1, 2))),
            "sumsq": rel_err(stats[1], (ref * ref).sum((0, 1, 2)))}
    dy = torch.randn_like(ref).bfloat16()
    (ref * dy.float()).sum().backward()
    dx = torch.empty_like(x)
    dw = torch.zeros(9, C, device="cuda")
    ops.dwconv3x3_bwd(x, w.reshape(9, C), dy.contiguous(), dx, dw, stride, False)
    gx = xi.grad[:, :, pt // 2:pt // 2 + H, pl // 2:pl // 2 + W].permute(0, 2, 3, 1)
    errs["dx"] = rel_err(dx, gx)
    errs["dw"] = rel_err(dw, wk.grad[:, 0].permute(1, 2, 0).reshape(9, C))
    torch.cuda.synchronize()
    return errs


CASES["conv_192_s2"] = (case_conv, {})
CASES["conv_192_s1"] = (case_conv, dict(stride=1))
CASES["conv_64_s2"] = (case_conv, dict(B=1, H=56, W=56, C=64))
CASES["conv_384_odd"] = (case_conv, dict(B=2, H=15, W=15, C=384))
CASES["ln"] = (case_ln, {})
CASES["ln_strided"] = (case_ln, dict(strided=True))
CASES["stem"] = (case_stem, {})
CASES["attn_fwd_197"] = (case_attn, dict(bwd=False))
CASES["attn_fwd_64"] = (case_attn, dict(Nq=64, Nk=64, bwd=False))
CASES["attn_197"] = (case_attn, {})
CASES["attn_577"] = (case_attn, dict(B=1, H=2, Nq=577, Nk=577))
CASES["attn_cross"] = (case_attn, dict(B=2, H=2, Nq=225, Nk=64))
CASES["attn_cvt2_fwd"] = (case_attn, dict(B=2, H=3, Nq=784, Nk=196, bwd=False))
CASES["attn_cvt2"] = (case_attn, dict(B=2, H=3, Nq=784, Nk=196))
CASES["attn_cvt1_fwd"] = (case_attn, dict(B=1, H=1, Nq=3136, Nk=784, bwd=False))
CASES["attn_cvt1"] = (case_attn, dict(B=1, H=1, Nq=3136, Nk=784))
CASES["attn_q300_k100"] = (case_attn, dict(B=1, H=1, Nq=300, Nk=100))


def case_attn_time(B=256, H=6, N=197, bwd=False, reps=20):
    import torch
    from sav import ops
    D = 64
    qkv = torch.randn(B, N, 3, H, D, device="cuda").bfloat16()
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    o2, lse = ops.attn_fwd(q, k, v)
    d_o = torch.randn(B, N, H, D, device="cuda").bfloat16()
    dqkv = torch.empty_like(qkv)
    if bwd:
        fn = lambda: ops.attn_bwd(q, k, v, o2, d_o, lse, dqkv[:, :, 0], dqkv[:, :, 1], dqkv[:, :, 2])
        flops, byt = 8.0 * B * H * N * N * D, 9 * B * N * H * D * 2
    else:
        fn = lambda: ops.attn_fwd(q, k, v)
        flops, byt = 4.0 * B * H * N * N * D, 5 * B * N * H * D * 2
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    return {"us": ms * 1e3, "tflops": flops / ms / 1e9, "gbs": byt / ms / 1e6}


TIMING = {
    "t_attn_fwd": dict(_attn=True),
    "t_attn_bwd": dict(_attn=True, bwd=True),
    "t_attn_fwd577": dict(_attn=True, B=64, H=12, N=577),
    "t_attn_bwd577": dict(_attn=True, B=64, H=12, N=577, bwd=True),
    "t_qkv_fwd": dict(N=1152, K=384, mode="plain"),
    "t_out_fwd": dict(N=384, K=384, mode="resid"),
    "t_fc1_fwd": dict(N=1536, K=384, mode="gelu2"),
    "t_fc2_fwd": dict(N=384, K=1536, mode="resid"),
    "t_out_fwd_ln": dict(N=384, K=384, mode="resid_ln"),
    "t_fc2_fwd_ln": dict(N=384, K=1536, mode="resid_ln"),
    "t_fc2_dx": dict(N=1536, K=384, mode="dgelu", b_mn=False),
    "t_fc2_dx_cs": dict(N=1536, K=384, mode="dgelu_cs", b_mn=False),
    "t_fc1_dx": dict(N=384, K=1536, mode="plain", b_mn=False),
    "t_qkv_dx": dict(N=384, K=1152, mode="plain", b_mn=False),
    "t_fc1_dw": dict(M=384, N=1536, K=50432, mode="dw", a_mn=True),
    "t_fc2_dw": dict(M=1536, N=384, K=50432, mode="dw", a_mn=True),
    "t_qkv_dw": dict(M=384, N=1152, K=50432, mode="dw", a_mn=True),
    "t_out_dw": dict(M=384, N=384, K=50432, mode="dw", a_mn=True),
    "t_sq_kk": dict(M=8192, N=8192, K=8192, mode="plain", b_mn=False),
    "t_sq_km": dict(M=8192, N=8192, K=8192, mode="plain", b_mn=True),
}


def main():
    if "--time" in sys.argv:
        only = os.environ.get("SAV_TIME_ONLY", "")
        for name, kw in TIMING.items():
            if only and not any(o in name for o in only.split(",")):
                continue
            try:
                r = subprocess.run([sys.executable, __file__, "--timecase", name], capture_output=True,
                                   text=True, timeout=120)
                line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
                print(f"{name:12s} {line[-1][7:] if line else (r.stderr or r.stdout).strip().splitlines()[-2:]}", flush=True)
            except subprocess.TimeoutExpired:
                print(f"{name:12s} HANG", flush=True)
        return 0
    if "--timecase" in sys.argv:
        kw = dict(TIMING[sys.argv[sys.argv.index("--timecase") + 1]])
        fn = case_attn_time if kw.pop("_attn", False) else case_gemm_time
        print("RESULT " + json.dumps(fn(**kw)))
        return 0
    ap = argparse.ArgumentParser()
    ap.add_argument("--case")
    ap.add_argument("--tol", type=float, default=2e-2)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    if args.case:
        fn, kw = CASES[args.case]
        print("RESULT " + json.dumps(fn(**kw)))
        return 0
    bad = 0
    results = {}
    for name in CASES:
        if args.only and args.only not in name:
            continue
        try:
            r = subprocess.run([sys.executable, __file__, "--case", name], capture_output=True,
                               text=True, timeout=45)
            line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
            if r.returncode == 0 and line:
                errs = json.loads(line[-1][7:])
                ok = all(e == e and e < args.tol for e in errs.values())
                results[name] = errs
                print(f"{'ok  ' if ok else 'FAIL'} {name:18s} " +
                      " ".join(f"{k}={v:.2e}" for k, v in errs.items()), flush=True)
                bad += not ok
            else:
                bad += 1
                tail = (r.stderr or r.stdout).strip().splitlines()[-3:]
                print(f"FAIL {name:18s} rc={r.returncode} {' | '.join(tail)}", flush=True)
        except subprocess.TimeoutExpired:
            bad += 1
            print(f"HANG {name:18s} (timeout)", flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "diag.json"), "w") as f:
        json.dump(results, f, indent=1)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
