"""Development tool (CPU): where does the bf16 path lose accuracy against the fp32 oracle?

Emulates the CUDA path's rounding points on top of the oracle's formulas (round-to-bf16 in forward
and of the gradient in backward) and reports the worst per-tensor max|d|/max|ref| with each group
of rounding points switched on/off.  Used to decide the residual-stream precision.

    python tests/tools/bf16_error_budget.py [--model deit_ti_patch16] [--batch 8] [--img 224]
"""
import argparse
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import vit_oracle as O  # noqa: E402


class _R(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, fwd, bwd):
        ctx.bwd = bwd
        return x.bfloat16().float() if fwd else x

    @staticmethod
    def backward(ctx, g):
        return (g.bfloat16().float() if ctx.bwd else g), None, None


class _AttnCore(torch.autograd.Function):
    """Mimics csrc/attn.cu: bf16 P for PV and dV, bf16 dS for dK/dQ, fp32 softmax statistics;
    delta = rowsum(dO * O) from the bf16-rounded O (mode 'o16'), from O kept in fp32 ('o32'), or
    the exact rowsum(P * dP) ('exact')."""

    @staticmethod
    def forward(ctx, q, k, v, scale, delta_mode):
        s = torch.einsum("bqhd,bkhd->bhqk", q, k) * scale
        p = O.softmax_last(s)
        o32 = torch.einsum("bhqk,bkhd->bqhd", p.bfloat16().float(), v)
        o = o32.bfloat16().float()
        ctx.save_for_backward(q, k, v, p, o, o32)
        ctx.scale, ctx.mode = scale, delta_mode
        return o

    @staticmethod
    def backward(ctx, d_o):
        q, k, v, p, o, o32 = ctx.saved_tensors
        d_o = d_o.bfloat16().float()
        dp = torch.einsum("bqhd,bkhd->bhqk", d_o, v)
        if ctx.mode == "o16":
            delta = (d_o * o).sum(-1)
        elif ctx.mode == "o32":
            delta = (d_o * o32).sum(-1)
        else:
            delta = None
        if delta is None:
            dl = (p * dp).sum(-1, keepdim=True)
        else:
            dl = delta.permute(0, 2, 1).unsqueeze(-1)
        ds = (p * (dp - dl) * ctx.scale).bfloat16().float()
        dv = torch.einsum("bhqk,bqhd->bkhd", p.bfloat16().float(), d_o)
        dk = torch.einsum("bhqk,bqhd->bkhd", ds, q)
        dq = torch.einsum("bhqk,bkhd->bqhd", ds, k)
        return dq.bfloat16().float(), dk.bfloat16().float(), dv.bfloat16().float(), None, None


ATTN_MODE = [None]
SITES = {}  # site name -> (round forward value, round gradient); missing = follow the group switch


def R(x, on=True, bwd=None, site=None):
    bwd = on if bwd is None else bwd
    if site in SITES:
        on, bwd = SITES[site]
    return _R.apply(x, on, bwd)


def forward(params, images_nhwc, cfg, act=True, stream=True, weights=True, attn=True):
    """act: non-residual activations (LN out, qkv, ctx, hidden) bf16; stream: residual stream x and
    its gradient bf16; weights: bf16 weight shadows; attn: P / dS rounded to bf16 inside the core."""
    Wt = (lambda w: w.bfloat16().float()) if weights else (lambda w: w)
    ph, pw = cfg["patch_shape"]
    H = cfg["num_heads"]
    b, Hh, Ww, c = images_nhwc.shape
    gh, gw = Hh // ph, Ww // pw
    x = images_nhwc.reshape(b, gh, ph, gw, pw, c).permute(0, 1, 3, 2, 4, 5).reshape(b, gh * gw, -1)
    x = R(x, act, False, site="img")
    tok = R(x @ Wt(params["PatchEmbedBlock_0"]["Dense_0"]["kernel"]), act, site="tok")
    enc = params["Encoder_0"]
    x = torch.cat([params["cls"].expand(b, -1, -1), tok], 1) + enc["AddAbsPosEmbed_0"]["pos_embed"]
    x = R(x, stream, site="stream")
    for i in range(cfg["num_layers"]):
        p = enc[f"EncoderBlock_{i}"]
        y = R(O.layer_norm(x, p["LayerNorm_0"]["scale"], p["LayerNorm_0"]["bias"]), act, site="ln")
        a = p["SelfAttentionBlock_0"]
        d = a["queries"]["kernel"].shape[-1]
        q = R(torch.einsum("bqc,chd->bqhd", y, Wt(a["queries"]["kernel"])), act, site="qkv")
        k = R(torch.einsum("bkc,chd->bkhd", y, Wt(a["keys"]["kernel"])), act, site="qkv")
        v = R(torch.einsum("bkc,chd->bkhd", y, Wt(a["values"]["kernel"])), act, site="qkv")
        if ATTN_MODE[0]:
            o = _AttnCore.apply(q, k, v, 1.0 / math.sqrt(d), ATTN_MODE[0])
        else:
            s = R(torch.einsum("bqhd,bkhd->bhqk", q, k) / math.sqrt(d), False, attn)
            pr = R(O.softmax_last(s), attn, False)
            o = R(torch.einsum("bhqk,bkhd->bqhd", pr, v), act, site="ctx")
        y = torch.einsum("bqhd,hdc->bqc", o, Wt(a["DenseGeneral_0"]["kernel"]))
        x = R(y + x, stream, site="stream")
        y = R(O.layer_norm(x, p["LayerNorm_1"]["scale"], p["LayerNorm_1"]["bias"]), act, site="ln")
        f = p["FFBlock_0"]
        h = R(y @ Wt(f["Dense_0"]["kernel"]) + f["Dense_0"]["bias"], act, site="pre")
        h = R(O.gelu_tanh(h), act, site="hid")
        y = h @ Wt(f["Dense_1"]["kernel"]) + f["Dense_1"]["bias"]
        x = R(x + y, stream, site="stream")
    x = R(O.layer_norm(x, enc["LayerNorm_0"]["scale"], enc["LayerNorm_0"]["bias"]), act, site="ln_f")
    return R(x[:, 0] @ Wt(params["Dense_0"]["kernel"]) + params["Dense_0"]["bias"], act, site="logits")


def run(cfg, params, images, labels, **kw):
    leaves = []
    p = O.tree_map(lambda t: (leaves.append(t.detach().clone().requires_grad_(True)) or leaves[-1]),
                   params)
    logits = forward(p, O.hwcn_to_nhwc(images), cfg, **kw)
    loss = O.loss_fn(logits, labels, cfg["num_classes"], 0.1)
    loss.backward()
    return logits.detach(), O.tree_map(lambda t: t.grad, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="deit_ti_patch16")
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--img", type=int, default=224)
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    cfg = O.vit_config(a.model)
    params = O.synthetic_vit_params(cfg, a.img, seed=42)
    images, labels = O.synthetic_batch(a.batch, a.img, seed=0)
    _, logits_ref, grads_ref = O.vit_loss_and_grads(params, images, labels, cfg, 0.1)
    ref = dict(O.tree_leaves(grads_ref))
    rel = lambda x, y: float((x - y).abs().max() / y.abs().max().clamp_min(1e-20))
    cases = {
        "all bf16 (current CUDA path)": dict(),
        "fp32 residual stream": dict(stream=False),
        "fp32 stream + fp32 weights": dict(stream=False, weights=False),
        "fp32 stream, attn P/dS fp32": dict(stream=False, attn=False),
        "only stream bf16": dict(act=False, weights=False, attn=False),
        "only weights bf16": dict(act=False, stream=False, attn=False),
        "only act bf16": dict(weights=False, stream=False, attn=False),
    }
    site_cases = {
        "fp32 stream; ln grad fp32": {"ln": (True, False)},
        "fp32 stream; qkv grad fp32": {"qkv": (True, False)},
        "fp32 stream; ctx grad fp32": {"ctx": (True, False)},
        "fp32 stream; pre fp32 both": {"pre": (False, False)},
        "fp32 stream; hid grad fp32": {"hid": (True, False)},
        "fp32 stream; all act grads fp32": {k: (True, False) for k in ("ln", "qkv", "ctx", "pre", "hid", "tok")},
        "fp32 stream; all act fwd fp32": {k: (False, True) for k in ("ln", "qkv", "ctx", "pre", "hid", "tok")},
    }
    for name, sites in site_cases.items():
        cases[name] = dict(stream=False, _sites=sites)
    cases["stream: fwd fp32, grad bf16"] = dict(_sites={"stream": (False, True)})
    cases["stream: fwd bf16, grad fp32"] = dict(_sites={"stream": (True, False)})
    for m in ("o16", "o32", "exact"):
        cases[f"fp32 stream, kernel-like attn, delta {m}"] = dict(stream=False, _attn=m)
    if a.only:
        cases = {k: v for k, v in cases.items() if a.only in k}
    for name, kw in cases.items():
        kw = dict(kw)
        ATTN_MODE[0] = kw.pop("_attn", None)
        SITES.clear()
        SITES.update(kw.pop("_sites", {}))
        logits, grads = run(cfg, params, images, labels, **kw)
        errs = sorted(((rel(g, ref[k]), k) for k, g in O.tree_leaves(grads)), reverse=True)
        over = sum(e > 2e-2 for e, _ in errs)
        print(f"{name:34s} logits {rel(logits, logits_ref):.2e}  worst grad {errs[0][0]:.2e} "
              f"({errs[0][1]})  median {errs[len(errs) // 2][0]:.2e}  >2e-2: {over}", flush=True)


if __name__ == "__main__":
    main()
