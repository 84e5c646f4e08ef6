"""Development tool (GPU, needs a SAV_GEMM_DEV=1 build): role timeline of CTA 0 of the GEMM kernel from clock64 stamps.

    python tests/tools/gemm_timeline.py [--M 50432 --N 384 --K 1536] [--b_mn]
Events: P1 tile start (producer); M10 tile start, M11 accumulator free, M12 first operands landed,
M13 all MMAs of the tile issued; E20 waiting for the accumulator, E21 accumulator full, E22 last
slab of the tile handed to TMA.
"""
import argparse
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "self-attention-experiments-vision_b200"))
from sav import _abi, ops  # noqa: E402

EV = {1: "P tile start", 10: "M tile start", 11: "M acc free", 12: "M first operands", 13: "M issued",
      20: "E wait acc", 21: "E acc full", 22: "E tile stored", 23: "E tmem read", 24: "E slab in smem",
      25: "E proxy fence", 26: "S slab ready", 27: "S slot recycled", 28: "S store issued"}
ROLES = ["producer", "mma", "epilogue", "store"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--M", type=int, default=50432)
    ap.add_argument("--N", type=int, default=384)
    ap.add_argument("--K", type=int, default=1536)
    ap.add_argument("--b_mn", action="store_true")
    ap.add_argument("--lines", type=int, default=80)
    ap.add_argument("--bias", action="store_true")
    ap.add_argument("--gelu2", action="store_true", help="FC1 forward epilogue: bias, pre-activation + GELU outputs")
    ap.add_argument("--role", default="")
    a = ap.parse_args()
    A = torch.randn(a.M, a.K, device="cuda").bfloat16()
    B = torch.randn((a.K, a.N) if a.b_mn else (a.N, a.K), device="cuda").bfloat16()
    out = torch.empty(a.M, a.N, device="cuda", dtype=torch.bfloat16)
    kw = {}
    if a.bias or a.gelu2:
        kw["bias"] = torch.randn(a.N, device="cuda")
    if a.gelu2:
        kw["act"] = 1
        kw["aux_out"] = torch.empty_like(out)
    for _ in range(3):
        ops.gemm(A, B, b_mn=a.b_mn, out=out, **kw)
    buf = torch.zeros(8 * 512, dtype=torch.int64, device="cuda")
    lib = ctypes.CDLL(_abi.LIB_PATH)
    lib.sav_set_debug_buffer.argtypes = [ctypes.c_void_p]
    lib.sav_set_debug_buffer(ctypes.c_void_p(buf.data_ptr()))
    ops.gemm(A, B, b_mn=a.b_mn, out=out, **kw)
    torch.cuda.synchronize()
    lib.sav_set_debug_buffer(ctypes.c_void_p(0))
    t = buf.cpu().view(8, 256, 2)
    t0 = min(int(t[r, 0, 1]) for r in range(len(ROLES)) if int(t[r, 0, 0]) != 0)
    events = []
    for r, name in enumerate(ROLES):
        for i in range(256):
            ev, clk = int(t[r, i, 0]), int(t[r, i, 1])
            if ev == 0:
                break
            events.append((clk - t0, name, EV.get(ev, str(ev))))
    events.sort()
    last = {}
    if a.role:
        events = [e for e in events if e[1] == a.role]
    for clk, name, ev in events[: a.lines]:
        d = clk - last.get(name, clk)
        last[name] = clk
        print(f"{clk:9d}  (+{d:6d})  {name:9s} {ev}")


if __name__ == "__main__":
    main()
