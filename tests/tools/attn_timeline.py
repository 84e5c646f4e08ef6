"""Development tool (GPU): phase timeline of the attention kernels for CTA 0, from clock64 stamps.

    python tests/tools/attn_timeline.py [--bwd] [--N 197] [--B 256] [--H 6]
"""
import argparse
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "self-attention-experiments-vision_b200"))
from sav import _abi, ops  # noqa: E402

EV = {1: "P kv_empty wait", 2: "P kv_empty ok", 10: "M kv_full wait", 11: "M kv_full ok", 12: "M s_free ok",
      13: "M S issued", 14: "M p_full ok", 15: "M PV issued", 20: "S s_full wait", 21: "S s_full ok",
      22: "S pass1 done", 40: "M sdp wait st_free", 41: "M st_free ok", 42: "M sdp issued", 43: "M grads wait pds", 44: "M pds_full ok", 45: "M grads issued", 50: "C wait st_full", 51: "C st_full ok", 52: "C math done", 53: "C pds_free ok", 54: "C smem written", 55: "C dkv_full ok", 56: "C dkv stored", 57: "C dq_full ok", 30: "E o_full g0", 31: "E o_full g1", 32: "E stored", 23: "S pass2 done", 24: "S o_full ok", 25: "S O read", 26: "S stored"}
ROLES = ["producer", "mma0", "mma1", "softmax0", "softmax1", "epilogue"]
if "--bwd" in sys.argv:
    # attn_bwd3_kernel (build with SAV_ATTN_DEV=1)
    ROLES = ["producer", "issuerA", "issuerB", "math_w4", "math_w8", "epilogue", "stats"]
    EV = {1: "P wait qdo_empty", 2: "P qdo_empty ok", 3: "P wait kv_empty", 4: "P kv_empty ok",
          10: "A wait kv_full", 11: "A kv_full ok", 14: "A wait qdo_full", 15: "A qdo_full ok",
          12: "A wait s_free", 13: "A s_free ok -> issue S/dP",
          20: "B wait pds_full", 21: "B pds_full ok", 22: "B dkv/dq_free ok -> issue grads",
          35: "M wait st_full", 36: "M st_full ok", 30: "M wait s_full", 31: "M s_full ok", 32: "M math done",
          33: "M pds_free ok", 34: "M chunk written",
          40: "E delta start", 41: "E delta done", 42: "E wait dkv_full", 43: "E dkv_full ok", 44: "E dV dK stored",
          45: "E wait dq_full", 46: "E dq_full ok",
          50: "S wait dl_ready", 51: "S dl_ready ok", 52: "S wait st_empty", 53: "S st_empty ok"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--N", type=int, default=197)
    ap.add_argument("--B", type=int, default=256)
    ap.add_argument("--H", type=int, default=6)
    ap.add_argument("--bwd", action="store_true")
    ap.add_argument("--items", type=int, default=4)
    a = ap.parse_args()
    qkv = torch.randn(a.B, a.N, 3, a.H, 64, device="cuda").bfloat16()
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    buf = torch.zeros(8 * 512, dtype=torch.int64, device="cuda")
    for _ in range(3):
        o2, lse = ops.attn_fwd(q, k, v)
    lib = ctypes.CDLL(_abi.LIB_PATH)
    lib.sav_set_debug_buffer.argtypes = [ctypes.c_void_p]
    lib.sav_set_debug_buffer(ctypes.c_void_p(buf.data_ptr()))
    if a.bwd:
        d_o = torch.randn(a.B, a.N, a.H, 64, device="cuda").bfloat16()
        dqkv = torch.empty_like(qkv)
        ops.attn_bwd(q, k, v, o2, d_o, lse, dqkv[:, :, 0], dqkv[:, :, 1], dqkv[:, :, 2])
    else:
        ops.attn_fwd(q, k, v)
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
    skip = int(os.environ.get("SKIP", "0"))
    for clk, name, ev in events[skip: skip + a.items * 60]:
        d = clk - last.get(name, clk)
        last[name] = clk
        print(f"{clk:9d}  (+{d:6d})  {name:9s} {ev}")


if __name__ == "__main__":
    main()
