"""Build libsav_b200.so (all hand-written sm_100a kernels + the C ABI) in-tree with nvcc.

Usage: python build.py [--force].  Objects are rebuilt only when their sources are newer.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libsav_b200.so")
SOURCES = ["host.cu", "gemm.cu", "attn.cu", "attn_fwd.cu", "attn_bwd.cu", "norm.cu", "stem.cu", "optim.cu",
           "attn_mat.cu", "elementwise.cu", "conv.cu", "conv2d.cu", "cnn.cu", "dp.cu", "attn_small.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default"]


def _newer(src_list, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in src_list)


def build(force=False, verbose=False):
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(HERE, "..", "include", "sav_abi.h"))
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    jobs = []
    objs = []
    # SAV_GEMM_DEV=1: compile the GEMM's development instrumentation in (SAV_GEMM_DEBUG switches and
    # the role timeline used by tests/tools/gemm_*.sh); the default build carries none of it
    dev = os.environ.get("SAV_GEMM_DEV", "0") == "1"
    marker = os.path.join(objdir, ".gemm_dev")
    was_dev = os.path.exists(marker)
    # SAV_ATTN_DEV=1: role timeline of the attention backward (tests/tools/attn_timeline.py --bwd)
    adev = os.environ.get("SAV_ATTN_DEV", "0") == "1"
    amarker = os.path.join(objdir, ".attn_dev")
    was_adev = os.path.exists(amarker)
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(objdir, s.replace(".cu", ".o"))
        objs.append(obj)
        extra = ["-DSAV_GEMM_DEV"] if (dev and s == "gemm.cu") else []
        if adev and s == "attn_bwd.cu":
            extra = ["-DSAV_ATTN_DEV"]
        if force or _newer([src] + headers, obj) or (s == "gemm.cu" and dev != was_dev) or \
                (s == "attn_bwd.cu" and adev != was_adev):
            jobs.append([NVCC] + FLAGS + extra + ["-c", src, "-o", obj])
    if adev:
        open(amarker, "w").close()
    elif was_adev:
        os.remove(amarker)
    if dev:
        open(marker, "w").close()
    elif was_dev:
        os.remove(marker)

    def run(cmd):
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(run, jobs))
    # the guarded XLA-FFI shim must keep preprocessing (it is an empty translation unit without
    # jaxlib's headers; with them it is built separately, see its header comment)
    shim = os.path.join(CSRC, "ffi", "sav_xla_ffi.cc")
    if os.path.exists(shim):
        run(["g++", "-std=c++17", "-fsyntax-only", "-I", os.path.join(HERE, "..", "include"), shim])
    if jobs or not os.path.exists(OUT):
        run([NVCC, "-shared", "-o", OUT] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"])
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
