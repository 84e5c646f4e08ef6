# Needs a dev build: SAV_GEMM_DEV=1 python self-attention-experiments-vision_b200/build.py
# Which part of the empty GEMM skeleton costs time?  (development; bits 32 = no proxy fence,
# 64 = no per-slab named barrier -- timing only, results are garbage)
export SAV_TIME_ONLY=t_fc1_fwd,t_qkv_fwd,t_fc1_dx
for d in 31 63 95 127; do echo "== debug=$d"; SAV_GEMM_DEBUG=$d timeout 100 python tests/gpu_diag.py --time; done
for c in 1 2; do echo "== ctas=$c debug=31"; SAV_GEMM_CTAS=$c SAV_GEMM_DEBUG=31 timeout 100 python tests/gpu_diag.py --time; done
