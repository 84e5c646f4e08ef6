# Needs a dev build: SAV_GEMM_DEV=1 python self-attention-experiments-vision_b200/build.py
# Where does the GEMM epilogue time go?  SAV_GEMM_DEBUG bits: 1 no TMA loads, 2 no MMAs,
# 4 no GELU math, 8 no TMEM read, 16 no global stores.
export SAV_TIME_ONLY=t_fc1_fwd,t_qkv_fwd,t_fc2_fwd,t_fc1_dx
for d in 0 3 7 11 19 27 31; do echo "== debug=$d"; SAV_GEMM_DEBUG=$d timeout 100 python tests/gpu_diag.py --time; done
