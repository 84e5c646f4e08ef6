# Needs a dev build: SAV_GEMM_DEV=1 python self-attention-experiments-vision_b200/build.py
# What bounds each block GEMM after the warp-uniform issue fix?  SAV_GEMM_DEBUG bits:
# 1 no TMA loads, 2 no MMAs, 8 no TMEM read, 16 no global stores.
export SAV_TIME_ONLY=t_qkv_fwd,t_out_fwd,t_fc1_fwd,t_fc2_fwd,t_fc2_dx,t_fc1_dx,t_qkv_dx,t_fc1_dw,t_fc2_dw,t_qkv_dw
for d in 0 1 2 16 24 26; do echo "== debug=$d"; SAV_GEMM_DEBUG=$d timeout 100 python tests/gpu_diag.py --time | cut -c1-70; done
