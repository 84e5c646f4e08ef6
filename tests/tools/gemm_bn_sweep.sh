export SAV_TIME_ONLY=t_qkv_fwd,t_out_fwd,t_fc2_fwd,t_fc1_dx,t_qkv_dx,t_qkv_dw,t_out_dw,t_fc2_dw
for bn in 0 128 256; do for c in 2 1; do echo "== BN=$bn ctas=$c"; SAV_GEMM_BN=$bn SAV_GEMM_CTAS=$c timeout 150 python tests/gpu_diag.py --time; done; done
