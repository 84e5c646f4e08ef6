timeout 300 python tests/gpu_diag.py --only gemm | grep -v "^ok" | tail -5
SAV_TIME_ONLY=t_qkv_fwd,t_out_fwd,t_fc1_fwd,t_fc2_fwd,t_fc2_dx,t_fc1_dx,t_qkv_dx,t_fc1_dw,t_fc2_dw,t_qkv_dw,t_out_dw,t_sq_kk timeout 200 python tests/gpu_diag.py --time | cut -c1-60
timeout 300 python bench.py --steps 10 --warmup 3 | cut -c1-200
