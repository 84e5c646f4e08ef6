timeout 200 python tests/gpu_diag.py --only gemm | grep -v "^ok" | tail -5
timeout 200 python tests/gpu_diag.py --only attn | grep -v "^ok" | tail -5
timeout 200 python tests/gpu_diag.py --time
timeout 300 python bench.py --steps 10 --warmup 3
