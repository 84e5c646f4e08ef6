timeout 200 python tests/gpu_diag.py --only gemm | tail -22
timeout 200 python tests/gpu_diag.py --time
SAV_GEMM_DEBUG=1 timeout 100 python tests/tools/gemm_timeline.py | grep "M issued\|M first" | head -6
