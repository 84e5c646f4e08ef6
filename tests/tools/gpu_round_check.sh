set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r01c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r01c_pytest.log
python tests/gpu_diag.py --time > gpurun_out/r01c_time.log 2>&1
python bench.py --steps 10 --warmup 3 > gpurun_out/r01c_bench.json 2> gpurun_out/r01c_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01c_ref.json 2> gpurun_out/r01c_ref.err
NCU="ncu --set full --clock-control none --import-source on"
timeout 300 $NCU -k regex:gemm_kernel --launch-skip 3 -c 1 -f -o gpurun_out/r01c_fc1_fwd python tests/gpu_diag.py --timecase t_fc1_fwd > gpurun_out/ncu1.log 2>&1
timeout 300 $NCU -k regex:attn_fwd --launch-skip 3 -c 1 -f -o gpurun_out/r01c_attn_fwd python tests/gpu_diag.py --timecase t_attn_fwd > gpurun_out/ncu2.log 2>&1
timeout 300 $NCU -k regex:attn_bwd2 --launch-skip 3 -c 1 -f -o gpurun_out/r01c_attn_bwd python tests/gpu_diag.py --timecase t_attn_bwd > gpurun_out/ncu3.log 2>&1
timeout 300 $NCU -k regex:gemm_kernel --launch-skip 3 -c 1 -f -o gpurun_out/r01c_fc2_fwd python tests/gpu_diag.py --timecase t_fc2_fwd > gpurun_out/ncu4.log 2>&1
timeout 300 $NCU -k regex:gemm_kernel --launch-skip 3 -c 1 -f -o gpurun_out/r01c_fc1_dw python tests/gpu_diag.py --timecase t_fc1_dw > gpurun_out/ncu5.log 2>&1
ls -la gpurun_out
