# One gpurun call: parity tests, bench (+ reference arm), launch list, variant timings, ncu captures.
set -x
T=${TAG:-r01d}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_ref.json 2> gpurun_out/${T}_ref.err
timeout 600 python tests/tools/variant_time.py > gpurun_out/${T}_variants.log 2>&1
timeout 300 python tests/gpu_diag.py --time > gpurun_out/${T}_time.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/${T}_launches_bench_nograph.csv python bench.py --steps 2 --warmup 1 --no-graph --no-cpu-baseline > gpurun_out/${T}_ncu_list.log 2>&1
NCU="ncu --set full --clock-control none --import-source on"
SAV_TIME_ONLY=t_fc1_fwd timeout 300 $NCU -k regex:gemm_kernel --launch-skip 3 -c 1 -f -o gpurun_out/${T}_fc1_fwd python tests/gpu_diag.py --time > gpurun_out/ncu1.log 2>&1
SAV_TIME_ONLY=t_attn_fwd timeout 300 $NCU -k regex:attn_fwd2 --launch-skip 3 -c 1 -f -o gpurun_out/${T}_attn_fwd python tests/gpu_diag.py --time > gpurun_out/ncu2.log 2>&1
SAV_TIME_ONLY=t_attn_bwd timeout 300 $NCU -k regex:attn_bwd2 --launch-skip 3 -c 1 -f -o gpurun_out/${T}_attn_bwd python tests/gpu_diag.py --time > gpurun_out/ncu3.log 2>&1
ls -la gpurun_out | tail -30
