# Round-end check of the committed state: parity tests, smoke(), bench, reference arm, launch list.
T=${TAG:-final}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${T}_smoke.log
python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_ref.json 2> gpurun_out/${T}_ref.err
timeout 300 python tests/gpu_diag.py --time > gpurun_out/${T}_time.log 2>&1
timeout 200 python tests/tools/membound_time.py > gpurun_out/${T}_membound.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/${T}_launches_bench_nograph.csv python bench.py --steps 2 --warmup 1 --no-graph --no-cpu-baseline > gpurun_out/${T}_ncu_list.log 2>&1
tail -2 gpurun_out/${T}_pytest.log; tail -1 gpurun_out/${T}_smoke.log; cut -c1-240 gpurun_out/${T}_bench.json
