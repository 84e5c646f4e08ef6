# Round-2 check: parity tests, smoke(), bench, isolated kernel times, launch list.
T=${TAG:-r02a}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${T}_smoke.log
python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
timeout 300 python tests/gpu_diag.py --time > gpurun_out/${T}_time.log 2>&1
timeout 200 python tests/tools/membound_time.py > gpurun_out/${T}_membound.log 2>&1
tail -5 gpurun_out/${T}_pytest.log; tail -2 gpurun_out/${T}_smoke.log; cut -c1-400 gpurun_out/${T}_bench.json
