#!/bin/bash
# A/B of the weight-gradient side stream (SAV_DW_STREAM): parity of the ViT step with it on, then the
# bench line with it off / on, twice each, same box.
out=gpurun_out; mkdir -p $out
tag=${1:-r02an}
SAV_DW_STREAM=1 timeout 600 python -m pytest tests/test_vit_parity_gpu.py tests/test_train_state_gpu.py -q -m gpu -x > $out/${tag}_dw_parity.log 2>&1
echo "parity exit $?" >> $out/${tag}_dw_parity.log
tail -3 $out/${tag}_dw_parity.log
for i in 1 2; do
  for v in 0 1; do
    SAV_DW_STREAM=$v timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > $out/${tag}_bench_dw${v}_${i}.json 2> $out/${tag}_bench_dw${v}_${i}.err
    python - <<PY
import json
try:
    d=json.loads(open("$out/${tag}_bench_dw${v}_${i}.json").read().strip().splitlines()[-1])
    print("dw_stream=$v run $i: ms_per_step", d["ms_per_step"], "value", d["value"], "e2e", d["e2e"]["value"], "clocks", d["clocks"])
except Exception as e:
    print("dw_stream=$v run $i failed", e)
PY
  done
done
