# A/B of an environment switch on the same box: bash tests/tools/gpu_ab_env.sh VAR a b
V=$1; A=$2; B=$3
for v in $A $B $A $B; do
  echo "== $V=$v"
  env $V=$v timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline | grep -o '"value": [0-9.]*, "unit": "images/s", "n_gpus"\|"ms_per_step": [0-9.]*'
done
