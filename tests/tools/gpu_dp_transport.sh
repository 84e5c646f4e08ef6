# gpurun --gpus N: data-parallel bench per gradient-exchange transport (sav/dp.py) and block count.
N=${NGPU:-2}
T=${TAG:-r02af}
run() {  # label, env...
  local label=$1; shift
  env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${T}_dp${N}_${label}.json 2> gpurun_out/${T}_dp${N}_${label}.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${T}_dp${N}_${label}.json").read().strip().splitlines()[-1])
    print("${label}", "N=$N", round(d["ms_per_step"], 3), "ms", round(d["value"]), "img/s  e2e", round(d["e2e"]["value"]), d["config"].get("grad_exchange"))
except Exception as e:
    print("${label}", "FAILED", e)
PY
}
run multimem SAV_DP_TRANSPORT=peer
run nccl SAV_DP_TRANSPORT=nccl
run p2p SAV_DP_TRANSPORT=peer SAV_DP_MULTICAST=0
for b in ${BLOCKS:-32 64 256}; do run multimem_b$b SAV_DP_TRANSPORT=peer SAV_DP_BLOCKS=$b; done
CUDA_VISIBLE_DEVICES=0 timeout 100 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${T}_dp1.json 2>/dev/null
python -c "import json; d=json.loads(open('gpurun_out/${T}_dp1.json').read().strip().splitlines()[-1]); print('single GPU', round(d['ms_per_step'],3), 'ms', round(d['value']), 'img/s')"
