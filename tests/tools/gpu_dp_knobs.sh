# gpurun --gpus 2: how much of the data-parallel step is the exchange, and does it disturb the
# persistent one-CTA-per-SM kernels it overlaps with?  (NCCL's kernels take SMs.)
run() { echo "== $*"; env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; l=[x for x in sys.stdin.read().splitlines() if x.startswith('{')]; d=json.loads(l[-1]); print(d['value'], d['ms_per_step'])"; }
run SAV_DP_OVERLAP=1
run SAV_DP_OVERLAP=0
run SAV_DP_OVERLAP=1 NCCL_MAX_CTAS=4
run SAV_DP_OVERLAP=1 NCCL_MAX_CTAS=2 SAV_DP_BUCKET_MB=32
run SAV_DP_OVERLAP=1 SAV_DP_BUCKET_MB=100
