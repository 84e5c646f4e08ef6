set -x
N=${NGPU:-8}
T=${TAG:-r01g}
mkdir -p gpurun_out
for n in $N 4; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${T}_bench_dp$n.json 2> gpurun_out/${T}_bench_dp$n.err
echo "rc=$?" >> gpurun_out/${T}_bench_dp$n.err
done
tail -3 gpurun_out/${T}_bench_dp$N.err
