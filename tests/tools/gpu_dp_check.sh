# gpurun --gpus N: A/B of the GEMM CTA-pair heuristic on one GPU, then the data-parallel bench.
set -x
N=${NGPU:-2}
T=${TAG:-r01e}
mkdir -p gpurun_out
for c in 1 0 1 0; do
  echo "== SAV_GEMM_CTAS=$c" >> gpurun_out/${T}_ab.log
  CUDA_VISIBLE_DEVICES=0 SAV_GEMM_CTAS=$c python bench.py --steps 20 --warmup 5 --no-cpu-baseline >> gpurun_out/${T}_ab.log 2>&1
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${T}_bench_dp$N.json 2> gpurun_out/${T}_bench_dp$N.err
echo "rc=$?" >> gpurun_out/${T}_bench_dp$N.err
tail -5 gpurun_out/${T}_bench_dp$N.err
