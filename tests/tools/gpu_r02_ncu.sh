# ncu counters of the step's heaviest kernels (one gpurun call): duration, DRAM bytes, the tensor-pipe
# counters that DO count tcgen05 (sm__pipe_tensor_cycles_active, sm__ops_path_tensor_op_utchmma_*),
# shared-memory bank conflicts, achieved occupancy.  Each case runs once without ncu first.
T=${TAG:-r02}
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32.sum,sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32.avg.pct_of_peak_sustained_elapsed,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,sm__throughput.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"
for c in t_attn_bwd:attn_bwd3 t_attn_fwd:attn_fwd2 t_fc1_fwd:gemm_kernel t_qkv_fwd:gemm_kernel t_fc2_dw:gemm_kernel t_fc2_fwd_ln:gemm_kernel t_out_fwd_ln:gemm_kernel t_fc2_dx_cs:gemm_kernel; do
  case=${c%%:*}; k=${c##*:}
  SAV_TIME_ONLY=$case timeout 120 python tests/gpu_diag.py --time > gpurun_out/${T}_plain_${case}.log 2>&1 &&
  SAV_TIME_ONLY=$case timeout 300 ncu --metrics $M --clock-control none -k regex:$k --launch-skip 3 -c 1 --csv --log-file gpurun_out/${T}_ncu_${case}.csv python tests/gpu_diag.py --time > gpurun_out/${T}_ncu_${case}.log 2>&1
done
timeout 120 python tests/tools/ln_once.py > gpurun_out/${T}_plain_ln.log 2>&1 &&
for k in ln_bwd_pieces ln_fwd_pieces; do
  timeout 300 ncu --metrics $M --clock-control none -k regex:$k --launch-skip 3 -c 1 --csv --log-file gpurun_out/${T}_ncu_${k}.csv python tests/tools/ln_once.py > gpurun_out/${T}_ncu_${k}.log 2>&1
done
# every launch of one bench step with its device time (shares of the step)
timeout 200 python bench.py --steps 2 --warmup 1 --no-graph --no-cpu-baseline > gpurun_out/${T}_plain_bench_nograph.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${T}_launches_bench_nograph.csv python bench.py --steps 2 --warmup 1 --no-graph --no-cpu-baseline > gpurun_out/${T}_ncu_list.log 2>&1
# one full-set capture of the kernel that changed last (FULL=case:kernel-regex:name), with source correlation
F=${FULL:-t_attn_bwd:attn_bwd3:attn_bwd3}
fc=${F%%:*}; fr=${F#*:}; fk=${fr%%:*}; fn=${fr##*:}
SAV_TIME_ONLY=$fc timeout 300 ncu --set full --clock-control none --import-source on -k regex:$fk --launch-skip 3 -c 1 -f -o gpurun_out/${T}_${fn}_full python tests/gpu_diag.py --time > gpurun_out/${T}_ncu_full.log 2>&1
ls gpurun_out | grep ${T}_ncu
