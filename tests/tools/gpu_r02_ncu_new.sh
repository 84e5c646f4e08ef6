# ncu counters of the kernels added late in round 2: the long-sequence attention backward (query
# groups + fp32 reduce-add + conversion) and the tensor-core talking-heads row kernels.
T=${TAG:-r02bk}
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread"
SAV_TIME_ONLY=t_attn_bwd577 timeout 300 ncu --metrics $M --clock-control none -k regex:"attn_bwd3|acc_to_bnhd|fill_kernel" --launch-skip 9 -c 4 --csv --log-file gpurun_out/${T}_ncu_t_attn_bwd577.csv python tests/gpu_diag.py --time > gpurun_out/${T}_ncu.log 2>&1
M2="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__inst_executed_pipe_tensor.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,smsp__issue_active.avg.pct_of_peak_sustained_active"
timeout 300 ncu --metrics $M2 --clock-control none -k regex:"thsoftmax_(fwd|bwd)_tc" --launch-skip 4 -c 2 --csv --log-file gpurun_out/${T}_ncu_thsoftmax_tc.csv python tests/tools/th_once.py --reps 2 > gpurun_out/${T}_ncu2.log 2>&1
tail -n 2 gpurun_out/${T}_ncu.log; tail -n 2 gpurun_out/${T}_ncu2.log
