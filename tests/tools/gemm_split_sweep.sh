# Development (GPU, -DSAV_GEMM_DEV build): operand stages vs slab slots for the GEMMs whose epilogue
# streams as many bytes as the operands.
export SAV_TIME_ONLY=${SAV_TIME_ONLY:-t_out_fwd,t_fc1_fwd,t_fc2_fwd,t_fc2_dx}
for st in 0 5 4 3; do
  echo "== SAV_GEMM_STAGES=$st"
  SAV_GEMM_STAGES=$st timeout 300 python tests/gpu_diag.py --time
done
