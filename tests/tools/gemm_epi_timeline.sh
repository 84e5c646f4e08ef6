# Needs a dev build: SAV_GEMM_DEV=1 python self-attention-experiments-vision_b200/build.py
# fine-grained epilogue timeline of CTA 0: QKV forward (bias) and FC1 forward (bias + GELU, two outputs)
echo "== qkv_fwd"; timeout 100 python tests/tools/gemm_timeline.py --M 50432 --N 1152 --K 384 --bias --role epilogue --lines 70
echo "== fc1_fwd"; timeout 100 python tests/tools/gemm_timeline.py --M 50432 --N 1536 --K 384 --gelu2 --role epilogue --lines 70
