# Needs a dev build: SAV_GEMM_DEV=1 python self-attention-experiments-vision_b200/build.py
# MMA-only cost per tile (24 k-blocks = 96 MMAs) by accumulator width and CTA-pair mode:
# loads, TMEM reads and stores disabled (bits 1+8+16), so the tile period is pure issue pacing.
#  +128 issues every MMA twice; +256 no full-barrier wait; +512 no per-k-block commit / empty wait
for dbg in 25 281 793; do for ctas in 1 2; do for bn in 128 256; do
  echo "== debug=$dbg ctas=$ctas bn=$bn"
  SAV_GEMM_CTAS=$ctas SAV_GEMM_BN=$bn SAV_GEMM_DEBUG=$dbg timeout 60 python tests/tools/gemm_timeline.py --M 50432 --N 1536 --K 1536 --lines 60 | grep "M issued" | head -3
done; done; done
