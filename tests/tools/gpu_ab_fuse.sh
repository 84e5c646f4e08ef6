# same-box A/B of the bias-gradient fusion into LayerNorm backward
for v in 1 0 1 0; do
  SAV_FUSE_BIAS_COLSUM=$v timeout 100 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('fuse=$v', round(d['value']), round(d['ms_per_step'],3), d['gpu_launches'])"
done
