# Needs a dev build: SAV_GEMM_DEV=1 python self-attention-experiments-vision_b200/build.py
export SAV_TIME_ONLY=t_sq,t_fc1_dw,t_fc2_dw,t_fc1_dx,t_fc1_fwd
for c in 1 2; do for d in 0 1 2; do echo "== ctas=$c debug=$d"; SAV_GEMM_CTAS=$c SAV_GEMM_DEBUG=$d timeout 200 python tests/gpu_diag.py --time; done; done
python - <<'PY'
import torch,time
a=torch.randn(8192,8192,device='cuda').bfloat16(); b=torch.randn(8192,8192,device='cuda').bfloat16()
for _ in range(3): a@b
e0,e1=torch.cuda.Event(True),torch.cuda.Event(True)
torch.cuda.synchronize(); e0.record()
for _ in range(20): a@b
e1.record(); torch.cuda.synchronize()
ms=e0.elapsed_time(e1)/20
print("cublas 8192^3", ms*1e3, "us", 2*8192**3/ms/1e9, "TF/s")
PY
