"""Development tool (GPU): dump our and the oracle's gradient of selected tensors for DeiT-S/16
224^2 B=32 to gpurun_out/ for offline error-pattern analysis."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "self-attention-experiments-vision_b200"))
import sav
from models import create_model
from oracle import vit_oracle as O
from sav.train import HWCN, make_loss_fn
name, batch, img = "deit_s_patch16", int(os.environ.get("B", 32)), 224
cfg = O.vit_config(name)
params = O.synthetic_vit_params(cfg, img_size=img, seed=42)
images, labels = O.synthetic_batch(batch, img_size=img, seed=0)
_, logits_ref, grads_ref = O.vit_loss_and_grads(params, images, labels, cfg, 0.1)
model = create_model(name)
x = HWCN(images.cuda())
variables = sav.load_params(model, params, x, is_training=True)
(loss, logits), grads = sav.value_and_grad(make_loss_fn(model, 0.1), has_aux=True)(variables["params"], x, labels.cuda())
torch.cuda.synchronize()
ours = dict(O.tree_leaves(grads)); ref = dict(O.tree_leaves(grads_ref))
rel = lambda a, b: float((a.float().cpu() - b).abs().max() / b.abs().max())
errs = sorted(((rel(ours[k], ref[k]), k) for k in ref), reverse=True)
print(errs[:8])
keep = [k for _, k in errs[:6]]
torch.save({k: (ours[k].float().cpu(), ref[k]) for k in keep}, os.path.join(ROOT, "gpurun_out", "qgrad_dump.pt"))
