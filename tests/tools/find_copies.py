"""Development tool (GPU): where does the training step still call into PyTorch copy kernels?
Wraps Tensor.contiguous / Tensor.copy_ / Tensor.to / Tensor.clone and prints the call sites that
actually launch a device copy during one eager DeiT-S step."""
import collections
import os
import sys
import traceback

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "self-attention-experiments-vision_b200"))


def main():
    import sav  # noqa: F401
    from models import create_model
    from oracle import vit_oracle as O
    from sav.train import HWCN, TrainState, train_step
    name = sys.argv[1] if len(sys.argv) > 1 else "deit_s_patch16"
    cfg = O.vit_config(name)
    B = 32
    images, labels = O.synthetic_batch(B, img_size=224, seed=0)
    model = create_model(name)
    x = HWCN(images.cuda())
    y = labels.cuda()
    variables = model.init({"params": 42}, x, is_training=True)
    variables = {"params": variables["params"].arena.load_tree(O.synthetic_vit_params(cfg, 224, seed=42))}
    state = TrainState(model, variables, lr=1e-3, weight_decay=1e-4, clip_grad=1.0)
    train_step(state, x, y, 0.1, None)
    torch.cuda.synchronize()
    sites = collections.Counter()
    orig_contig, orig_copy, orig_clone = torch.Tensor.contiguous, torch.Tensor.copy_, torch.Tensor.clone

    def site():
        for fr in reversed(traceback.extract_stack()[:-2]):
            if "self-attention-experiments-vision_b200" in fr.filename:
                return f"{os.path.basename(fr.filename)}:{fr.lineno} {fr.line}"
        return "?"

    def contig(t, *a, **k):
        if t.is_cuda and not t.is_contiguous():
            sites["contiguous " + site()] += 1
        return orig_contig(t, *a, **k)

    def copy_(t, *a, **k):
        if t.is_cuda:
            sites["copy_ " + site()] += 1
        return orig_copy(t, *a, **k)

    def clone(t, *a, **k):
        if t.is_cuda:
            sites["clone " + site()] += 1
        return orig_clone(t, *a, **k)
    torch.Tensor.contiguous, torch.Tensor.copy_, torch.Tensor.clone = contig, copy_, clone
    try:
        train_step(state, x, y, 0.1, None)
        torch.cuda.synchronize()
    finally:
        torch.Tensor.contiguous, torch.Tensor.copy_, torch.Tensor.clone = orig_contig, orig_copy, orig_clone
    for s, n in sites.most_common():
        print(f"{n:4d}  {s}")


if __name__ == "__main__":
    main()
