"""Regenerates tests/golden/vit_micro.npz: a REGRESSION fixture of the oracle itself.

The reference ships no golden vectors and cannot be imported here (jax/flax absent),
so this fixture does not pin the oracle to the reference -- oracle/vit_oracle.py says "parity
unpinned" -- it pins the oracle against accidental edits, and gives the GPU tests a second,
file-based target.  Run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import vit_oracle as O  # noqa: E402

CFG = dict(num_layers=2, num_heads=1, embed_dim=64, patch_shape=(16, 16), num_classes=16,
           expand_ratio=4)
IMG, BATCH, SMOOTH = 48, 5, 0.1


def compute():
    params = O.synthetic_vit_params(CFG, IMG, seed=7)
    images, labels = O.synthetic_batch(BATCH, IMG, num_classes=CFG["num_classes"], seed=3)
    loss, logits, grads = O.vit_loss_and_grads(params, images, labels, CFG, SMOOTH)
    out = {"loss": loss.numpy(), "logits": logits.numpy()}
    for k, g in O.tree_leaves(grads):
        out["grad:" + k] = g.numpy()
    return out


if __name__ == "__main__":
    np.savez_compressed(os.path.join(os.path.dirname(__file__), "vit_micro.npz"), **compute())
    print("written")
