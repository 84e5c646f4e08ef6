"""Times fwd+bwd of the variant configs of BASELINE.json on one B200 (CUDA events, after warm-up):
CaiT-S24 224^2 (B=128), ViT-B/16 384^2 (B=64, 577 tokens), BoTNet MHSA block (B=256, 14x14x512).
Development / DESIGN.md numbers -- not the bench metric.

    python tests/tools/variant_time.py [cait|vitb384|bot|botnet|cvt|mixer|tnt|ceit]
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "self-attention-experiments-vision_b200"))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def model_case(name, batch, img, rngs=None):
    import sav
    from models import create_model
    from sav import _abi
    from sav import functional as F
    from sav.train import HWCN
    model = create_model(name)
    images = torch.randn((img, img, 3, batch), device="cuda")
    labels = torch.randint(0, 1000, (batch,), device="cuda", dtype=torch.int32)
    variables = model.init({"params": 0, **(rngs or {})}, HWCN(images), is_training=True)
    arena = variables["params"].arena
    # non-degenerate head (the reference's zero-init head makes every other gradient zero)
    arena.master.normal_(0.0, 0.02)
    arena.refresh_shadow()

    def loss_fn(p, im, lb):
        logits = model.apply({"params": p}, im, is_training=True, rngs=rngs)
        return F.softmax_cross_entropy(logits, lb, 0.1), logits
    step = sav.value_and_grad(loss_fn, has_aux=True)
    n0 = _abi.launch_count()
    step(variables["params"], HWCN(images), labels)
    launches = _abi.launch_count() - n0
    ms = timed(lambda: step(variables["params"], HWCN(images), labels))
    return {"case": f"{name} {img}^2 B={batch} fwd+bwd", "ms": ms, "images_per_s": batch / ms * 1e3,
            "launches": launches}


def bot_case(batch=256):
    import sav
    from models import BoTMHSA
    from sav.tape import Var
    m = BoTMHSA(num_heads=4, head_ch=128)
    x = torch.randn((batch, 14, 14, 512), device="cuda").bfloat16()
    cot = torch.randn((batch, 14, 14, 512), device="cuda").bfloat16()
    variables = m.init({"params": 0}, Var(x), device="cuda")

    class Cot:
        def __init__(self, out):
            self.out = out

        def seed(self):
            self.out.g = cot

    def fn(p, xv):
        return Cot(m.apply({"params": p}, xv))
    step = sav.value_and_grad(fn)
    ms = timed(lambda: step(variables["params"], Var(x)))
    flops = 3 * (2 * batch * 196 * 512 * 1536 + 4 * batch * 4 * 196 * 196 * 128)
    return {"case": f"BoTMHSA 14x14x512 h4 d128 B={batch} fwd+bwd", "ms": ms, "tflops": flops / ms / 1e9}


if __name__ == "__main__":
    which = sys.argv[1:] or ["cait", "vitb384", "bot", "botnet", "cvt", "mixer", "tnt", "ceit"]
    for w in which:
        if w == "cait":
            r = model_case("cait_s_24", 128, 224, rngs={"stochastic_depth": 1})
        elif w == "vitb384":
            r = model_case("vit_b_patch16", 64, 384)
        elif w == "botnet":
            r = model_case("botnet_t3", 64, 224)      # BASELINE config 5 as a whole model
        elif w == "cvt":
            r = model_case("cvt-13", 64, 224)
        elif w == "mixer":
            r = model_case("mixer_b_patch16", 128, 224)
        elif w == "tnt":
            r = model_case("tnt_s_patch16", 64, 224)
        elif w == "ceit":
            r = model_case("ceit_s", 64, 224)
        else:
            r = bot_case()
        print("RESULT " + json.dumps(r), flush=True)
