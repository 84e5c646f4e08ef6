"""The train.py-shaped driver (reference train.py:130-255): the reference's option names parse, the
synthetic feeder produces loader-shaped batches, and (GPU) a short run trains, evaluates and
writes a Flax-layout checkpoint."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRIVER = os.path.join(ROOT, "self-attention-experiments-vision_b200", "train.py")

REFERENCE_OPTIONS = ["--data_dir", "--img_size", "--num_epochs", "--batch_size", "--label_smoothing",
                     "--augmentation", "--model_name", "--lr", "--weight_decay", "--clip_grad",
                     "--checkpoint_dir", "--seed"]


def test_cli_keeps_the_reference_option_names_and_defaults():
    """train.py:130-190."""
    out = subprocess.run([sys.executable, DRIVER, "--help"], capture_output=True, text=True, check=True).stdout
    for opt in REFERENCE_OPTIONS:
        assert opt in out, opt
    import importlib.util
    spec = importlib.util.spec_from_file_location("sav_train_driver", DRIVER)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    defaults = {p.name: p.default for p in mod.main.params}
    assert defaults["img_size"] == 224 and defaults["num_epochs"] == 300 and defaults["batch_size"] == 32
    assert defaults["label_smoothing"] == 0.1 and defaults["lr"] == 5e-4 and defaults["weight_decay"] == 1e-4
    assert defaults["clip_grad"] is None and defaults["seed"] == 42


def test_reference_schedule_and_feeder_shapes():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("the feeder pins host memory (needs a CUDA runtime)")
    import importlib.util
    spec = importlib.util.spec_from_file_location("sav_train_driver", DRIVER)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    it = mod.build_train_input(None, 8, 32, "cutmix_mixup_randaugment_405")
    b = next(it)
    assert tuple(b["images"].shape) == (32, 32, 3, 8) and b["images"].dtype == torch.float32   # H W C N
    assert b["labels"].dtype == torch.int32 and set(b) == {"images", "labels", "mix_labels", "ratio"}
    assert set(next(mod.build_eval_input(None, 8, 32))) == {"images", "labels"}


def test_warmup_cosine_host_formula_matches_oracle():
    from oracle import vit_oracle as O
    from sav.train import reference_schedule
    s = reference_schedule(5e-4, 256, 100)          # train.py:214-220
    ref = O.warmup_cosine_decay_schedule(0.0, 5e-4 * 256 / 512, 500, 3000, 1e-5)
    for c in (0, 1, 250, 499, 500, 501, 1750, 2999, 3000, 10 ** 5):
        assert abs(s(c) - ref(c)) <= 1e-12


@pytest.mark.gpu
def test_driver_trains_evaluates_and_checkpoints(tmp_path):
    cmd = [sys.executable, DRIVER, "--model_name", "deit_ti_patch16", "--img_size", "64", "--batch_size", "8",
           "--num_epochs", "11", "--steps_per_epoch", "2", "--clip_grad", "1.0", "--lr", "1e-3",
           "--checkpoint_dir", str(tmp_path)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert lines[-1]["done"] and lines[-1]["steps"] == 22
    assert any("eval/loss" in l for l in lines) and all(l["train/loss"] == l["train/loss"] for l in lines if "train/loss" in l)
    files = sorted(os.listdir(tmp_path))
    assert files and all(f.startswith("checkpoint_") for f in files) and len(files) <= 3
    from sav import checkpoint as C
    sd = C.restore_checkpoint(str(tmp_path))
    assert set(sd["opt_state"]) == {"0", "1", "2", "3"} and "Encoder_0" in sd["params"]["params"]
