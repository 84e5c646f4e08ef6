"""Training driver with the reference's command line (train.py:130-255), running on the B200 path.

Same option names and defaults as the reference's click CLI (--data_dir --img_size --num_epochs
--batch_size --label_smoothing --augmentation --model_name --lr --weight_decay --clip_grad
--checkpoint_dir --seed); same loop shape: schedule = warm-up (5 epochs) + cosine to epoch 30 with
base_lr = lr * batch_size / 512 (train.py:214-220), `for step, batch in zip(range(steps_total),
train_input)` (train.py:233), checkpoint every 10 epochs (train.py:236-237, keep=3), eval every 5
epochs (train.py:239-252).

What differs, and why: the reference's loader (`input_pipeline.load`, TensorFlow /
tfds) does not exist and is out of scope, so `build_train_input` / `build_eval_input` are a
synthetic feeder producing batches of the loader's shape -- images [H, W, C, N] float32
(transpose=True), labels int32, and for the cutmix / mixup augmentations `mix_labels` + `ratio`
(train.py:84-87) -- from pinned host memory; --fake_data is therefore the only data source.  The LR
sign is descent.  One process per GPU (torchrun) replaces pmap; wandb logging is a JSON line on
stdout.  The step itself is sav.train.GraphedTrainStep: one CUDA graph per step, the next batch
copied host->device while the current step runs.

    python train.py --model_name deit_s_patch16 --batch_size 256 --num_epochs 1 --steps_per_epoch 20
"""
import json
import os
import sys
import time

import click
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
if HERE not in sys.path:
    sys.path.insert(0, HERE)

import sav  # noqa: E402,F401
from models import create_model  # noqa: E402
from sav import checkpoint as checkpoints  # noqa: E402
from sav.train import (GraphedTrainStep, HWCN, TrainState, eval_step as _eval_step,  # noqa: E402
                       reference_schedule, train_step as _train_step)

IMAGES_PER_EPOCH = 1281167   # ImageNet-1k train split (train.py:203)


class FakeImageNet:
    """Synthetic stand-in for input_pipeline.load(..., transpose=True, fake_data=True): an endless
    iterator of host batches {'images': [H,W,3,N] float32, 'labels': [N] int32} (+ 'mix_labels',
    'ratio' when the augmentation mixes images: cutmix / mixup, train.py:84-87).  A small pool of
    pinned batches is cycled so the feeder never limits the step."""

    def __init__(self, batch_size, img_size, num_classes=1000, mix=False, seed=0, pool=4, training=True):
        g = torch.Generator().manual_seed(seed)
        self.batches = []
        for _ in range(pool):
            b = {"images": torch.randn((img_size, img_size, 3, batch_size), generator=g).pin_memory(),
                 "labels": torch.randint(0, num_classes, (batch_size,), generator=g, dtype=torch.int32).pin_memory()}
            if mix and training:
                b["mix_labels"] = torch.randint(0, num_classes, (batch_size,), generator=g,
                                                dtype=torch.int32).pin_memory()
                b["ratio"] = torch.rand((batch_size,), generator=g).pin_memory()
            self.batches.append(b)
        self.i = 0

    def __iter__(self):
        return self

    def __next__(self):
        b = self.batches[self.i % len(self.batches)]
        self.i += 1
        return b


def _world():
    return int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), \
        int(os.environ.get("LOCAL_RANK", "0"))


def build_train_input(data_dir, batch_size, img_size, augmentation, seed=0):
    """train.py:41-58.  batch_size is the GLOBAL batch; each rank feeds batch_size / world."""
    world, rank, _ = _world()
    per_device, ragged = divmod(batch_size, world)
    if ragged:
        raise ValueError(f'Batch size {batch_size} must be divisible by num devices {world}')
    mix = augmentation is not None and ("cutmix" in augmentation or "mixup" in augmentation)
    return FakeImageNet(per_device, img_size, mix=mix, seed=seed * 1000 + rank, training=True)


def build_eval_input(data_dir, batch_size, img_size, seed=0):
    """train.py:61-74."""
    world, rank, _ = _world()
    return FakeImageNet(batch_size // world, img_size, seed=seed * 1000 + 500 + rank, training=False)


def create_train_state(rng, model, img_size, lr_schedule_fn, weight_decay, max_norm, device="cuda"):
    """train.py:22-38: optax.chain(clip_by_global_norm, scale_by_adam, additive_weight_decay,
    scale_by_schedule) + model.init on a [1, S, S, 3] input -> TrainState."""
    x = torch.ones((1, img_size, img_size, 3), device=device)
    variables = model.init({"params": rng, "stochastic_depth": rng}, x, is_training=False)
    return TrainState(model, variables, lr=lr_schedule_fn, weight_decay=weight_decay,
                      clip_grad=max_norm, seed=rng)


def train_step(train_state, batch, label_smoothing, dp=None):
    """train.py:77-109 on device-resident batch tensors (eager form; the loop below uses the graphed
    form of the same function)."""
    return _train_step(train_state, HWCN(batch["images"]), batch["labels"], label_smoothing, dp,
                       batch.get("mix_labels"), batch.get("ratio"))


def eval_step(train_state, batch, dp=None):
    """train.py:112-120."""
    return _eval_step(train_state, HWCN(batch["images"]), batch["labels"], dp)


def save_checkpoint(train_state, dir):
    """train.py:123-127: process 0 writes checkpoint_<step>, keep=3."""
    if _world()[1] == 0:
        checkpoints.save_checkpoint(dir, train_state, int(train_state.step), keep=3, overwrite=True)


@click.command()
@click.option('--data_dir', type=str, default=None, help='unused: the TF/tfds loader is out of scope')
@click.option('--img_size', type=int, default=224)
@click.option('--num_epochs', type=int, default=300)
@click.option('--batch_size', type=int, default=32)
@click.option('--label_smoothing', type=float, default=0.1)
@click.option('--augmentation', type=str, default='cutmix_mixup_randaugment_405')
@click.option('--model_name', type=str, default='vit_b_patch16')
@click.option('--lr', type=float, default=5e-4)
@click.option('--weight_decay', type=float, default=1e-4)
@click.option('--clip_grad', type=float, default=None)
@click.option('--checkpoint_dir', type=str, default=None)
@click.option('--seed', type=int, default=42)
@click.option('--fake_data/--no-fake_data', default=True, help='synthetic batches (the only source)')
@click.option('--steps_per_epoch', type=int, default=None,
              help='override images_per_epoch // batch_size (short runs)')
def main(data_dir, img_size, num_epochs, batch_size, label_smoothing, augmentation, model_name, lr,
         weight_decay, clip_grad, checkpoint_dir, seed, fake_data, steps_per_epoch):
    if not fake_data:
        raise click.UsageError("only --fake_data is available: the reference's TensorFlow input "
                               "pipeline (data/**) is out of scope and broken as written (SURVEY D14)")
    world, rank, local = _world()
    torch.cuda.set_device(local)
    dp = None
    if world > 1:
        import torch.distributed as dist
        from sav.dp import DataParallel
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        dp = DataParallel()

    steps_per_epoch = steps_per_epoch or IMAGES_PER_EPOCH // batch_size        # train.py:205
    steps_per_eval = steps_per_epoch * 5
    steps_per_checkpoint = steps_per_epoch * 10
    steps_total = steps_per_epoch * num_epochs

    train_input = build_train_input(data_dir, batch_size, img_size, augmentation, seed)
    eval_input = build_eval_input(data_dir, batch_size, img_size, seed)
    schedule_fn = reference_schedule(lr, batch_size, steps_per_epoch)          # train.py:214-220
    model = create_model(model_name=model_name, num_classes=1000, dtype=torch.bfloat16)
    train_state = create_train_state(seed, model, img_size, schedule_fn, weight_decay, clip_grad)
    if dp is not None:
        dp.broadcast_params(train_state.arena)                                 # replicate, train.py:228

    # device-resident batch buffers the captured step reads; host batches stream through them
    first = next(train_input)
    dev = {k: v.cuda(non_blocking=True) for k, v in first.items()}
    eval_dev = {k: v.cuda(non_blocking=True) for k, v in next(eval_input).items()}
    train_fn = GraphedTrainStep(train_state, HWCN(dev["images"]), dev["labels"], label_smoothing, dp,
                                mix_labels=dev.get("mix_labels"), ratio=dev.get("ratio"))
    t0 = time.perf_counter()
    train_fn.prefetch(first["images"], first["labels"])
    staged = first
    for step, batch in zip(range(steps_total), train_input):                   # train.py:233
        for k in ("mix_labels", "ratio"):                                      # small: plain copies
            if k in staged:
                dev[k].copy_(staged[k], non_blocking=True)
        loss = train_fn.step_prefetched()                                      # the staged batch
        train_fn.prefetch(batch["images"], batch["labels"])                    # next one, overlapped
        staged = batch
        if step % 50 == 0 and rank == 0:
            print(json.dumps({"step": step, "train/loss": loss.item(), "lr": train_state.current_lr(),
                              "images_per_s": (step + 1) * batch_size / (time.perf_counter() - t0)}),
                  flush=True)
        if checkpoint_dir and step % steps_per_checkpoint == 0 and step != 0:  # train.py:236-237
            torch.cuda.synchronize()
            save_checkpoint(train_state, checkpoint_dir)
        if step % steps_per_eval == 0 and step != 0:                           # train.py:239-252
            b = next(eval_input)
            for k, v in b.items():
                eval_dev[k].copy_(v, non_blocking=True)
            ev = eval_step(train_state, eval_dev, dp)
            if rank == 0:
                print(json.dumps({"step": step, "eval/loss": ev.item()}), flush=True)
    torch.cuda.synchronize()
    if checkpoint_dir:
        save_checkpoint(train_state, checkpoint_dir)
    if rank == 0:
        print(json.dumps({"done": True, "steps": steps_total, "final_loss": loss.item() if steps_total else None,
                          "seconds": time.perf_counter() - t0}), flush=True)


if __name__ == '__main__':
    main()
