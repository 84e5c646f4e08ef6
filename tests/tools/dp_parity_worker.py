"""Worker of tests/test_dp_gpu.py (launched under torchrun, one rank per GPU): every rank runs the
data-parallel training-step gradient path (value_and_grad with the DataParallel hook: bucketed NCCL
all-reduce overlapped with backward) on ITS shard of a global batch; rank 0 compares the reduced
gradient arena with (a) the fp32 oracle's gradients of the whole global batch and (b) a single-rank
run of the whole batch on the same kernels.  Prints one JSON line."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "self-attention-experiments-vision_b200"))


def main():
    import sav
    from models import create_model
    from oracle import vit_oracle as O
    from sav.dp import DataParallel
    from sav.train import HWCN, make_loss_fn
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    name, per_rank, img = "deit_ti_patch16", 4, 224
    cfg = O.vit_config(name)
    params = O.synthetic_vit_params(cfg, img_size=img, seed=42)
    images, labels = O.synthetic_batch(per_rank * world, img_size=img, seed=0)
    model = create_model(name)
    lo, hi = rank * per_rank, (rank + 1) * per_rank
    x = HWCN(images[..., lo:hi].contiguous().cuda())
    y = labels[lo:hi].cuda()
    variables = sav.load_params(model, params, x, is_training=True)
    dp = DataParallel(bucket_mb=4.0)
    dp.broadcast_params(variables["params"].arena)
    loss_fn = make_loss_fn(model, 0.1, world)
    for _ in range(2):      # first pass learns the bucket schedule, second runs overlapped
        (loss, _), grads = sav.value_and_grad(loss_fn, has_aux=True, grad_hook=dp)(variables["params"], x, y)
    torch.cuda.synchronize()
    out = {"world": world, "buckets": dp.stats["buckets"], "transport": dp.stats.get("transport", dp.transport)}
    if rank == 0:
        _, _, grads_ref = O.vit_loss_and_grads(params, images, labels, cfg, 0.1)
        ours = dict(O.tree_leaves(grads))
        rel = lambda a, b: float((a.float().cpu() - b).abs().max() / b.abs().max().clamp_min(1e-20))
        errs = {k: rel(ours[k], g) for k, g in O.tree_leaves(grads_ref)}
        out["worst_vs_oracle"] = max(errs.values())
        out["worst_l2_vs_oracle"] = max(float((ours[k].float().cpu() - g).norm() / g.norm()) for k, g in O.tree_leaves(grads_ref))
        # the same global batch on ONE rank through the same kernels
        xg = HWCN(images.cuda())
        v1 = sav.load_params(model, params, xg, is_training=True)
        (_, _), g1 = sav.value_and_grad(make_loss_fn(model, 0.1, 1), has_aux=True)(v1["params"], xg, labels.cuda())
        torch.cuda.synchronize()
        single = dict(O.tree_leaves(g1))
        vs = sorted(((rel(ours[k], single[k].float().cpu()), k) for k in single), reverse=True)
        out["worst_vs_single_rank"] = vs[0][0]
        out["worst_vs_single_rank_tensors"] = vs[:3]
    dist.barrier()
    # every rank must hold the same reduced gradients
    chk = variables["params"].arena.grad.double().sum().reshape(1)
    gathered = [torch.zeros_like(chk) for _ in range(world)]
    dist.all_gather(gathered, chk)
    if rank == 0:
        out["checksums_equal"] = all(float(g) == float(gathered[0]) for g in gathered)
        print("DPRESULT " + json.dumps(out), flush=True)
    dist.barrier()
    torch.cuda.synchronize()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
