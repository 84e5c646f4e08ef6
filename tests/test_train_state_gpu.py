"""Training-step host logic on the device (train.py:22-38,77-120): the mixup branch of the loss,
the device-resident optax count + warm-up-cosine schedule, drop-path masks that resample on CUDA
graph replay, an unsynchronised graphed loop against eager synchronised steps, and eval_step."""
import math

import pytest
import torch

from oracle import vit_oracle as O


def _tiny_state(lr, seed=42, clip=1.0, name="deit_ti_patch16", img=64, batch=4):
    import sav
    from models import create_model
    from sav.train import HWCN, TrainState
    cfg = O.vit_config(name)
    images, labels = O.synthetic_batch(batch, img_size=img, seed=0)
    model = create_model(name)
    x = HWCN(images.cuda())
    variables = sav.load_params(model, O.synthetic_vit_params(cfg, img, seed=seed), x, is_training=True)
    return TrainState(model, variables, lr=lr, weight_decay=1e-4, clip_grad=clip), x, labels.cuda(), cfg


@pytest.mark.gpu
def test_mixup_loss_and_dlogits_match_oracle():
    """train.py:84-87: y = ratio*onehot(l) + (1-ratio)*onehot(l2), then smooth_labels, CE, mean."""
    from sav import ops
    g = torch.Generator().manual_seed(5)
    B, classes, alpha = 37, 1000, 0.1
    logits = (torch.randn((B, classes), generator=g) * 2).bfloat16()
    labels = torch.randint(0, classes, (B,), generator=g, dtype=torch.int32)
    mix = torch.randint(0, classes, (B,), generator=g, dtype=torch.int32)
    mix[:5] = labels[:5]                                   # partner == label is a legal draw
    ratio = torch.rand((B,), generator=g)
    ratio[5], ratio[6] = 0.0, 1.0
    lr = logits.float().requires_grad_(True)
    loss_ref = O.loss_fn(lr, labels, classes, alpha, 1, mix, ratio)
    loss_ref.backward()
    dl, stats = ops.softmax_xent(logits.cuda(), labels.cuda(), alpha, 1.0 / B, mix.cuda(), ratio.cuda())
    torch.cuda.synchronize()
    assert abs(float(stats[0]) / B - float(loss_ref)) <= 1e-5 * abs(float(loss_ref))
    err = float((dl.float().cpu() - lr.grad).abs().max() / lr.grad.abs().max())
    assert err <= 4e-3, err                                 # bf16 rounding of the stored dlogits
    # without the mix keys the same entry point is the plain-label loss
    lr2 = logits.float().requires_grad_(True)
    ref2 = O.loss_fn(lr2, labels, classes, alpha)
    dl2, st2 = ops.softmax_xent(logits.cuda(), labels.cuda(), alpha, 1.0 / B)
    assert abs(float(st2[0]) / B - float(ref2)) <= 1e-5 * abs(float(ref2))


@pytest.mark.gpu
def test_device_schedule_matches_optax_restatement():
    """sav_schedule_step against the oracle's optax.warmup_cosine_decay_schedule and Adam bias
    corrections: count read before the increment, t = count + 1."""
    from sav import ops
    from sav.train import WarmupCosine
    init, peak, warm, decay, end = 0.0, 2.5e-4, 5, 30, 1e-5
    ref = O.warmup_cosine_decay_schedule(init, peak, warm, decay, end)
    host = WarmupCosine(init, peak, warm, decay, end)
    step = torch.zeros(1, dtype=torch.int64, device="cuda")
    hyper = torch.zeros(3, device="cuda")
    b1, b2 = 0.9, 0.999
    for count in range(40):
        ops.schedule_step(step, hyper, 1, init, peak, warm, decay, end, b1, b2)
        h = hyper.cpu()
        assert int(step.item()) == count + 1
        assert abs(float(h[0]) - ref(count)) <= 2e-6 * peak + 1e-12, (count, float(h[0]), ref(count))
        assert abs(host(count) - ref(count)) <= 1e-12
        assert abs(float(h[1]) - (1 - b1 ** (count + 1))) <= 1e-6
        # fp32 like optax (1 - decay**count in float32): two ulps of the power
        import numpy as np
        bc2 = np.float32(1) - np.power(np.float32(b2), np.float32(count + 1))
        assert abs(float(h[2]) - float(bc2)) <= 2.4e-7
    assert ref(0) == 0.0 and abs(ref(warm) - peak) < 1e-12 and abs(ref(10 ** 6) - end) < 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("lr_kind", ["cosine", "callable"])
def test_unsynchronised_graphed_steps_equal_eager_synchronised_steps(lr_kind):
    """N graph replays issued back to back with NO host synchronisation must leave the parameters
    where N eager, fully synchronised steps leave them: the schedule / bias corrections must not be
    read from host memory that a later iteration has already overwritten."""
    from sav.train import GraphedTrainStep, WarmupCosine, train_step
    N = 12

    def lr():
        if lr_kind == "cosine":
            return WarmupCosine(0.0, 3e-3, 4, 20, 1e-5)
        return lambda count: 3e-3 * (0.5 ** count)       # changes fast: a stale value would show

    s1, x1, y1, _ = _tiny_state(lr())
    for _ in range(N + 2):                                # GraphedTrainStep runs 2 eager warm-ups
        s1.advance_schedule()
        train_step(s1, x1, y1, 0.1, None)
        torch.cuda.synchronize()
    s2, x2, y2, _ = _tiny_state(lr())
    g = GraphedTrainStep(s2, x2, y2, 0.1, None, warmup=2)
    for _ in range(N):
        g()                                               # no sync, no .item()
    torch.cuda.synchronize()
    assert s1.step == s2.step == N + 2
    assert int(s1.step_dev.item()) == int(s2.step_dev.item()) == N + 2
    a, b = s1.arena.master, s2.arena.master
    # Two separately built states: dW split-K partials are combined by fp32 reduce-adds in arrival
    # order, and Adam's m / sqrt(v) turns last-bit gradient differences of near-zero gradients into
    # O(lr) parameter differences on a few elements -- so compare in relative L2.  A schedule that
    # is stale by ONE step moves every parameter by sum_t |lr_t - lr_{t+1}| ~ 3e-3 against an rms
    # parameter of ~0.07: relative L2 >= 4e-2.
    l2 = float((a - b).norm() / a.norm())
    print("unsynchronised vs eager: relative L2", l2, "max", float((a - b).abs().max()))
    assert l2 <= 4e-3, l2
    assert float((s1.hyper - s2.hyper).abs().max()) <= 1e-7 * 3e-3 + 1e-9


@pytest.mark.gpu
def test_graphed_cait_step_resamples_drop_path_masks():
    """CaiT (stoch_depth_rate > 0) trains through the shipped graphed step, and every replay draws
    new per-sample masks from the device Philox stream keyed by the device step counter."""
    import sav
    from models import CaiT
    from sav import rng
    from sav.train import HWCN, GraphedTrainStep, TrainState
    from oracle import variants_oracle as V
    cfg = dict(num_layers=2, num_heads=4, embed_dim=192, layerscale_eps=1e-5, num_layers_token_only=2,
               patch_shape=(16, 16), num_classes=1000, expand_ratio=4)
    B, img = 16, 64
    images, labels = O.synthetic_batch(B, img_size=img, seed=0)
    model = CaiT(num_classes=1000, num_layers=2, num_layers_token_only=2, num_heads=4, embed_dim=192,
                 patch_shape=(16, 16), stoch_depth_rate=0.3, layerscale_eps=1e-5)
    x = HWCN(images.cuda())
    variables = sav.load_params(model, V.synthetic_cait_params(cfg, img_size=img, seed=42, layerscale=0.1),
                                x, is_training=False)
    state = TrainState(model, variables, lr=0.0, weight_decay=0.0, clip_grad=None, seed=77)
    g = GraphedTrainStep(state, x, labels.cuda(), 0.1, None, warmup=2)
    losses = []
    for _ in range(6):
        losses.append(round(g().item(), 5))
    # lr = 0: the parameters never move, so different losses can only come from different masks
    assert len(set(losses)) >= 4, losses
    # and the masks are the documented stream: (seed, step counter, call index)
    f_a = rng.drop_path_factors(77, 3, 2, B, 0.7)
    f_b = rng.drop_path_factors(77, 4, 2, B, 0.7)
    assert not torch.equal(f_a[0], f_b[0]) and not torch.equal(f_a[0], f_a[1])
    assert torch.equal(f_a[0], rng.drop_path_factors(77, 3, 1, B, 0.7)[0])
    vals = set(torch.cat(f_a + f_b).cpu().tolist())
    assert vals <= {0.0, float(torch.tensor(1.0 / 0.7, dtype=torch.float32))} and len(vals) == 2
    # keep-probability sanity over a long draw
    big = rng.drop_path_factors(1, 0, 1, 100_000, 0.7)[0]
    assert abs(float((big > 0).float().mean()) - 0.7) < 0.01


@pytest.mark.gpu
def test_eval_step_matches_oracle():
    """train.py:112-120: is_training=False forward, plain one-hot CE, mean."""
    from sav.train import eval_step
    state, x, y, cfg = _tiny_state(1e-3, batch=6)
    loss = eval_step(state, x, y)
    images, labels = O.synthetic_batch(6, img_size=64, seed=0)
    params = O.synthetic_vit_params(cfg, 64, seed=42)
    logits = O.vit_forward(params, O.hwcn_to_nhwc(images), cfg["num_layers"], cfg["num_heads"],
                           cfg["patch_shape"])
    ref = float(O.loss_fn(logits, labels, 1000, 0.0))
    assert abs(loss.item() - ref) <= 2e-2 * abs(ref), (loss.item(), ref)


@pytest.mark.gpu
def test_batchnorm_model_trains_under_graph_capture_with_running_averages(tmp_path):
    """CvT / BoTNet / CeiT carry flax's 'batch_stats' collection.  The reference's train.py applies the
    model without it (and so cannot train them); TrainState here keeps the running averages,
    the captured step updates them in place on the device (momentum 0.9, leff.py:35-39), eval_step
    normalises with them (use_running_average), and the checkpoint stores them beside the parameters.
    Also: a state WITHOUT the collection (the reference's call shape) still captures -- its scratch
    statistics are created once, not by a host -> device copy inside the graph."""
    import sav
    from models import CeiT
    from oracle import variants_oracle as V
    from sav import checkpoint as C
    from sav.train import HWCN, GraphedTrainStep, TrainState, eval_step
    cfg = V.ceit_config("ceit_t")
    cfg["num_layers"] = 2
    params = V.synthetic_ceit_params(cfg, seed=7)
    images, labels = O.synthetic_batch(8, img_size=224, seed=0)
    x, y = HWCN(images.cuda()), labels.cuda()

    def make(with_stats):
        model = CeiT(num_classes=1000, num_layers=2, num_heads=cfg["num_heads"], embed_dim=cfg["embed_dim"])
        variables = sav.load_params(model, params, x, is_training=True)
        if not with_stats:
            variables = {"params": variables["params"]}
        return TrainState(model, variables, lr=1e-5, weight_decay=1e-4, clip_grad=1.0)

    st = make(True)
    key = ("Image2TokenBlock_0", "BatchNorm_0")
    ra_mean = st.batch_stats[key[0]][key[1]]["mean"]
    ra_var = st.batch_stats[key[0]][key[1]]["var"]
    assert float(ra_mean.abs().max()) == 0.0 and float((ra_var - 1).abs().max()) == 0.0
    # batch statistics of the stem convolution from the oracle: the running averages must move
    # towards them as 1 - 0.9^t
    conv = V.conv2d_nobias(O.hwcn_to_nhwc(images), params["Image2TokenBlock_0"]["Conv_0"]["kernel"], 2, 4)
    b_mean, b_var = conv.mean(dim=(0, 1, 2)), conv.var(dim=(0, 1, 2), unbiased=False)
    step = GraphedTrainStep(st, x, y, 0.1, warmup=1)          # 1 eager step (capturing executes nothing)
    t = 1
    for _ in range(3):
        loss = step()
        t += 1
    torch.cuda.synchronize()
    assert math.isfinite(loss.item())
    w = 1.0 - 0.9 ** t
    # (lr = 1e-5: the convolution kernel barely moves over the four steps)
    assert float((ra_mean.cpu() - w * b_mean).abs().max()) <= 0.02 * float(b_mean.abs().max()) + 1e-3
    assert float((ra_var.cpu() - (w * b_var + (1 - w))).abs().max()) <= 0.02 * float(b_var.abs().max()) + 1e-3
    # eval mode normalises with the running averages: finite, and different from train mode's value
    ev = eval_step(st, x, y).item()
    assert math.isfinite(ev)
    # checkpoint round trip keeps the collection beside the parameters (reference layout: whole
    # variables under `params`)
    C.save_checkpoint(str(tmp_path), st, int(st.step))
    sd = C.restore_checkpoint(str(tmp_path))
    assert set(sd["params"]) == {"params", "batch_stats"}
    st2 = make(True)
    C.restore_checkpoint(str(tmp_path), st2)
    assert torch.equal(st2.batch_stats[key[0]][key[1]]["mean"], ra_mean)
    assert abs(eval_step(st2, x, y).item() - ev) <= 1e-6 * abs(ev)
    # the reference's call shape (no collection) still captures and replays
    st3 = make(False)
    step3 = GraphedTrainStep(st3, x, y, 0.1, warmup=1)
    l3 = step3()
    torch.cuda.synchronize()
    assert math.isfinite(l3.item())
