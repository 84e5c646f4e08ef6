"""The graphed training step fed from pinned host memory through the double-buffered staging API
(GraphedTrainStep.prefetch / step_prefetched) must produce the losses of the same steps run on
device-resident inputs: the staging only moves bytes (train.py:80-91 semantics unchanged).

Two separately built training states are compared, and the weight-gradient GEMMs combine their
split-K partial tiles with fp32 reduce-adds in arrival order, so the parameters after a step agree
to fp32 rounding, not bit for bit: the losses are compared to 1e-3 relative (they differ by ~6e-6;
feeding the wrong batch changes them by tens of percent: 4.39 / 7.06 / 6.92 for the three batches)."""
import pytest
import torch


def _setup(seed_params=42):
    from models import create_model
    from oracle import vit_oracle as O
    from sav.train import HWCN, GraphedTrainStep, TrainState
    cfg = O.vit_config("deit_ti_patch16")
    model = create_model("deit_ti_patch16")
    batches = [O.synthetic_batch(4, 64, seed=s) for s in range(3)]
    images = batches[0][0].cuda()
    labels = batches[0][1].cuda()
    variables = model.init({"params": 0}, HWCN(images), is_training=True)
    variables = {"params": variables["params"].arena.load_tree(O.synthetic_vit_params(cfg, 64, seed=seed_params))}
    state = TrainState(model, variables, lr=1e-3, weight_decay=1e-4, clip_grad=1.0)
    g = GraphedTrainStep(state, HWCN(images), labels, 0.1, None)
    return g, images, labels, batches


@pytest.mark.gpu
def test_prefetched_steps_equal_device_resident_steps():
    g1, images, labels, batches = _setup()
    direct = []
    for im, lb in batches:
        images.copy_(im.cuda())
        labels.copy_(lb.cuda())
        direct.append(float(g1().item()))
    g2, _, _, _ = _setup()
    host = [(im.pin_memory(), lb.pin_memory()) for im, lb in batches]
    fed = []
    g2.prefetch(*host[0])
    for i in range(len(host)):
        loss = g2.step_prefetched()
        if i + 1 < len(host):
            g2.prefetch(*host[i + 1])      # overlaps the step just launched
        fed.append(float(loss.item()))
    assert len({round(v, 1) for v in direct}) == 3          # the batches are distinguishable
    assert all(abs(a - b) <= 1e-3 * abs(b) for a, b in zip(fed, direct)), (fed, direct)


@pytest.mark.gpu
def test_prefetch_protocol_errors():
    g, _, _, batches = _setup()
    im, lb = batches[0][0].pin_memory(), batches[0][1].pin_memory()
    with pytest.raises(RuntimeError):
        g.step_prefetched()
    g.prefetch(im, lb)
    with pytest.raises(RuntimeError):
        g.prefetch(im, lb)
    g.step_prefetched()
    torch.cuda.synchronize()
