"""GPU checks of the 'next' rows of SURVEY 8f that are built: top-k correctness (utils.py:20-37)
against a literal argsort / isin restatement, and the checkpoint round trip (train.py:123-127)
through a real TrainState (params, Adam moments, step)."""
import pytest
import torch


def _topk_ref(logits, labels, ks):
    """utils.py:20-31 literally: argsort ascending (stable), last k indices, isin(label)."""
    order = torch.argsort(logits, dim=-1, stable=True)
    return {k: (order[:, -k:] == labels[:, None]).any(-1).float() for k in ks}


@pytest.mark.gpu
def test_topk_correct_matches_argsort_restatement_including_ties():
    from utils import topk_correct
    g = torch.Generator().manual_seed(0)
    B, classes = 67, 1000
    logits = torch.randn((B, classes), generator=g).bfloat16()
    logits[:, 10:20] = logits[:, 10:11]          # exact ties (bf16 makes them common anyway)
    labels = torch.randint(0, classes, (B,), generator=g, dtype=torch.int32)
    labels[:10] = torch.arange(10, 20, dtype=torch.int32)   # labels inside the tie group
    # make some labels the arg-max so both outcomes occur
    for b in range(20, 40):
        logits[b, labels[b]] = 10.0
    mask = (torch.rand((B,), generator=g) > 0.3).float()
    out = topk_correct(logits.cuda(), labels.cuda(), mask=mask.cuda(), prefix="val_", topk=(1, 5))
    ref = _topk_ref(logits.float(), labels.long(), (1, 5))
    assert set(out) == {"val_top_1_acc", "val_top_5_acc"}
    for k in (1, 5):
        assert torch.equal(out[f"val_top_{k}_acc"].cpu(), ref[k] * mask), k
    assert 0 < float(ref[1].sum()) < B


@pytest.mark.gpu
def test_checkpoint_roundtrip_through_train_state(tmp_path):
    import sav
    from models import create_model
    from oracle import vit_oracle as O
    from sav import checkpoint as C
    from sav.train import HWCN, TrainState, train_step
    cfg = O.vit_config("deit_ti_patch16")
    images, labels = O.synthetic_batch(4, img_size=64, seed=0)
    model = create_model("deit_ti_patch16")
    x = HWCN(images.cuda())
    variables = sav.load_params(model, O.synthetic_vit_params(cfg, 64, seed=42), x, is_training=True)
    state = TrainState(model, variables, lr=1e-3, weight_decay=1e-4, clip_grad=1.0)
    for _ in range(2):
        state.advance_schedule()
        train_step(state, x, labels.cuda(), 0.1, None)
    torch.cuda.synchronize()
    path = C.save_checkpoint(str(tmp_path), state, state.step)
    before = {k: v.clone() for k, v in (("p", state.arena.master), ("m", state.arena.m), ("v", state.arena.v))}
    # perturb everything, then restore
    state.arena.master.add_(1.0)
    state.arena.m.zero_()
    state.arena.v.zero_()
    state.set_step(0)
    sd = C.restore_checkpoint(path, state)
    torch.cuda.synchronize()
    assert state.step == 2 and int(sd["step"]) == 2 and int(state.step_dev.item()) == 2
    # every parameter leaf and moment is restored bit-exactly (fp32 -> bytes -> fp32)
    for path_, p in state.arena.params.items():
        for name, flat in (("p", state.arena.master), ("m", state.arena.m), ("v", state.arena.v)):
            assert torch.equal(p._view(flat), p._view(before[name])), (path_, name)
    # the bf16 shadow was refreshed from the restored master
    k = next(iter(state.arena.params.values()))
    assert torch.equal(k.shadow.float(), k.master.bfloat16().float())
    # the reference TrainState nesting: whole variables under params, optax.chain states by index
    assert "Encoder_0" in sd["params"]["params"] and set(sd["opt_state"]) == {"0", "1", "2", "3"}
    assert "Encoder_0" in sd["opt_state"]["1"]["mu"]["params"] and int(sd["opt_state"]["3"]["count"]) == 2
