"""Development tool (GPU): where the graph-replayed DeiT-S training step spends its time.

Every C-ABI launch inside the captured step is followed by an external timing event (an event-record
node of the graph), so one replay yields the device time of every SLOT = kernel duration + the gap
to the next kernel, in the real step (power-capped clocks, warm/cold caches as they really are).
Comparing the slot times with the same kernels timed alone (tests/gpu_diag.py --time, ncu launch
list) splits the step into kernel time, inter-kernel gap and clock derate.

    python tests/tools/step_timeline.py [--batch 256] [--replays 10] > profiles/r02_step_timeline.txt
"""
import argparse
import collections
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "self-attention-experiments-vision_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--replays", type=int, default=10)
    ap.add_argument("--model", default="deit_s_patch16")
    ap.add_argument("--img", type=int, default=224)
    a = ap.parse_args()
    import sav  # noqa: F401
    from models import create_model
    from oracle import vit_oracle as O
    from sav import _abi
    from sav.train import HWCN, GraphedTrainStep, TrainState

    lib = _abi.lib()
    log = []          # (entry point, event) in launch order, filled while `recording`
    state = {"on": False}

    class Wrapped:
        def __init__(self, name, fn):
            self.name, self.fn = name, fn

        def __call__(self, *args):
            n0 = lib.sav_launch_count()
            rc = self.fn(*args)
            if state["on"] and lib.sav_launch_count() > n0:
                ev = torch.cuda.Event(enable_timing=True, external=True)
                ev.record(torch.cuda.current_stream())
                log.append((self.name, ev))
            return rc

    class Proxy:
        def __getattr__(self, name):
            fn = getattr(lib, name)
            return Wrapped(name, fn) if name.startswith("sav_") and name not in (
                "sav_launch_count", "sav_last_error", "sav_abi_version") else fn
    _abi._lib = Proxy()

    torch.cuda.set_device(0)
    images, labels = O.synthetic_batch(a.batch, a.img, seed=0)
    model = create_model(a.model)
    x = HWCN(images.cuda())
    try:
        cfg = O.vit_config(a.model)
        variables = sav.load_params(model, O.synthetic_vit_params(cfg, a.img, seed=42), x, is_training=True)
    except KeyError:
        # any other create_model name (cait / cvt / botnet / mixer / tnt / ceit): random parameters
        variables = model.init({"params": 0, "stochastic_depth": 1}, x, is_training=True)
        variables["params"].arena.master.normal_(0.0, 0.02)
        variables["params"].arena.refresh_shadow()
    st = TrainState(model, variables, lr=1e-3, weight_decay=1e-4, clip_grad=1.0, seed=1)
    # warm-up eagerly (unrecorded), then capture with recording on
    orig_init = GraphedTrainStep.__init__
    step = GraphedTrainStep.__new__(GraphedTrainStep)
    step.state, step.images, step.labels = st, x, labels.cuda()
    step.mix_labels = step.ratio = None
    from sav.train import train_step
    step.stream = torch.cuda.Stream()
    with torch.cuda.stream(step.stream):
        for _ in range(2):
            st.advance_schedule()
            train_step(st, x, step.labels, 0.1, None)
    torch.cuda.synchronize()
    step.graph = torch.cuda.CUDAGraph()
    start = torch.cuda.Event(enable_timing=True, external=True)
    with torch.cuda.graph(step.graph, stream=step.stream):
        start.record(torch.cuda.current_stream())
        state["on"] = True
        step.loss = train_step(st, x, step.labels, 0.1, None)
        state["on"] = False
    names = [n for n, _ in log]
    acc = [0.0] * len(log)
    total = 0.0
    for r in range(a.replays + 3):
        st.advance_schedule()
        step.graph.replay()
        torch.cuda.synchronize()
        if r < 3:
            continue
        prev = start
        for i, (_, ev) in enumerate(log):
            acc[i] += prev.elapsed_time(ev) * 1e3
            prev = ev
        total += start.elapsed_time(log[-1][1]) * 1e3
    slot = [v / a.replays for v in acc]
    total /= a.replays
    print(f"# {a.model} {a.img}^2 batch {a.batch}: {len(log)} launching ABI calls per step, step (first..last event) "
          f"{total / 1e3:.3f} ms, mean of {a.replays} replays; slot = kernel(s) of the call + gap to the next")
    by = collections.OrderedDict()
    for n, v in zip(names, slot):
        c = by.setdefault(n, [0, 0.0, 1e9, 0.0])
        c[0] += 1; c[1] += v; c[2] = min(c[2], v); c[3] = max(c[3], v)
    print(f"{'entry point':28s} {'calls':>5s} {'total us':>10s} {'share':>7s} {'mean us':>9s} {'min':>8s} {'max':>8s}")
    for n, (cnt, tot, mn, mx) in sorted(by.items(), key=lambda kv: -kv[1][1]):
        print(f"{n:28s} {cnt:5d} {tot:10.1f} {tot / total * 100:6.1f}% {tot / cnt:9.1f} {mn:8.1f} {mx:8.1f}")
    print("\n# launch order (one encoder block of the backward pass, from the middle of the step)")
    mid = len(log) // 2 + 20
    for i in range(mid, min(mid + 24, len(log))):
        print(f"{i:4d} {names[i]:28s} {slot[i]:8.1f} us")
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "step_timeline_slots.csv"), "w") as f:
        f.write("index,entry_point,slot_us\n")
        for i, (n, v) in enumerate(zip(names, slot)):
            f.write(f"{i},{n},{v:.2f}\n")


if __name__ == "__main__":
    main()
