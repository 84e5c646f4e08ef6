"""Training-step host logic mirroring train.py:22-38,77-109: state creation, the loss function,
value_and_grad, data-parallel gradient mean, and the optax chain as one fused optimizer kernel."""
import torch

from . import functional as F
from . import ops
from .tape import value_and_grad


class HWCN:
    """Marks a raw fp32 image batch in the reference's loader layout [H,W,C,N] (train.py:80) so the
    'H W C N -> N H W C' rearrange + bf16 cast are absorbed by the patchify kernel."""

    def __init__(self, images):
        self.images = images
        self.shape = (images.shape[3], images.shape[0], images.shape[1], images.shape[2])


def load_params(model, tree, *example_args, device="cuda", **example_kw):
    """Build the model's arena with init() and copy a foreign nested dict (same names/shapes,
    e.g. the oracle's or a reference checkpoint's) into it.  Returns {'params': ParamTree}."""
    variables = model.init({"params": 0}, *example_args, device=device, **example_kw)
    arena = variables["params"].arena
    out = {"params": arena.load_tree(tree)}
    if "batch_stats" in variables:      # BatchNorm running averages as init() made them (0 / 1)
        out["batch_stats"] = variables["batch_stats"]
    return out


class WarmupCosine:
    """optax.warmup_cosine_decay_schedule(init_value, peak_value, warmup_steps, decay_steps,
    end_value) (train.py:215-220).  Evaluated on the device by sav_schedule_step inside the
    captured step; __call__ is the same formula on the host (logging, tests)."""

    def __init__(self, init_value, peak_value, warmup_steps, decay_steps, end_value=0.0):
        self.init_value, self.peak_value = float(init_value), float(peak_value)
        self.warmup_steps, self.decay_steps = int(warmup_steps), int(decay_steps)
        self.end_value = float(end_value)

    def __call__(self, count):
        import math
        count = int(count)
        if count < self.warmup_steps:
            frac = 1.0 - count / self.warmup_steps
            return (self.init_value - self.peak_value) * frac + self.peak_value
        D = max(self.decay_steps - self.warmup_steps, 1)
        c = min(count - self.warmup_steps, D)
        cosv = 0.5 * (1.0 + math.cos(math.pi * c / D))
        alpha = self.end_value / self.peak_value if self.peak_value != 0.0 else 0.0
        return self.peak_value * ((1.0 - alpha) * cosv + alpha)


def warmup_cosine_decay_schedule(init_value, peak_value, warmup_steps, decay_steps, end_value=0.0):
    """Name-compatible with the optax call in train.py:215-220."""
    return WarmupCosine(init_value, peak_value, warmup_steps, decay_steps, end_value)


def reference_schedule(lr, batch_size, steps_per_epoch):
    """train.py:214-220: base_lr = lr * bs / 512, 5 warm-up epochs, cosine to epoch 30, end 1e-5."""
    return WarmupCosine(0.0, lr * (batch_size / 512), 5 * steps_per_epoch, 30 * steps_per_epoch, 1e-5)


def make_loss_fn(model, label_smoothing=0.0, world_size=1, rngs=None, reference_double_scale=False,
                 batch_stats=None):
    """train.py:79-92.  `batch` form: loss_fn(params, images, labels, mix_labels=None, ratio=None)
    (the 'mix_labels' / 'ratio' keys of train.py:84-87).  reference_double_scale reproduces the
    reference's `loss / device_count` FOLLOWED by pmean (net 1/world^2); the default is
    the plain global mean.  batch_stats: the BatchNorm running averages of a CvT / BoTNet / CeiT
    state; they are updated in place on the device (apply(..., mutable=['batch_stats']) -- the
    reference's own train.py omits the collection and cannot train those models)."""
    def loss_fn(params, images, labels, mix_labels=None, ratio=None):
        if batch_stats is not None:
            logits, _ = model.apply({"params": params, "batch_stats": batch_stats}, images,
                                    is_training=True, rngs=rngs, mutable=["batch_stats"])
        else:
            logits = model.apply({"params": params}, images, is_training=True, rngs=rngs)
        w = world_size * world_size if reference_double_scale else world_size
        loss = F.softmax_cross_entropy(logits, labels, label_smoothing, w, mix_labels, ratio)
        return loss, logits
    return loss_fn


_HOST_RING = 16


class TrainState:
    """flax.training.train_state.TrainState stand-in (train.py:33-37): step, apply_fn, params,
    optimizer hyper-parameters of the optax chain in train.py:25-27.

    The optax `count` lives ON THE DEVICE (`step_dev`, int64): every apply_gradients() enqueues
    sav_schedule_step, which increments it and writes {lr(count), 1-b1^t, 1-b2^t} for the Adam
    kernel behind it.  A captured step therefore advances its own schedule on replay and never
    reads host memory at execution time (no pinned-buffer race in unsynchronised loops).  lr may be
    a float, a WarmupCosine (both evaluated on the device) or any Python callable count -> lr; the
    callable's values travel through a ring of pinned slots, each guarded by an event before it is
    reused."""

    def __init__(self, model, variables, lr=5e-4, weight_decay=1e-4, clip_grad=None,
                 b1=0.9, b2=0.999, eps=1e-8, seed=0):
        self.model = model
        self.apply_fn = model.apply
        self.params = variables["params"]
        # flax 'batch_stats' collection (BatchNorm running averages) when the model has one
        self.batch_stats = variables.get("batch_stats") if isinstance(variables, dict) else None
        self.arena = self.params.arena
        self.arena.ensure_moments()
        self.step = 0                       # host mirror of step_dev
        self.lr, self.wd, self.clip = lr, weight_decay, clip_grad
        self.b1, self.b2, self.eps = b1, b2, eps
        dev = self.arena.device
        self.sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.hyper = torch.zeros(3, dtype=torch.float32, device=dev)
        self.step_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        from .rng import DeviceRng
        self.rngs = {"stochastic_depth": DeviceRng(seed, self.step_dev)}
        self._ring = None

    @property
    def _mode(self):
        if isinstance(self.lr, WarmupCosine):
            return 1
        return 2 if callable(self.lr) else 0

    def set_step(self, step):
        """Restore the optax count (checkpoint resume)."""
        self.step = int(step)
        self.step_dev.fill_(int(step))

    def advance_schedule(self):
        """Host bookkeeping for the NEXT apply_gradients (possibly a graph replay).  With a device
        schedule this only moves the host mirror; an arbitrary callable's lr(count) is staged
        through an event-guarded pinned ring."""
        count = self.step
        self.step += 1
        if self._mode == 2:
            if self._ring is None:
                self._ring = [(torch.zeros(1, dtype=torch.float32).pin_memory(), torch.cuda.Event())
                              for _ in range(_HOST_RING)]
                self._ring_i = 0
            slot, ev = self._ring[self._ring_i % _HOST_RING]
            if self._ring_i >= _HOST_RING:
                ev.synchronize()            # the copy that last read this slot has executed
            self._ring_i += 1
            slot[0] = float(self.lr(count))
            self.hyper[0:1].copy_(slot, non_blocking=True)
            ev.record(torch.cuda.current_stream())

    def current_lr(self):
        """lr the last enqueued step used (host formula; optax evaluates the schedule at count-1+... =
        the count before the increment)."""
        c = max(self.step - 1, 0)
        return float(self.lr(c)) if callable(self.lr) else float(self.lr)

    def apply_gradients(self):
        """clip -> Adam -> +wd*p -> *(-lr) -> apply (train.py:25-27,100), one pass over the arena."""
        a = self.arena
        lr = self.lr
        if isinstance(lr, WarmupCosine):
            ops.schedule_step(self.step_dev, self.hyper, 1, lr.init_value, lr.peak_value,
                              lr.warmup_steps, lr.decay_steps, lr.end_value, self.b1, self.b2)
        else:
            ops.schedule_step(self.step_dev, self.hyper, 2 if callable(lr) else 0, 0.0,
                              0.0 if callable(lr) else float(lr), 0, 0, 0.0, self.b1, self.b2)
        use_clip = self.clip is not None and self.clip > 0
        if use_clip:
            ops.fill(self.sumsq, 0.0)
            ops.sumsq(a.grad, self.sumsq)
        ops.adam_step(a.master, a.grad, a.m, a.v, a.shadow, self.sumsq,
                      float(self.clip) if use_clip else 0.0, 0.0, self.b1, self.b2,
                      self.eps, self.wd, 1, hyper=self.hyper)


def train_step(state, images, labels, label_smoothing=0.0, dp=None, mix_labels=None, ratio=None):
    """train.py:77-109: loss + grads -> pmean over ranks -> apply_gradients.  Returns the LossVar
    (device-resident loss sum and top-1 count).  mix_labels / ratio: the optional mixup keys of the
    batch (train.py:84-87)."""
    world = dp.world_size if dp is not None else 1
    double = bool(getattr(dp, "reference_double_scale", False))
    loss_fn = make_loss_fn(state.model, label_smoothing, world, rngs=state.rngs,
                           reference_double_scale=double, batch_stats=getattr(state, "batch_stats", None))
    # dp is the gradient hook: finished buckets are all-reduced while backward is still running
    (loss, _logits), _grads = value_and_grad(loss_fn, has_aux=True, grad_hook=dp)(
        state.params, images, labels, mix_labels, ratio)
    state.apply_gradients()
    return loss


class EvalLoss:
    """Device-resident eval loss: psum_r(mean_b(loss) / world) = sum of per-sample losses over the
    global batch / (B * world).  .item() syncs."""

    def __init__(self, loss_sum, denom, stats):
        self.loss_sum, self.denom, self.stats = loss_sum, denom, stats

    def item(self):
        return float(self.loss_sum.item()) / self.denom


def eval_step(state, images, labels, dp=None):
    """train.py:112-120: logits with is_training=False, plain one-hot softmax cross entropy (no
    smoothing, no mixup), mean over the local batch, / device_count, psum over ranks.  The
    collective (train.py:120) is one all-reduce of the per-rank loss sum."""
    variables = {"params": state.params}
    if getattr(state, "batch_stats", None) is not None:
        variables["batch_stats"] = state.batch_stats      # BatchNorm: use_running_average (eval mode)
    logits = state.apply_fn(variables, images, is_training=False)
    lv = logits.v
    B = lv.shape[0]
    _dl, stats = ops.softmax_xent(lv.contiguous(), labels, 0.0, 0.0)
    world = dp.world_size if dp is not None else 1
    loss_sum = stats[0:1]
    if dp is not None and world > 1:
        loss_sum = loss_sum.clone()
        dp.all_reduce_sum(loss_sum)
    return EvalLoss(loss_sum, B * world, stats)


class GraphedTrainStep:
    """One full training step (patchify .. optimizer) captured into a CUDA graph and replayed:
    launch-bound host work disappears (SURVEY 3.1's XLA executable, done with streams + graphs)."""

    def __init__(self, state, images, labels, label_smoothing=0.0, dp=None, warmup=2,
                 mix_labels=None, ratio=None):
        self.state, self.images, self.labels = state, images, labels
        self.mix_labels, self.ratio = mix_labels, ratio
        self.stream = torch.cuda.Stream()
        self.stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.stream):
            for _ in range(warmup):
                state.advance_schedule()
                self.loss = train_step(state, images, labels, label_smoothing, dp, mix_labels, ratio)
        torch.cuda.current_stream().wait_stream(self.stream)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        from . import _abi
        n0 = _abi.launch_count()
        with torch.cuda.graph(self.graph, stream=self.stream):
            self.loss = train_step(state, images, labels, label_smoothing, dp, mix_labels, ratio)
        self.launches_per_step = _abi.launch_count() - n0

    def __call__(self):
        self.state.advance_schedule()
        self.graph.replay()
        return self.loss

    # ---- host-fed steps.  The graph reads fixed device buffers, so a host batch goes through one
    # of two staging slots: prefetch() copies pinned host memory into the free slot on a copy
    # stream (a DMA engine, no SMs), step_prefetched() moves the staged batch into the graph's
    # input buffers device-to-device (154 MB: ~50 us) and replays.  Calling prefetch() for batch
    # i+1 right after launching step i hides the PCIe transfer (2.4 ms at batch 256) behind the
    # step, which is what the reference's loader does on the host side (input_pipeline.py:344).
    def _staging(self):
        if getattr(self, "_stage", None) is None:
            img = self.images.images if isinstance(self.images, HWCN) else self.images
            self._img = img
            self._stage = [(torch.empty_like(img), torch.empty_like(self.labels)) for _ in range(2)]
            self._copy_stream = torch.cuda.Stream()
            self._ready = [torch.cuda.Event(), torch.cuda.Event()]
            self._consumed = [torch.cuda.Event(), torch.cuda.Event()]
            self._slot, self._pending = 0, None
        return self._stage

    def prefetch(self, images_host, labels_host):
        """Start copying the next batch (pinned host tensors, same shapes / dtypes as the graph's
        inputs) into the free staging slot; returns immediately."""
        self._staging()
        if self._pending is not None:
            raise RuntimeError("prefetch(): the previously staged batch has not been consumed")
        s = self._slot
        with torch.cuda.stream(self._copy_stream):
            self._copy_stream.wait_event(self._consumed[s])   # the step that read this slot is past it
            self._stage[s][0].copy_(images_host, non_blocking=True)
            self._stage[s][1].copy_(labels_host, non_blocking=True)
            self._ready[s].record(self._copy_stream)
        self._pending = s
        self._slot ^= 1

    def step_prefetched(self):
        """One training step on the batch staged by the last prefetch()."""
        self._staging()
        s = self._pending
        if s is None:
            raise RuntimeError("step_prefetched(): call prefetch() first")
        cur = torch.cuda.current_stream()
        cur.wait_event(self._ready[s])
        self._img.copy_(self._stage[s][0], non_blocking=True)
        self.labels.copy_(self._stage[s][1], non_blocking=True)
        self._consumed[s].record(cur)
        self._pending = None
        return self()
