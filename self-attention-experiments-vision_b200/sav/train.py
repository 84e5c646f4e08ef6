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
    return {"params": arena.load_tree(tree)}


def make_loss_fn(model, label_smoothing=0.0, world_size=1):
    """train.py:79-92."""
    def loss_fn(params, images, labels):
        logits = model.apply({"params": params}, images, is_training=True)
        loss = F.softmax_cross_entropy(logits, labels, label_smoothing, world_size)
        return loss, logits
    return loss_fn


class TrainState:
    """flax.training.train_state.TrainState stand-in (train.py:33-37): step, apply_fn, params,
    optimizer hyper-parameters of the optax chain in train.py:25-27."""

    def __init__(self, model, variables, lr=5e-4, weight_decay=1e-4, clip_grad=None,
                 b1=0.9, b2=0.999, eps=1e-8):
        self.model = model
        self.apply_fn = model.apply
        self.params = variables["params"]
        self.arena = self.params.arena
        self.arena.ensure_moments()
        self.step = 0
        self.lr, self.wd, self.clip = lr, weight_decay, clip_grad
        self.b1, self.b2, self.eps = b1, b2, eps
        self.sumsq = torch.zeros(1, dtype=torch.float32, device=self.arena.device)
        # {lr, 1-b1^t, 1-b2^t}: lives on the device so a captured step can be replayed
        self.hyper = torch.zeros(3, dtype=torch.float32, device=self.arena.device)
        self._hyper_host = torch.zeros(3, dtype=torch.float32).pin_memory()

    def advance_schedule(self):
        """Host side of scale_by_schedule / Adam bias correction for the NEXT apply_gradients;
        enqueue before the (possibly graph-replayed) step."""
        self.step += 1
        lr_t = self.lr(self.step) if callable(self.lr) else self.lr
        self._hyper_host[0] = float(lr_t)
        self._hyper_host[1] = 1.0 - self.b1 ** self.step
        self._hyper_host[2] = 1.0 - self.b2 ** self.step
        self.hyper.copy_(self._hyper_host, non_blocking=True)

    def apply_gradients(self):
        """clip -> Adam -> +wd*p -> *(-lr) -> apply (train.py:25-27,100), one pass over the arena.
        Reads lr / bias corrections from self.hyper (call advance_schedule() first)."""
        a = self.arena
        use_clip = self.clip is not None and self.clip > 0
        if use_clip:
            ops.fill(self.sumsq, 0.0)
            ops.sumsq(a.grad, self.sumsq)
        ops.adam_step(a.master, a.grad, a.m, a.v, a.shadow, self.sumsq,
                      float(self.clip) if use_clip else 0.0, 0.0, self.b1, self.b2,
                      self.eps, self.wd, 1, hyper=self.hyper)


def train_step(state, images, labels, label_smoothing=0.0, dp=None):
    """train.py:77-109: loss + grads -> pmean over ranks -> apply_gradients.  Returns the LossVar
    (device-resident loss sum and top-1 count)."""
    world = dp.world_size if dp is not None else 1
    loss_fn = make_loss_fn(state.model, label_smoothing, world)
    # dp is the gradient hook: finished buckets are all-reduced while backward is still running
    (loss, _logits), _grads = value_and_grad(loss_fn, has_aux=True, grad_hook=dp)(
        state.params, images, labels)
    state.apply_gradients()
    return loss


class GraphedTrainStep:
    """One full training step (patchify .. optimizer) captured into a CUDA graph and replayed:
    launch-bound host work disappears (SURVEY 3.1's XLA executable, done with streams + graphs)."""

    def __init__(self, state, images, labels, label_smoothing=0.0, dp=None, warmup=2):
        self.state, self.images, self.labels = state, images, labels
        self.stream = torch.cuda.Stream()
        self.stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.stream):
            for _ in range(warmup):
                state.advance_schedule()
                self.loss = train_step(state, images, labels, label_smoothing, dp)
        torch.cuda.current_stream().wait_stream(self.stream)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        from . import _abi
        n0 = _abi.launch_count()
        with torch.cuda.graph(self.graph, stream=self.stream):
            self.loss = train_step(state, images, labels, label_smoothing, dp)
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
