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

    def apply_gradients(self, lr=None):
        a = self.arena
        self.step += 1
        use_clip = self.clip is not None and self.clip > 0
        if use_clip:
            ops.fill(self.sumsq, 0.0)
            ops.sumsq(a.grad, self.sumsq)
        lr_t = lr(self.step) if callable(lr) else (lr if lr is not None else
                                                   (self.lr(self.step) if callable(self.lr) else self.lr))
        ops.adam_step(a.master, a.grad, a.m, a.v, a.shadow, self.sumsq,
                      float(self.clip) if use_clip else 0.0, float(lr_t), self.b1, self.b2,
                      self.eps, self.wd, self.step)


def train_step(state, images, labels, label_smoothing=0.0, dp=None):
    """train.py:77-109: loss + grads -> pmean over ranks -> apply_gradients.  Returns the LossVar
    (device-resident loss sum and top-1 count)."""
    world = dp.world_size if dp is not None else 1
    loss_fn = make_loss_fn(state.model, label_smoothing, world)
    (loss, _logits), _grads = value_and_grad(loss_fn, has_aux=True)(state.params, images, labels)
    if dp is not None:
        dp.all_reduce_grads(state.arena)
    state.apply_gradients()
    return loss
