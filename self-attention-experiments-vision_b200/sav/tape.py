"""Define-by-run reverse-mode tape (the stand-in for jax.value_and_grad, train.py:94-95).

Every forward op appends a closure; `backward()` runs them in reverse.  Gradient fan-in is never a
separate add kernel on the hot path: closures hand the already-accumulated gradient of their input
to the producing kernel (`dx_add` of LayerNorm-backward, `residual` of the dX GEMM epilogue).
"""
import threading

_state = threading.local()


def current_tape():
    return getattr(_state, "tape", None)


class Tape:
    def __init__(self):
        self.nodes = []

    def record(self, fn):
        self.nodes.append(fn)

    def __enter__(self):
        self._prev = current_tape()
        _state.tape = self
        return self

    def __exit__(self, *exc):
        _state.tape = self._prev

    def backward(self):
        nodes, self.nodes = self.nodes, []
        for fn in reversed(nodes):
            fn()


class Var:
    """A device activation (bf16 unless noted) plus its gradient slot."""
    __slots__ = ("_v", "g", "requires_grad", "pre_act", "g_is_pre", "shape3")

    def __init__(self, v=None, requires_grad=True):
        self._v = v
        self.g = None
        self.requires_grad = requires_grad
        self.pre_act = None    # bf16 pre-activation saved by a fused GELU epilogue
        self.g_is_pre = False  # g already multiplied by gelu'(pre_act)
        self.shape3 = None

    @property
    def v(self):
        return self._v

    @property
    def shape(self):
        return tuple(self.v.shape)

    @property
    def ndim(self):
        return self.v.dim()

    @property
    def dtype(self):
        return self.v.dtype

    def take_grad(self):
        g, self.g = self.g, None
        return g

    def __add__(self, other):
        from . import functional as F
        return F.add(self, other)

    __radd__ = __add__

    def __getitem__(self, idx):
        from . import functional as F
        return F.getitem(self, idx)


def record(fn):
    t = current_tape()
    if t is not None:
        t.record(fn)


def value_and_grad(fn, has_aux=False):
    """value_and_grad(loss_fn)(params) -> (value, grads).  `params` must be a ParamTree (arena-backed);
    grads is a ParamTree of views into the arena's fp32 gradient buffer.  The loss function must
    return a `LossVar` produced by functional.softmax_cross_entropy (it seeds the backward)."""
    def wrapped(params, *args, **kw):
        arena = params.arena
        arena.begin_grad()
        with Tape() as tape:
            out = fn(params, *args, **kw)
            loss = out[0] if has_aux else out
            loss.seed()
            tape.backward()
        grads = arena.grad_tree()
        return ((loss, out[1]), grads) if has_aux else (loss, grads)
    return wrapped
