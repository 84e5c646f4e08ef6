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
        self.pending = []

    def record(self, fn):
        self.nodes.append(fn)

    def defer(self, var):
        """A Var whose producing kernel has not been launched yet (it waits for a consumer to fuse
        with); flush() launches whatever is still pending so that every forward op is on the tape
        before the backward pass starts."""
        self.pending.append(var)

    def flush(self):
        pending, self.pending = self.pending, []
        for var in pending:
            var.v

    def __enter__(self):
        self._prev = current_tape()
        _state.tape = self
        return self

    def __exit__(self, *exc):
        _state.tape = self._prev

    def backward(self, after_node=None):
        self.flush()
        nodes, self.nodes = self.nodes, []
        for i, fn in enumerate(reversed(nodes)):
            fn()
            if after_node is not None:
                after_node(i)


class Var:
    """A device activation (bf16 unless noted) plus its gradient slot."""
    __slots__ = ("_v", "g", "g16", "requires_grad", "pre_act", "g_is_pre", "shape3")

    def __init__(self, v=None, requires_grad=True):
        self._v = v
        self.g = None      # gradient: bf16 for activations, fp32 for residual-stream Vars
        self.g16 = None    # residual-stream Vars only: bf16 copy of g (MMA operand)
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

    def take_grad16(self):
        g, self.g16 = self.g16, None
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


def value_and_grad(fn, has_aux=False, grad_hook=None):
    """value_and_grad(loss_fn)(params) -> (value, grads).  `params` must be a ParamTree (arena-backed);
    grads is a ParamTree of views into the arena's fp32 gradient buffer.  The loss function must
    return a `LossVar` produced by functional.softmax_cross_entropy (it seeds the backward).
    grad_hook (dp.DataParallel): begin(arena) / wrote(path) / after_node(i) / finish() -- lets the
    gradient all-reduce of finished buckets overlap the rest of the backward pass."""
    def wrapped(params, *args, **kw):
        arena = params.arena
        arena.begin_grad()
        arena.grad_hook = grad_hook
        if grad_hook is not None:
            grad_hook.begin(arena)
        try:
            with Tape() as tape:
                out = fn(params, *args, **kw)
                loss = out[0] if has_aux else out
                tape.flush()
                loss.seed()
                tape.backward(grad_hook.after_node if grad_hook is not None else None)
            if grad_hook is not None:
                grad_hook.finish()
        finally:
            arena.grad_hook = None
        grads = arena.grad_tree()
        return ((loss, out[1]), grads) if has_aux else (loss, grads)
    return wrapped
