"""Minimal flax.linen-compatible module surface (Flax itself is not installable here).

Provides exactly what the reference's model code and train.py touch: dataclass-style Module fields,
`@compact`, Flax auto-naming (`ClassName_index`) of submodules created inside `__call__`,
`self.param(name, init, shape)`, `self.make_rng`, and `init / apply / init_with_output` with
`{'params': tree}` variables (train.py:29-37,82,114; models/vit_test.py:25).  Parameters live in a
flat device arena (arena.py); activations are `Var`s whose arithmetic is the C-ABI kernels.
"""
import dataclasses
import math
import threading

import torch

from . import functional as F
from .arena import ArenaBuilder, ParamTree
from .tape import Var

_ctx = threading.local()


def _stack():
    if not hasattr(_ctx, "stack"):
        _ctx.stack = []
    return _ctx.stack


class _Run:
    """State of one init/apply call."""

    def __init__(self, mode, arena=None, rngs=None, device=None):
        self.mode = mode            # 'init' | 'apply'
        self.arena = arena
        self.builder = ArenaBuilder() if mode == "init" else None
        self.rngs = rngs or {}
        self.device = device
        self.init_params = {}       # path -> temporary Param-like (init mode)
        self.gens = {}
        self.batch_stats = {}       # path -> fp32 device tensor (flax 'batch_stats' collection)
        # flax updates BatchNorm running averages only under apply(..., mutable=['batch_stats']); during
        # init() the collection is created (mean 0, var 1) but not updated (is_initializing guard)
        self.mutable_stats = False
        self.had_batch_stats = False


def _run():
    return getattr(_ctx, "run", None)


def compact(fn):
    fn._compact = True
    return fn


# ----------------------------------------------------------------------------- initializers
class initializers:
    @staticmethod
    def zeros(gen, shape):
        return torch.zeros(shape)

    @staticmethod
    def ones(gen, shape):
        return torch.ones(shape)

    @staticmethod
    def normal(stddev=1e-2):
        return lambda gen, shape: torch.randn(shape, generator=gen) * stddev

    @staticmethod
    def lecun_normal(in_axes=1):
        """variance_scaling(1.0, 'fan_in', 'truncated_normal') -- flax Dense default."""
        def init(gen, shape):
            fan_in = 1
            for s in shape[:in_axes]:
                fan_in *= s
            std = math.sqrt(1.0 / max(fan_in, 1)) / 0.87962566103423978
            t = torch.empty(shape)
            torch.nn.init.trunc_normal_(t, 0.0, 1.0, -2.0, 2.0, generator=gen)
            return t * std
        return init

    @staticmethod
    def orthogonal():
        def init(gen, shape):
            a = torch.randn(shape, generator=gen)
            q, r = torch.linalg.qr(a)
            return q * torch.sign(torch.diagonal(r))
        return init

    @staticmethod
    def full(value):
        return lambda gen, shape: torch.full(shape, float(value))


class _TempParam:
    """Init-mode stand-in for arena.Param: standalone tensors, no gradients."""

    def __init__(self, master, shadow):
        self.master, self.shadow = master, shadow
        self.grad = None

    def touch_grad(self):
        return False


class Module:
    name = None

    def __init_subclass__(cls, **kw):
        super().__init_subclass__(**kw)
        ann = dict(cls.__dict__.get("__annotations__", {}))
        ann.pop("name", None)
        cls.__annotations__ = {**ann, "_kw_only": dataclasses.KW_ONLY, "name": str}
        if "name" not in cls.__dict__:
            cls.name = None
        dataclasses.dataclass(cls, eq=False, repr=False, kw_only=False)
        call = cls.__dict__.get("__call__")
        if call is not None and not getattr(call, "_wrapped", False):
            cls.__call__ = _wrap_call(call)

    def __post_init__(self):
        st = _stack()
        object.__setattr__(self, "_path", None)
        if st:
            parent = st[-1]
            if self.name is None:
                base = type(self).__name__
                idx = parent["counters"].get(base, 0)
                parent["counters"][base] = idx + 1
                object.__setattr__(self, "name", f"{base}_{idx}")
            object.__setattr__(self, "_path", parent["path"] + (self.name,))

    # ---- parameters -------------------------------------------------------------------------
    def _scope_path(self):
        return _stack()[-1]["path"]

    def param(self, name, init_fn, *shape_args):
        shape = shape_args[0] if len(shape_args) == 1 and isinstance(shape_args[0], (tuple, list)) \
            else tuple(shape_args)
        return self._param(self._scope_path() + (name,), init_fn, tuple(shape))

    def _param(self, path, init_fn, shape, fused=None):
        run = _run()
        if run.mode == "apply":
            try:
                return run.arena.params[path]
            except KeyError:
                raise KeyError(f"parameter {'/'.join(path)} not found in variables") from None
        if path in run.init_params:
            return run.init_params[path]
        gen = self._gen("params")
        if fused is None:
            val = init_fn(gen, shape)
            run.builder.add_leaf(path, shape, val)
            p = self._temp(val, run)
            run.init_params[path] = p
            return p
        raise AssertionError

    def fused_params(self, names, leaf, init_fn, shape, lead=()):
        """Allocate `len(names)` same-shaped kernels [K, *rest] as one [K, n, *rest] block so they
        can be used as a single GEMM weight; each stays visible as `<name>/<leaf>`."""
        run = _run()
        base_path = self._scope_path()
        paths = [base_path + (n, leaf) for n in names]
        K, rest = shape[0], tuple(shape[1:])
        inner = 1
        for s in rest:
            inner *= s
        n = len(names)
        if run.mode == "apply":
            ps = [run.arena.params[p] for p in paths]
            a, off = run.arena, ps[0].offset
            w2 = torch.as_strided(a.shadow, (K, n * inner), (n * inner, 1), off)
            g2 = torch.as_strided(a.grad, (K, n * inner), (n * inner, 1), off)
            return ps, F.Weight(w2, g2, ps)
        if paths[0] in run.init_params:
            return run.init_params[paths[0] + ("#fused",)]
        gen = self._gen("params")
        vals = [init_fn(gen, shape) for _ in names]
        block = torch.stack([v.reshape(K, inner) for v in vals], dim=1)  # [K, n, inner]
        base = run.builder.alloc(block.numel())
        stride = (n * inner,) + tuple(_contig(rest))
        master = self._to_dev(block, run)
        shadow = _shadow_of(master)
        out = []
        for i, (p, v) in enumerate(zip(paths, vals)):
            run.builder.add_leaf(p, tuple(lead) + tuple(shape), v.reshape(tuple(lead) + tuple(shape)),
                                 base=base, stride=(0,) * len(lead) + stride, rel_offset=i * inner)
            tp = _TempParam(master[:, i].view((K,) + rest) if not rest == () else master[:, i],
                            shadow[:, i].view((K,) + rest))
            run.init_params[p] = tp
            out.append(tp)
        res = (out, F.Weight(shadow.view(K, n * inner), None, out))
        run.init_params[paths[0] + ("#fused",)] = res
        return res

    def stacked_params(self, names, rows, width, init_fn, pad_to=8):
        """Allocate several [rows_i, width] tables back to back (row count padded to `pad_to` with
        zero rows that receive zero gradients) so that they can be ONE K-major GEMM operand
        [sum(rows) padded, width]; each stays visible as its own leaf (botnet.py:124-132)."""
        run = _run()
        base_path = self._scope_path()
        paths = [base_path + (n,) for n in names]
        total = sum(rows)
        padded = (total + pad_to - 1) // pad_to * pad_to
        if run.mode == "apply":
            ps = [run.arena.params[p] for p in paths]
            a, off = run.arena, ps[0].offset
            w2 = torch.as_strided(a.shadow, (padded, width), (width, 1), off)
            g2 = torch.as_strided(a.grad, (padded, width), (width, 1), off)
            m2 = torch.as_strided(a.master, (padded, width), (width, 1), off)
            return ps, F.Weight(w2, g2, ps, wm=m2)
        key = paths[0] + ("#stacked",)
        if key in run.init_params:
            return run.init_params[key]
        gen = self._gen("params")
        vals = [init_fn(gen, (r, width)) for r in rows]
        block = torch.zeros((padded, width))
        block[:total] = torch.cat(vals, dim=0)
        base = run.builder.alloc(padded * width)
        master = self._to_dev(block, run)
        shadow = _shadow_of(master)
        out, r0 = [], 0
        for p, v, r in zip(paths, vals, rows):
            run.builder.add_leaf(p, (r, width), v, base=base, stride=(width, 1), rel_offset=r0 * width)
            tp = _TempParam(master[r0:r0 + r], shadow[r0:r0 + r])
            run.init_params[p] = tp
            out.append(tp)
            r0 += r
        res = (out, F.Weight(shadow, None, out, wm=master))
        run.init_params[key] = res
        return res

    def variable(self, collection, name, init_fn, shape):
        """self.variable('batch_stats', name, init, shape) of flax.linen: a non-trainable fp32 state
        tensor.  Created during init(); during apply() it is taken from variables['batch_stats']."""
        if collection != "batch_stats":
            raise NotImplementedError(f"variable collection '{collection}'")
        run = _run()
        path = self._scope_path() + (name,)
        if path not in run.batch_stats:
            if run.mode == "apply" and run.had_batch_stats:
                raise KeyError(f"batch_stats/{'/'.join(path)} not found in variables")
            # apply() without the collection (train.py's call shape): a scratch state that
            # the training-mode statistics never read.  It is created once per arena and reused by
            # later apply() calls: a captured step must not issue a host -> device copy.
            cache = getattr(run.arena, "_scratch_stats", None) if run.arena is not None else None
            if cache is None and run.arena is not None:
                cache = run.arena._scratch_stats = {}
            if cache is not None and path in cache:
                run.batch_stats[path] = cache[path]
            else:
                run.batch_stats[path] = self._to_dev(init_fn(None, tuple(shape)), run)
                if cache is not None:
                    cache[path] = run.batch_stats[path]
        return run.batch_stats[path]

    def is_mutable_collection(self, collection):
        return collection == "batch_stats" and _run().mutable_stats

    def _to_dev(self, val, run):
        if run.device.type == "meta":
            return torch.empty(val.shape, dtype=torch.float32, device="meta")
        return val.to(device=run.device, dtype=torch.float32)

    def _temp(self, val, run):
        m = self._to_dev(val, run)
        return _TempParam(m, _shadow_of(m))

    def _gen(self, stream):
        run = _run()
        if stream not in run.gens:
            seed = run.rngs.get(stream)
            if seed is None:
                raise KeyError(f"{type(self).__name__} needs an rng named '{stream}'")
            g = torch.Generator()
            g.manual_seed(int(seed) if not torch.is_tensor(seed) else int(seed.flatten()[0]))
            run.gens[stream] = g
        return run.gens[stream]

    def make_rng(self, stream):
        return self._gen(stream)

    def make_device_rng(self, stream):
        """Device-side stream (sav/rng.py): rngs[stream] is an int seed or a rng.DeviceRng carrying
        the training step's device counter; one instance per apply() so call indices restart."""
        from .rng import DeviceRng
        run = _run()
        key = ("device", stream)
        if key not in run.gens:
            seed = run.rngs.get(stream)
            if seed is None:
                raise KeyError(f"{type(self).__name__} needs an rng named '{stream}'")
            if isinstance(seed, DeviceRng):
                run.gens[key] = DeviceRng(seed.seed, seed.step_dev, seed.step)
            else:
                run.gens[key] = DeviceRng(int(seed) if not torch.is_tensor(seed) else int(seed.flatten()[0]))
        return run.gens[key]

    # ---- init / apply -------------------------------------------------------------------------
    def _execute(self, run, args, kw):
        prev = _run()
        _ctx.run = run
        saved_stack = _ctx.stack if hasattr(_ctx, "stack") else []
        _ctx.stack = []
        try:
            return self(*args, **kw)
        finally:
            _ctx.run = prev
            _ctx.stack = saved_stack

    def init_with_output(self, rngs, *args, device=None, **kw):
        if not isinstance(rngs, dict):
            rngs = {"params": rngs}
        device = torch.device(device) if device is not None else _default_device(args)
        run = _Run("init", rngs=rngs, device=device)
        args = tuple(_to_run_device(a, device) for a in args)
        out = self._execute(run, args, kw)
        arena = run.builder.build(device)
        variables = {"params": arena.tree()}
        if run.batch_stats:
            variables["batch_stats"] = _nest(run.batch_stats)
        return out, variables

    def init(self, rngs, *args, **kw):
        return self.init_with_output(rngs, *args, **kw)[1]

    def apply(self, variables, *args, rngs=None, mutable=None, **kw):
        params = variables["params"] if "params" in variables else variables
        if not isinstance(params, ParamTree) or params.arena is None:
            raise TypeError("apply() needs the arena-backed tree returned by init(); use "
                            "sav.load_params(model, tree, example_input) for foreign trees")
        run = _Run("apply", arena=params.arena, rngs=rngs or {}, device=params.arena.device)
        if isinstance(variables, dict) and "batch_stats" in variables:
            run.batch_stats = _flatten(variables["batch_stats"])
            run.had_batch_stats = True
        want_stats = bool(mutable) and ("batch_stats" in mutable or mutable is True)
        run.mutable_stats = want_stats
        args = tuple(_to_run_device(a, params.arena.device) for a in args)
        out = self._execute(run, args, kw)
        if mutable:
            # flax: apply(..., mutable=[...]) returns (output, updated collections); the running
            # averages are updated in place on the device, the same tensors are handed back
            return out, ({"batch_stats": _nest(run.batch_stats)} if want_stats else {})
        return out


def _nest(flat):
    root = {}
    for path, v in flat.items():
        d = root
        for k in path[:-1]:
            d = d.setdefault(k, {})
        d[path[-1]] = v
    return root


def _flatten(tree, prefix=()):
    out = {}
    for k, v in tree.items():
        if isinstance(v, dict):
            out.update(_flatten(v, prefix + (k,)))
        else:
            out[prefix + (k,)] = v
    return out


def _shadow_of(master):
    """bf16 copy of an init-mode fp32 tensor through the library's own cast kernel (no torch op)."""
    if not master.is_cuda:
        return master.to(torch.bfloat16)     # meta (shape-only) path
    from . import ops
    out = torch.empty(master.shape, dtype=torch.bfloat16, device=master.device)
    if master.numel():
        ops.cast_bf16(master.contiguous().view(-1), out.view(-1))
    return out


def _contig(shape):
    st, acc = [], 1
    for s in reversed(shape):
        st.append(acc)
        acc *= s
    return list(reversed(st))


def _default_device(args):
    for a in args:
        if torch.is_tensor(a) and (a.is_cuda or a.is_meta):
            return a.device
    return torch.device("cuda") if torch.cuda.is_available() else torch.device("meta")


def _to_run_device(a, device):
    if torch.is_tensor(a) and a.device != device:
        if device.type == "meta":
            return torch.empty(a.shape, dtype=a.dtype, device="meta")
        return a.to(device)
    return a


def _wrap_call(call):
    def wrapped(self, *args, **kw):
        st = _stack()
        path = self._path if (st and self._path is not None) else (() if not st else st[-1]["path"] + (self.name,))
        st.append({"path": path, "counters": {}})
        try:
            return call(self, *args, **kw)
        finally:
            st.pop()
    wrapped._wrapped = True
    wrapped._compact = True
    return wrapped


# ----------------------------------------------------------------------------- stock layers
class Dense(Module):
    """flax.linen.Dense (ff.py:24,27; patch_embed.py:23-25; vit.py:96-98)."""
    features: int = 0
    use_bias: bool = True
    dtype: object = None
    kernel_init: object = None
    bias_init: object = None

    def __call__(self, inputs):
        K = inputs.shape[-1]
        kernel = self.param("kernel", self.kernel_init or initializers.lecun_normal(), (K, self.features))
        bias = self.param("bias", self.bias_init or initializers.zeros, (self.features,)) \
            if self.use_bias else None
        if isinstance(inputs, F.TokenMix):      # Dense over the token axis (mlp_mixer.py:18-23)
            return inputs.dense(kernel, bias, self.features)
        return F.linear(inputs, F.Weight(kernel.shadow, kernel.grad, [kernel], wm=kernel.master), bias)


class LayerNorm(Module):
    """flax.linen.LayerNorm (vit.py:19,26,57): epsilon 1e-6, scale ones, bias zeros."""
    epsilon: float = 1e-6
    dtype: object = None
    stream_out: bool = False     # (not a flax field) the OUTPUT is the fp32 residual stream: post-LN
                                 # blocks add their branches to the normalised tensor (ceit.py:36-44)

    def __call__(self, inputs):
        C = inputs.shape[-1]
        scale = self.param("scale", initializers.ones, (C,))
        bias = self.param("bias", initializers.zeros, (C,))
        if self.stream_out:
            return F.layernorm_stream(inputs, scale, bias, self.epsilon)
        return F.layernorm(inputs, scale, bias, self.epsilon)


class Dropout(Module):
    """flax.linen.Dropout; every reference model uses rate 0 (vit.py:68-69) => identity."""
    rate: float = 0.0
    deterministic: object = None

    def __call__(self, inputs, deterministic=None):
        det = deterministic if deterministic is not None else self.deterministic
        if self.rate == 0.0 or det:
            return inputs
        raise NotImplementedError("dropout with rate > 0 in training mode")


class _Activation:
    """Callable activation with a `kind` the fused kernels understand (BatchNorm + activation, the
    closing activation of a bottleneck block).  Called directly it is the standalone kernel."""

    def __init__(self, kind):
        self.kind = kind

    def __call__(self, x):
        return F.activation(x, self.kind)


class activation:
    gelu = staticmethod(F.gelu)
    swish = _Activation("swish")
    relu = _Activation("relu")
    sigmoid = _Activation("sigmoid")


gelu = F.gelu
swish = activation.swish
relu = activation.relu
sigmoid = activation.sigmoid


def _act_kind(fn):
    if fn is None:
        return "none"
    if fn is F.gelu:
        return "gelu"
    kind = getattr(fn, "kind", None)
    if kind is None:
        raise NotImplementedError("conv-stage activation must be one of nn.swish / nn.relu / nn.sigmoid")
    return kind


class Conv(Module):
    """flax.linen.Conv on channel-last bf16 activations (botnet.py:280-285): im2col + tcgen05 GEMM,
    or the GEMM alone for 1x1 stride-1 kernels.  kernel [kh, kw, cin, cout]."""
    features: int = 0
    kernel_size: tuple = (1, 1)
    strides: tuple = (1, 1)
    padding: object = 'SAME'
    use_bias: bool = True
    dtype: object = None
    precision: object = None
    kernel_init: object = None

    def __call__(self, inputs):
        kh, kw = self.kernel_size
        sh, sw = self.strides if not isinstance(self.strides, int) else (self.strides, self.strides)
        if kh != kw or sh != sw:
            raise NotImplementedError("nn.Conv: square kernels / equal strides only")
        cin = inputs.shape[-1]
        kernel = self.param("kernel", self.kernel_init or initializers.lecun_normal(in_axes=3),
                            (kh, kw, cin, self.features))
        pad = self.padding
        if not isinstance(pad, str):
            lo = {int(p[0]) for p in pad} | {int(p[1]) for p in pad}     # [(3, 3), (3, 3)], botnet.py:294
            if len(lo) != 1:
                raise NotImplementedError("nn.Conv: symmetric explicit padding only")
            pad = lo.pop()
        bias = self.param("bias", initializers.zeros, (self.features,)) if self.use_bias else None
        return F.conv2d(inputs, kernel, kh, sh, pad, bias=bias)


class BatchNorm(Module):
    """flax.linen.BatchNorm over the last axis (botnet.py:286-290): batch statistics (biased variance)
    unless use_running_average; running averages in the 'batch_stats' collection.  `act` fuses the
    activation that follows the norm in the reference (y = norm()(y); y = activation_fn(y))."""
    use_running_average: bool = False
    momentum: float = 0.99
    epsilon: float = 1e-5
    dtype: object = None
    scale_init: object = None
    bias_init: object = None

    def __call__(self, inputs, act=None):
        C = inputs.shape[-1]
        scale = self.param("scale", self.scale_init or initializers.ones, (C,))
        bias = self.param("bias", self.bias_init or initializers.zeros, (C,))
        ra_mean = self.variable("batch_stats", "mean", initializers.zeros, (C,))
        ra_var = self.variable("batch_stats", "var", initializers.ones, (C,))
        return F.batchnorm_act(inputs, scale, bias, self.epsilon, _act_kind(act), running=(ra_mean, ra_var),
                               momentum=self.momentum, training=not self.use_running_average,
                               update=self.is_mutable_collection("batch_stats"))
