"""Flat parameter arena: fp32 master, bf16 compute shadow, fp32 gradients (and Adam moments) as
single contiguous device buffers laid out for 180 GB HBM -- one cast kernel refreshes every
shadow, one kernel does the optimizer step, NCCL all-reduces contiguous gradient ranges.

The user-visible parameter tree keeps the reference's Flax auto-names; its leaves are
(possibly strided) views into the arena, e.g. queries/keys/values kernels [C,H,d] are views of one
fused [C,3,H,d] block so the QKV projection is a single GEMM.
"""
import torch

ALIGN = 64  # elements: keeps fp32 and bf16 views 16-byte aligned


class Param:
    __slots__ = ("path", "arena", "offset", "shape", "stride")

    def __init__(self, path, arena, offset, shape, stride):
        self.path, self.arena, self.offset = path, arena, offset
        self.shape, self.stride = tuple(shape), tuple(stride)

    def _view(self, flat):
        return torch.as_strided(flat, self.shape, self.stride, self.offset)

    @property
    def master(self):
        return self._view(self.arena.master)

    @property
    def shadow(self):
        return self._view(self.arena.shadow)

    @property
    def grad(self):
        return self._view(self.arena.grad)

    def touch_grad(self):
        """Record that a kernel accumulating into this leaf's gradient is about to be enqueued.
        Every gradient kernel ACCUMULATES (TMA reduce-add / red.add) into the arena's fp32 gradient
        buffer, which begin_grad() zero-fills once per step; the record feeds the data-parallel
        bucket schedule (dp.py)."""
        a = self.arena
        a._written.add(self.path)
        if a.grad_hook is not None:
            a.grad_hook.wrote(self.path)


class ParamTree(dict):
    """Nested dict of parameter tensors carrying the arena they are views of."""
    arena = None


def _contig_stride(shape):
    st, acc = [], 1
    for s in reversed(shape):
        st.append(acc)
        acc *= s
    return tuple(reversed(st))


class ArenaBuilder:
    def __init__(self):
        self.size = 0
        self.leaves = []   # (path, offset, shape, stride)
        self.inits = []    # (offset, flat fp32 cpu/gpu tensor)

    def alloc(self, numel):
        base = self.size
        self.size += (numel + ALIGN - 1) // ALIGN * ALIGN
        return base

    def add_leaf(self, path, shape, init, base=None, stride=None, rel_offset=0):
        shape = tuple(shape)
        if base is None:
            numel = 1
            for s in shape:
                numel *= s
            base = self.alloc(numel)
        if stride is None:
            stride = _contig_stride(shape)
        self.leaves.append((tuple(path), base + rel_offset, shape, tuple(stride)))
        self.inits.append((tuple(path), init))

    def build(self, device):
        arena = ParamArena(max(self.size, ALIGN), device)
        for (path, off, shape, stride), (_, init) in zip(self.leaves, self.inits):
            p = Param(path, arena, off, shape, stride)
            arena.params[path] = p
            if init is not None and not arena.master.is_meta:
                p.master.copy_(init.to(device=device, dtype=torch.float32))
        arena.refresh_shadow()
        return arena


class ParamArena:
    def __init__(self, size, device):
        self.size = size
        self.device = torch.device(device)
        self.master = torch.zeros(size, dtype=torch.float32, device=device)
        self.shadow = torch.zeros(size, dtype=torch.bfloat16, device=device)
        self.grad = torch.zeros(size, dtype=torch.float32, device=device)
        self.m = None
        self.v = None
        self.params = {}
        self._written = set()
        self.grad_hook = None

    @property
    def abstract(self):
        return self.master.is_meta

    def refresh_shadow(self):
        if self.abstract:
            return
        if self.master.is_cuda:
            from . import ops
            ops.cast_bf16(self.master, self.shadow)
        else:
            raise RuntimeError("parameter arena must live on a CUDA device (no CPU fallback)")

    def begin_grad(self):
        """Zero the whole gradient arena with one fill kernel (all gradient kernels accumulate)."""
        self._written = set()
        if not self.abstract:
            from . import ops
            ops.fill(self.grad, 0.0)

    def _tree(self, attr):
        root = ParamTree()
        root.arena = self
        for path, p in self.params.items():
            d = root
            for k in path[:-1]:
                d = d.setdefault(k, {})
            d[path[-1]] = getattr(p, attr)
        return root

    def tree(self):
        return self._tree("master")

    def _tree_of(self, flat):
        """Tree of views into another flat buffer with the arena's layout (Adam moments)."""
        root = ParamTree()
        root.arena = self
        for path, p in self.params.items():
            d = root
            for k in path[:-1]:
                d = d.setdefault(k, {})
            d[path[-1]] = p._view(flat)
        return root

    def load_tree_into(self, flat, tree):
        """Copy a nested dict with the parameter tree's names / shapes into `flat` (arena layout)."""
        def walk(d, prefix):
            for k, v in d.items():
                if isinstance(v, dict):
                    walk(v, prefix + (k,))
                else:
                    p = self.params[prefix + (k,)]
                    p._view(flat).copy_(v.to(device=self.device, dtype=torch.float32))
        walk(tree, ())

    def grad_tree(self):
        return self._tree("grad")

    def load_tree(self, tree):
        """Copy values from a foreign nested dict (e.g. the oracle's parameters) into the arena."""
        seen = set()

        def walk(d, prefix):
            for k, v in d.items():
                if isinstance(v, dict):
                    walk(v, prefix + (k,))
                else:
                    path = prefix + (k,)
                    if path not in self.params:
                        raise KeyError(f"unexpected parameter {'/'.join(path)}")
                    p = self.params[path]
                    if tuple(v.shape) != p.shape:
                        raise ValueError(f"{'/'.join(path)}: shape {tuple(v.shape)} != {p.shape}")
                    p.master.copy_(v.to(device=self.device, dtype=torch.float32))
                    seen.add(path)
        walk(tree, ())
        missing = set(self.params) - seen
        if missing:
            raise KeyError(f"missing parameters: {sorted('/'.join(m) for m in missing)[:5]}")
        self.refresh_shadow()
        return self.tree()

    def ensure_moments(self):
        if self.m is None:
            self.m = torch.zeros_like(self.master)
            self.v = torch.zeros_like(self.master)
