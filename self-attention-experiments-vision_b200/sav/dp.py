"""Batch-sharded data parallelism: the replacement of train.py's pmap + lax.pmean (train.py:96,
228-231).  One process per GPU; parameters and optimizer state replicated; the only
data-path collective is the gradient all-reduce.

The flat fp32 gradient arena is cut into contiguous buckets.  Backward writes gradients in reverse
forward order (head first, patch embedding last), so buckets are numbered from the END of the arena.
A first "learning" backward pass records which tape node is the last writer of every leaf; from
then on, as soon as the last writer of a bucket has been enqueued, that bucket's NCCL all-reduce is
issued asynchronously (torch's NCCL process group runs it on its own stream behind an event on the
compute stream), so the exchange over NVLink/NVSwitch overlaps the remaining backward kernels.  The
optimizer waits on all buckets.  Everything is CUDA-graph capturable (fork/join through events).

Transport.  On NVLink / NVSwitch boxes the exchange is not a library collective: the gradient arena
is re-homed into symmetric memory (torch.distributed._symmetric_memory does the allocation and the
handle exchange -- plumbing), and ONE kernel of ours (csrc/dp.cu, `sav_dp_allreduce_sum_f32`) sums
the whole arena in place right after the last backward kernel: `multimem.ld_reduce` / `multimem.st`
through the switch (NVLS) when the allocation has a multicast address, peer loads / stores otherwise.
It runs when nothing else does, so it does not share SMs with the persistent one-CTA-per-SM backward
kernels the way NCCL's overlapped CTAs do.  `transport="nccl"` (or a box where
the symmetric-memory rendezvous fails) keeps the bucketed, overlapped NCCL path.

Works on CPU tensors with the gloo backend too (tests/test_dp_cpu.py): only the transport differs.
"""
import os
import warnings

import torch
import torch.distributed as dist


def plan_buckets(leaves, bucket_elems):
    """leaves: iterable of (path, offset, span) describing arena ranges [offset, offset+span).
    Returns buckets [(lo, hi, [paths])] covering the arena from its END to its start, each closed
    as soon as it holds >= bucket_elems elements.  Ranges that overlap (interleaved views of one
    fused block, e.g. queries/keys/values) are merged into one range first."""
    spans = sorted(((off, off + span, path) for path, off, span in leaves), key=lambda t: t[0])
    merged = []  # [lo, hi, [paths]]
    for lo, hi, path in spans:
        if merged and lo < merged[-1][1]:
            merged[-1][1] = max(merged[-1][1], hi)
            merged[-1][2].append(path)
        else:
            merged.append([lo, hi, [path]])
    buckets = []
    cur_hi, cur_lo, cur_paths = None, None, []
    for lo, hi, paths in reversed(merged):
        if cur_hi is None:
            cur_hi = hi
        cur_lo = lo
        cur_paths.extend(paths)
        if cur_hi - cur_lo >= bucket_elems:
            buckets.append((cur_lo, cur_hi, cur_paths))
            cur_hi, cur_paths = None, []
    if cur_paths:
        buckets.append((cur_lo, cur_hi, cur_paths))
    return buckets


def _leaf_span(p):
    """extent (in elements) of a possibly strided arena view"""
    span = 1
    for s, st in zip(p.shape, p.stride):
        span += (s - 1) * st
    return span


class DataParallel:
    """Gradient hook for sav.tape.value_and_grad: begin() / wrote(path) / after_node(i) / finish()."""

    def __init__(self, bucket_mb=16.0, group=None, overlap=True, reference_double_scale=False,
                 transport=None):
        if not dist.is_initialized():
            raise RuntimeError("DataParallel needs an initialised torch.distributed process group")
        self.group = group
        # "peer" = our kernel over symmetric memory (default on CUDA), "nccl" = bucketed NCCL
        self.transport = transport or os.environ.get("SAV_DP_TRANSPORT", "peer")
        self._symm = None         # symmetric-memory handle of the re-homed gradient arena
        self.world_size = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.bucket_elems = int(bucket_mb * (1 << 20) / 4)
        self.overlap = overlap
        # train.py:91 divides the loss by device_count and train.py:96 pmean's the gradients on top
        # (net 1/world^2).  Default: the true global mean (1/world).  The switch
        # reproduces the reference's scaling -- it matters under clip_by_global_norm, whose
        # threshold then sees gradients `world` times smaller.
        self.reference_double_scale = bool(reference_double_scale)
        self.buckets = None       # [(lo, hi, paths)]
        self.fire_after = None    # node index -> [bucket ids]; None until learned
        self._last_writer = {}
        self._node = -1
        self._works = []
        self._arena = None
        self.stats = {"buckets": 0, "bytes": 0}

    # -- setup ------------------------------------------------------------------------------
    def broadcast_params(self, arena, src=0):
        """flax.jax_utils.replicate (train.py:228): every rank starts from rank 0's parameters."""
        dist.broadcast(arena.master, src=src, group=self.group)
        arena.refresh_shadow()
        self.attach(arena)

    def attach(self, arena):
        """Re-home the arena's gradient buffer into symmetric memory (collective; call it on every
        rank before the first step / graph capture).  Falls back to the NCCL transport, with a
        warning, when symmetric memory cannot be set up on this box."""
        if self.transport != "peer" or self._symm is not None or not arena.grad.is_cuda:
            if not arena.grad.is_cuda:
                self.transport = "nccl"     # gloo / CPU tests
            return
        ok = torch.ones(1, device=arena.grad.device)
        try:
            import torch.distributed._symmetric_memory as symm_mem
            buf = symm_mem.empty(arena.size, dtype=torch.float32, device=arena.grad.device)
            hdl = symm_mem.rendezvous(buf, self.group if self.group is not None else dist.group.WORLD)
            buf.zero_()
            self._symm = (buf, hdl)
        except Exception as e:   # noqa: BLE001 -- any failure means "not on this box"
            warnings.warn(f"sav.dp: symmetric-memory transport unavailable ({type(e).__name__}: {e}); "
                          f"using NCCL")
            ok.zero_()
        # all ranks take the same branch
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
        if float(ok.item()) < 1.0:
            self._symm = None
            self.transport = "nccl"
            return
        buf, hdl = self._symm
        arena.grad = buf                      # Param.grad views are taken from arena.grad on access
        self._arena_attached = arena
        mc = int(hdl.multicast_ptr) and os.environ.get("SAV_DP_MULTICAST", "1") == "1"
        self.stats["transport"] = "peer/multimem" if mc else "peer/p2p"

    def _plan(self, arena):
        leaves = [(path, p.offset, _leaf_span(p)) for path, p in arena.params.items()]
        self.buckets = plan_buckets(leaves, self.bucket_elems)
        self.stats.update({"buckets": len(self.buckets),
                           "bytes": sum((hi - lo) * 4 for lo, hi, _ in self.buckets)})

    # -- hook protocol (called by tape.value_and_grad) -------------------------------------------
    def begin(self, arena):
        if self.transport == "peer" and self._symm is None and not arena.grad.is_cuda:
            self.transport = "nccl"         # gloo / CPU tensors
        if self.transport == "peer" and self._symm is None:
            raise RuntimeError("DataParallel: call attach(arena) (or broadcast_params) on every rank "
                               "before the first step")
        if self.buckets is None or arena is not self._arena:
            self._arena = arena
            self._plan(arena)
            self.fire_after = None
            self._last_writer = {}
        self._node = -1
        self._works = []
        self._fired = set()

    def wrote(self, path):
        if self.fire_after is None:
            self._last_writer[path] = self._node + 1  # node currently executing

    def after_node(self, i):
        self._node = i
        if self.fire_after is not None and self.overlap and self.transport == "nccl":
            for b in self.fire_after.get(i, ()):
                self._launch(b)

    def finish(self):
        """Called after the last backward node: reduce what is left and make the compute stream
        wait for every bucket (never-written leaves are already zero: arena.begin_grad())."""
        if self.transport == "peer":
            # one kernel over the whole arena, on the compute stream, after the last gradient kernel
            from . import ops
            buf, hdl = self._symm
            if self._arena.grad.data_ptr() != buf.data_ptr():
                raise RuntimeError("DataParallel: the arena's gradient buffer is not the symmetric one")
            n = (self._arena.size + 3) // 4 * 4
            mc = hdl.multicast_ptr if os.environ.get("SAV_DP_MULTICAST", "1") == "1" else 0
            ops.dp_allreduce_sum(mc, hdl.buffer_ptrs_dev, hdl.signal_pad_ptrs_dev,
                                 hdl.signal_pad_size, 0, min(n, buf.numel()), self.rank, self.world_size,
                                 int(os.environ.get("SAV_DP_BLOCKS", "0")))
            self.stats["bytes"] = self._arena.size * 4
            return
        if self.fire_after is None:
            # learning pass: derive the firing schedule for the following steps
            sched = {}
            for b, (_, _, paths) in enumerate(self.buckets):
                last = max(self._last_writer.get(p, self._node) for p in paths)
                sched.setdefault(last, []).append(b)
            self.fire_after = sched
        for b in range(len(self.buckets)):
            if b not in self._fired:
                self._launch(b)
        for w in self._works:
            w.wait()
        self._works = []

    def _launch(self, b):
        lo, hi, _ = self.buckets[b]
        self._fired.add(b)
        # the loss gradient is already scaled by 1/(B*world) (functional.softmax_cross_entropy),
        # so SUM over ranks is the global-batch mean: lax.pmean(grads) of train.py:96
        self._works.append(dist.all_reduce(self._arena.grad[lo:hi], op=dist.ReduceOp.SUM,
                                           group=self.group, async_op=True))

    # -- non-hook convenience (eval path, train.py:120 psum) ---------------------------------------
    def all_reduce_sum(self, t):
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t
