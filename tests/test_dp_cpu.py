"""Data-parallel host logic on CPU: bucket planning and the world_size-2 gradient exchange over
gloo (the N>1 path of SURVEY 8e; NCCL replaces gloo on the GPU box, nothing else changes)."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "self-attention-experiments-vision_b200")


def test_plan_buckets_reverse_order_and_cover():
    from sav.dp import plan_buckets
    leaves = [(("a",), 0, 100), (("b",), 128, 50), (("c",), 192, 300), (("d",), 512, 10)]
    buckets = plan_buckets(leaves, 200)
    # first bucket is at the END of the arena (gradients of the last layers arrive first)
    assert buckets[0][1] == 522 and buckets[0][2][0] == ("d",)
    lo = min(b[0] for b in buckets)
    hi = max(b[1] for b in buckets)
    assert (lo, hi) == (0, 522)
    # buckets tile the range without overlap
    spans = sorted((b[0], b[1]) for b in buckets)
    for (l0, h0), (l1, h1) in zip(spans, spans[1:]):
        assert h0 <= l1
    assert sorted(p for b in buckets for p in b[2]) == [("a",), ("b",), ("c",), ("d",)]


def test_plan_buckets_merges_interleaved_views():
    """queries/keys/values are strided views of one fused [C,3,H,d] block: never split them."""
    from sav.dp import plan_buckets
    leaves = [(("q",), 0, 1000), (("k",), 64, 1000), (("v",), 128, 1000), (("o",), 1152, 100)]
    buckets = plan_buckets(leaves, 1)
    qkv = [b for b in buckets if ("q",) in b[2]][0]
    assert set(qkv[2]) == {("q",), ("k",), ("v",)} and (qkv[0], qkv[1]) == (0, 1128)


class _FakeParam:
    def __init__(self, path, arena, offset, n):
        self.path, self.arena, self.offset, self.shape, self.stride = path, arena, offset, (n,), (1,)

    @property
    def grad(self):
        return self.arena.grad[self.offset:self.offset + self.shape[0]]

    def touch_grad(self):
        self.arena._written.add(self.path)
        if self.arena.grad_hook is not None:
            self.arena.grad_hook.wrote(self.path)


class _FakeArena:
    """The slice of ParamArena that DataParallel touches, on CPU tensors."""

    def __init__(self, sizes):
        self.params, off = {}, 0
        for i, n in enumerate(sizes):
            self.params[(f"p{i}",)] = _FakeParam((f"p{i}",), self, off, n)
            off += (n + 63) // 64 * 64
        self.grad = torch.zeros(off)
        self.master = torch.zeros(off)
        self._written = set()
        self.grad_hook = None

    def refresh_shadow(self):
        pass


def _worker(rank, world, port, q):
    sys.path.insert(0, PKG)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sav.dp import DataParallel
        arena = _FakeArena([300, 70, 500, 64])
        dp = DataParallel(bucket_mb=400 * 4 / (1 << 20))
        # replicate: rank 0's parameters win
        arena.master.fill_(float(rank + 1))
        dp.broadcast_params(arena)
        assert torch.all(arena.master == 1.0)
        results = []
        for step in range(3):  # step 0 learns the schedule, steps 1-2 fire buckets mid-backward
            arena.grad.zero_()
            arena._written = set()
            arena.grad_hook = dp
            dp.begin(arena)
            order = list(reversed(list(arena.params.values())))  # backward visits leaves in reverse
            fired_before_end = 0
            for node, p in enumerate(order):
                p.touch_grad()
                p.grad.fill_(float((rank + 1) * (node + 1)))
                dp.after_node(node)
                if node < len(order) - 1:
                    fired_before_end = len(dp._fired)
            dp.finish()
            arena.grad_hook = None
            expect = [sum((r + 1) * (node + 1) for r in range(world)) for node in range(len(order))]
            ok = all(torch.all(p.grad == e) for p, e in zip(order, expect))
            results.append((ok, fired_before_end, len(dp.buckets)))
        q.put((rank, results))
    finally:
        dist.destroy_process_group()


def test_two_rank_bucketed_allreduce_overlaps_backward():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, results in out.items():
        assert all(ok for ok, _, _ in results), (rank, results)
        n_buckets = results[0][2]
        assert n_buckets >= 2
        assert results[0][1] == 0           # learning pass: everything reduced at the end
        assert results[1][1] >= 1           # afterwards buckets fire before backward has finished
