"""Data parallelism over complexes.

Complexes of a batch share no edges (pyg batching is a disjoint union), so the path shards by whole
complex: each rank runs forward+backward on its shard with no data-path collective, and the only
exchange is the parameter-gradient all-reduce (42.7 MB fp32), through torch.distributed (NCCL over
NVLink on GPUs, gloo in the CPU tests).  The reference has no distributed code at all (train.py:72
hard-codes cuda:0).

``GradientBuckets`` keeps the gradients *inside* a few flat fp32 buffers (``p.grad`` is a view), one
bucket per TransBlockV2 plus one for everything else, and launches each bucket's all-reduce from a
post-accumulate hook as soon as backward has produced the bucket's last gradient: the reduction of
blocks 2 and 1 overlaps the backward of blocks 1 and 0 of the protein pass, and there is no copy in
or out of the buckets (round 1 staged 230 per-parameter copies around one flat all-reduce).
"""
from __future__ import annotations

import re
from typing import Callable, Dict, Iterable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_by_edges(edge_counts: Sequence[int], world_size: int, equal_counts: bool = False) -> List[List[int]]:
    """Greedy longest-processing-time assignment of complexes to ranks, balancing EDGE counts
    (per-complex cost is proportional to edges, not atoms).  Deterministic; returns index lists.
    ``equal_counts``: no rank takes more than ceil(n / world_size) complexes (edges balanced within that), so a
    loader that draws a fixed number of complexes per rank and step walks every shard at the same pace and one
    step is one global batch."""
    order = sorted(range(len(edge_counts)), key=lambda i: (-edge_counts[i], i))
    loads = [0] * world_size
    shards: List[List[int]] = [[] for _ in range(world_size)]
    cap = -(-len(edge_counts) // world_size) if equal_counts else len(edge_counts)
    for i in order:
        r = min((k for k in range(world_size) if len(shards[k]) < cap), key=lambda k: (loads[k], k))
        shards[r].append(i)
        loads[r] += edge_counts[i]
    return [sorted(s) for s in shards]


def _default_bucket_of(name: str) -> str:
    m = re.search(r"(?:^|\.)blocks\.(\d+)\.", name)
    return f"block{m.group(1)}" if m else "rest"


class _Bucket:
    def __init__(self, key: str, named: List[Tuple[str, torch.nn.Parameter]]):
        self.key = key
        self.named = named
        self.flat: Optional[torch.Tensor] = None
        self.ready = 0
        self.handle = None
        self.launched = False


class GradientBuckets:
    """Gradients as views of per-bucket flat buffers + overlapped all-reduce.

    Use: ``gb = GradientBuckets(model.named_parameters())`` once; per step ``gb.zero_grad()`` (instead of the
    optimiser's), forward, backward (buckets reduce as they complete), ``gb.finish()`` (waits; gradients are then
    the sum - or mean - over ranks on EVERY rank, also for a parameter this rank's shard gave no gradient), step.

    Parameters that receive no gradient on any rank in the first step (``sphere_embedding*``: the int64 seed quirk,
    Embedding.py:262) are found once, globally, and moved to a bucket of their own, ``idle``, which is reduced LAST and
    only from ``finish()``, never from a hook.  They stay reduced every step (a few KB of zeros), so a parameter that
    wakes up later on some rank (a data-dependent branch, a hetero edge set that was empty in step 1, an unfrozen
    layer) is summed over the ranks like any other and the ranks cannot drift apart.  Their ``.grad`` is a view of
    zeros rather than ``None``: with Adam and ``weight_decay=0`` (config/train.yml:207) the update of such a parameter
    is exactly zero, as in the reference where the optimiser skips it; with weight decay it would decay.

    One ``backward()`` per step is assumed when ``overlap`` is on: a bucket is reduced the moment its last gradient
    arrives, and a gradient arriving for a bucket that was already reduced in the same step raises (a backward after
    ``finish()`` and before the next ``zero_grad()`` is outside the cycle and left alone).  For gradient accumulation (several
    backward passes per step) build with ``overlap=False``: everything is then reduced in ``finish()``.
    """

    def __init__(self, named_params: Iterable[Tuple[str, torch.nn.Parameter]], group=None, average: bool = True,
                 bucket_of: Callable[[str], str] = _default_bucket_of, overlap: bool = True):
        self.group = group
        self.average = average
        self.overlap = overlap
        self.idle: List[str] = []
        self._open = True     # inside a zero_grad() ... finish() cycle (hooks outside one are ignored)
        self._bucket_of = bucket_of
        self._named = [(n, p) for n, p in named_params if p.requires_grad]
        self._touched: Dict[int, bool] = {id(p): False for _, p in self._named}
        self._owner: Dict[int, _Bucket] = {}
        self._hooks = []
        self._settled = False
        self.buckets: List[_Bucket] = []
        self._build(self._named)
        for _, p in self._named:
            self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad))

    # -- construction --------------------------------------------------------------------------
    def _build(self, named):
        groups: Dict[str, list] = {}
        for n, p in named:
            groups.setdefault(self._bucket_of(n), []).append((n, p))
        self.buckets, self._owner = [], {}
        for key, items in groups.items():
            b = _Bucket(key, items)
            total = sum(p.numel() for _, p in items)
            dev = items[0][1].device
            b.flat = torch.zeros(total, dtype=torch.float32, device=dev)
            off = 0
            for _, p in items:
                view = b.flat[off:off + p.numel()].view_as(p)
                if p.grad is not None:   # built after a backward: keep what is there
                    view.copy_(p.grad)
                    self._touched[id(p)] = True
                p.grad = view
                off += p.numel()
                self._owner[id(p)] = b
            self.buckets.append(b)
        # NCCL needs the same collective order on every rank: buckets are reduced strictly in this sequence (the order
        # backward completes them in: last block first, everything else last), never in completion order
        def rank_of(b):
            m = re.match(r"block(\d+)$", b.key)
            return (0, -int(m.group(1))) if m else ((2, 0) if b.key == "idle" else (1, 0))
        self.buckets.sort(key=rank_of)
        self._next = 0

    def nbytes(self) -> int:
        return sum(b.flat.numel() * 4 for b in self.buckets)

    def _distributed(self) -> bool:
        return dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1

    # -- per step ------------------------------------------------------------------------------
    def zero_grad(self) -> None:
        self._open = True
        for b in self.buckets:
            b.flat.zero_()
            b.ready, b.handle, b.launched = 0, None, False
            self._next = 0
            off = 0
            for _, p in b.named:   # re-attach: an optimiser's zero_grad(set_to_none=True) may have dropped the views
                if p.grad is None or p.grad.data_ptr() != b.flat.data_ptr() + 4 * off:
                    p.grad = b.flat[off:off + p.numel()].view_as(p)
                off += p.numel()

    def _on_grad(self, p) -> None:
        self._touched[id(p)] = True
        b = self._owner.get(id(p))
        if b is None or not self._open:
            return   # a backward outside zero_grad() ... finish() (an evaluation pass, a parity probe) is not reduced
        if b.launched:
            raise RuntimeError(
                f"GradientBuckets: a gradient arrived for bucket '{b.key}' after its all-reduce was launched (a second "
                "backward() in one step?).  Build GradientBuckets(..., overlap=False) to reduce everything in finish().")
        b.ready += 1
        if self._settled and self.overlap:
            self._launch_ready()

    def _launch_ready(self, force: bool = False) -> None:
        while self._next < len(self.buckets):
            b = self.buckets[self._next]
            if not force and (b.ready != len(b.named) or b.key == "idle"):
                return
            self._launch(b)
            self._next += 1

    def _launch(self, b: _Bucket) -> None:
        b.launched = True
        if self._distributed():
            b.handle = dist.all_reduce(b.flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)

    def finish(self) -> None:
        """Launch what backward did not (buckets with a parameter that got no gradient on this rank), wait for all
        reductions, average.  After the first call the participating set is settled globally."""
        self._launch_ready(force=True)
        for b in self.buckets:
            if b.handle is not None:
                b.handle.wait()
                b.handle = None
            if self.average and self._distributed():
                b.flat.div_(dist.get_world_size(self.group))
        if not self._settled:
            self._settle()
        self._open = False

    def _settle(self) -> None:
        flags = torch.tensor([1.0 if self._touched[id(p)] else 0.0 for _, p in self._named],
                             device=self._named[0][1].device if self._named else "cpu")
        if self._distributed():
            dist.all_reduce(flags, op=dist.ReduceOp.MAX, group=self.group)
        live = flags.cpu().tolist()
        dead = {id(p) for (n, p), f in zip(self._named, live) if f <= 0}
        self._settled = True
        self.idle = [n for n, p in self._named if id(p) in dead]
        if dead:
            saved = {id(p): p.grad.clone() for _, p in self._named}
            inner = self._bucket_of
            self._bucket_of = lambda name, _idle=set(self.idle): "idle" if name in _idle else inner(name)
            self._build(self._named)
            for _, p in self._named:   # this step's (already reduced) gradients move to the rebuilt buckets
                p.grad.copy_(saved[id(p)])

    @property
    def dropped(self) -> List[str]:
        """Round-2 name of ``idle`` (the parameters are no longer dropped from the reduction)."""
        return self.idle

    # round-1 name, kept for callers that reduce after backward in one go
    def allreduce_(self) -> None:
        self.finish()


class GradientBucket(GradientBuckets):
    """Round-1 interface: one flat bucket, reduced by ``allreduce_`` after backward."""

    def __init__(self, params, group=None, average: bool = True):
        params = list(params)
        super().__init__(((f"p{i}", p) for i, p in enumerate(params)), group=group, average=average,
                         bucket_of=lambda name: "all")

    def allreduce_(self, group=None, average: Optional[bool] = None) -> None:
        if average is not None:
            self.average = average
        if group is not None:
            self.group = group
        # callers of this interface may have replaced or dropped the gradient views (zero_grad(set_to_none=True)):
        # bring whatever is in p.grad into the bucket, zeros where this rank has no gradient
        for b in self.buckets:
            off = 0
            for _, p in b.named:
                view = b.flat[off:off + p.numel()].view_as(p)
                if p.grad is None:
                    view.zero_()
                elif p.grad.data_ptr() != view.data_ptr():
                    view.copy_(p.grad)
                p.grad = view
                off += p.numel()
        self.finish()
