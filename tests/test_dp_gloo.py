"""Data-parallel host logic on CPU: world_size-2 gloo run of the sharding + flat-bucket gradient all-reduce
(singa_b200/parallel.py).  The summed gradients of two shards must equal the single-process gradients."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from singa_b200.parallel import GradientBucket, GradientBuckets, shard_by_edges


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _model():
    torch.manual_seed(0)
    m = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.SiLU(), torch.nn.Linear(5, 3))
    m.unused = torch.nn.Parameter(torch.zeros(4))  # never receives a gradient (like sphere_embedding)
    return m


def _data():
    g = torch.Generator().manual_seed(1)
    sizes = [5, 9, 2, 7, 3, 8]  # "edge counts" of six complexes
    return sizes, [torch.randn(n, 6, generator=g) for n in sizes]


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sizes, xs = _data()
    mine = shard_by_edges(sizes, world)[rank]
    m = _model()
    loss = sum(m(xs[i]).pow(2).sum() for i in mine)
    loss.backward()
    GradientBucket(m.parameters()).allreduce_(average=False)
    if rank == 0:
        torch.save({n: p.grad for n, p in m.named_parameters()}, out)
    dist.destroy_process_group()


def test_shards_are_balanced_and_complete():
    sizes = [5, 9, 2, 7, 3, 8, 1]
    shards = shard_by_edges(sizes, 3)
    assert sorted(i for s in shards for i in s) == list(range(7))
    loads = [sum(sizes[i] for i in s) for s in shards]
    assert max(loads) - min(loads) <= max(sizes)
    assert shard_by_edges(sizes, 3) == shards  # deterministic
    # equal counts per rank (what the Trainer uses): 8 complexes over 4 ranks = 2 each, whatever the edge counts
    sizes = [100, 1, 1, 1, 1, 1, 1, 50]
    assert sorted(len(s) for s in shard_by_edges(sizes, 4)) != [2, 2, 2, 2]
    eq = shard_by_edges(sizes, 4, equal_counts=True)
    assert [len(s) for s in eq] == [2, 2, 2, 2] and sorted(i for s in eq for i in s) == list(range(8))
    assert [len(s) for s in shard_by_edges([3, 2, 1], 2, equal_counts=True)] == [1, 2]   # cap = ceil(3 / 2)


def test_two_rank_allreduce_matches_single_process(tmp_path):
    out = str(tmp_path / "grads.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    sizes, xs = _data()
    m = _model()
    sum(m(x).pow(2).sum() for x in xs).backward()
    for n, p in m.named_parameters():
        if p.grad is None:   # never used anywhere: stays in the reduction as zeros (parallel.GradientBuckets, "idle")
            assert not got[n].any(), n
        else:
            assert torch.allclose(got[n], p.grad, rtol=1e-5, atol=1e-6), n


class _TwoBlock(torch.nn.Module):
    """Parameter names shaped like the real model: two 'blocks' plus a tail and a never-used table."""

    def __init__(self):
        super().__init__()
        torch.manual_seed(3)
        self.blocks = torch.nn.ModuleList([torch.nn.Linear(6, 6), torch.nn.Linear(6, 6)])
        self.head = torch.nn.Linear(6, 2)
        self.side = torch.nn.Linear(6, 2)    # only rank 1's data goes through it
        self.unused = torch.nn.Parameter(torch.zeros(4))

    def forward(self, x, use_side):
        h = self.blocks[1](torch.tanh(self.blocks[0](x)))
        out = self.head(h).pow(2).sum()
        if use_side:
            out = out + self.side(h).sum()
        return out


def _worker_buckets(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sizes, xs = _data()
    mine = shard_by_edges(sizes, world)[rank]
    m = _TwoBlock()
    gb = GradientBuckets(m.named_parameters(), average=False)
    steps = []
    for step in range(3):          # step 0 settles the participating set, steps 1-2 reduce from the hooks
        gb.zero_grad()
        sum(m(xs[i], use_side=(rank == 1)) for i in mine).backward()
        gb.finish()
        steps.append({n: (None if p.grad is None else p.grad.clone()) for n, p in m.named_parameters()})
    if True:
        torch.save(dict(steps=steps, dropped=gb.dropped, keys=[b.key for b in gb.buckets]), out + f".{rank}")
    dist.destroy_process_group()


def test_bucketed_overlapped_allreduce_with_a_rank_missing_a_gradient(tmp_path):
    """GradientBuckets: gradients live in per-block flat buffers, buckets reduce in a fixed order as backward
    completes them; a parameter without a gradient on ONE rank (``side`` on rank 0) still ends with the summed
    gradient on every rank, and a parameter without a gradient on ANY rank moves to the ``idle`` bucket,
    reduced last, with a zero gradient on every rank (it is no longer dropped from the reduction)."""
    out = str(tmp_path / "gb.pt")
    mp.spawn(_worker_buckets, args=(2, _free_port(), out), nprocs=2, join=True)
    r0, r1 = torch.load(out + ".0"), torch.load(out + ".1")
    assert r0["dropped"] == ["unused"] and r1["dropped"] == ["unused"]
    assert r0["keys"] == ["block1", "block0", "rest", "idle"]
    sizes, xs = _data()
    shards = shard_by_edges(sizes, 2)
    m = _TwoBlock()
    (sum(m(xs[i], False) for i in shards[0]) + sum(m(xs[i], True) for i in shards[1])).backward()
    for step in range(3):
        for n, p in m.named_parameters():
            a, b = r0["steps"][step][n], r1["steps"][step][n]
            if n == "unused":
                assert not a.any() and not b.any()
                continue
            assert torch.equal(a, b), (step, n)          # every rank holds the same reduced gradient
            assert torch.allclose(a, p.grad, rtol=1e-5, atol=1e-6), (step, n)


def _worker_late_parameter(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sizes, xs = _data()
    mine = shard_by_edges(sizes, world)[rank]
    m = _TwoBlock()
    gb = GradientBuckets(m.named_parameters(), average=False)
    steps = []
    for step in range(4):          # ``side`` is used by nobody in steps 0-1 and wakes up on rank 1 only in step 2
        gb.zero_grad()
        sum(m(xs[i], use_side=(rank == 1 and step >= 2)) for i in mine).backward()
        gb.finish()
        steps.append({n: p.grad.clone() for n, p in m.named_parameters()})
    # gradient accumulation: two backward passes in one step need overlap=False; with overlap on the second one raises
    gb.zero_grad()
    sum(m(xs[i], False) for i in mine).backward()
    try:
        sum(m(xs[i], False) for i in mine).backward()
        raised = False
    except RuntimeError as e:
        raised = "overlap=False" in str(e)
    gb.finish()
    sum(m(xs[i], False) for i in mine).backward()      # after finish(): outside the cycle, neither reduced nor an error
    for h in gb._hooks:
        h.remove()
    for p in m.parameters():
        p.grad = None
    gb2 = GradientBuckets(m.named_parameters(), average=False, overlap=False)
    acc = []
    for step in range(2):
        gb2.zero_grad()
        for _ in range(2):
            sum(m(xs[i], rank == 1) for i in mine).backward()
        gb2.finish()
        acc.append({n: p.grad.clone() for n, p in m.named_parameters()})
    torch.save(dict(steps=steps, idle=gb.idle, raised=raised, acc=acc), out + f".{rank}")
    dist.destroy_process_group()


def test_parameter_that_wakes_up_after_the_first_step_is_still_reduced(tmp_path):
    """a parameter with no gradient anywhere in step 0 used to be dropped for good; when it later got a
    gradient on one rank it was never all-reduced and the ranks diverged.  It now lives in the ``idle`` bucket, which
    is reduced from finish() every step.  Also: a second backward() in a step raises under overlap, and
    overlap=False sums both passes."""
    out = str(tmp_path / "late.pt")
    mp.spawn(_worker_late_parameter, args=(2, _free_port(), out), nprocs=2, join=True)
    r0, r1 = torch.load(out + ".0"), torch.load(out + ".1")
    assert sorted(r0["idle"]) == ["side.bias", "side.weight", "unused"] == sorted(r1["idle"])
    assert r0["raised"] and r1["raised"]
    sizes, xs = _data()
    shards = shard_by_edges(sizes, 2)
    for step in range(4):
        m = _TwoBlock()
        (sum(m(xs[i], False) for i in shards[0]) + sum(m(xs[i], step >= 2) for i in shards[1])).backward()
        for n, p in m.named_parameters():
            a, b = r0["steps"][step][n], r1["steps"][step][n]
            assert torch.equal(a, b), (step, n)
            want = torch.zeros_like(p) if p.grad is None else p.grad
            assert torch.allclose(a, want, rtol=1e-5, atol=1e-6), (step, n)
    assert bool(r0["steps"][2]["side.weight"].any())
    m = _TwoBlock()
    (2 * sum(m(xs[i], False) for i in shards[0]) + 2 * sum(m(xs[i], True) for i in shards[1])).backward()
    for step in range(2):
        for n, p in m.named_parameters():
            a, b = r0["acc"][step][n], r1["acc"][step][n]
            assert torch.equal(a, b), (step, n)
            want = torch.zeros_like(p) if p.grad is None else p.grad
            assert torch.allclose(a, want, rtol=1e-5, atol=1e-6), (step, n)


def _settle_worker(rank, world, port, out):
    """bench.settle_steps with a step that contains a collective and allocation counters that settle after a
    DIFFERENT number of steps on the two ranks (rank 0: 5 steps, rank 1: 2)."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world, timeout=__import__("datetime").timedelta(seconds=60))
    grows_until = 5 if rank == 0 else 2
    state = {"steps": 0, "allocs": 0}

    def fn():
        state["steps"] += 1
        if state["steps"] <= grows_until:
            state["allocs"] += 1
        t = torch.ones(3)
        dist.all_reduce(t)            # the gradient all-reduce of a step
        assert float(t[0]) == world

    def any_rank(flag):
        t = torch.tensor([1.0 if flag else 0.0])
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return bool(t.item() > 0)

    done = bench.settle_steps(fn, 3, 12, lambda: state["allocs"], any_rank)
    dist.barrier()                    # what the timed region does next: must not meet a stray all-reduce
    torch.save(done, f"{out}.{rank}")
    dist.destroy_process_group()


def test_warmup_until_steady_runs_the_same_number_of_steps_on_every_rank(tmp_path):
    """Regression: the bench's settle loop stopped per rank; at N = 8 two ranks were still in a warm-up step's
    all-reduce while six sat in the timed region's barrier (NCCL watchdog timeout after 600 s)."""
    out = str(tmp_path / "done")
    mp.spawn(_settle_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    d0, d1 = torch.load(out + ".0"), torch.load(out + ".1")
    assert d0 == d1 == 6              # rank 0's last growing step is 5: step 6 is the first steady one everywhere
    import bench
    state = {"n": 0}
    assert bench.settle_steps(lambda: None, 3, 12, lambda: state["n"]) == 3      # single rank, steady from the start
    def grow():
        state["n"] += 1
    assert bench.settle_steps(grow, 3, 4, lambda: state["n"]) == 7               # never steady: bounded by lo + hi
