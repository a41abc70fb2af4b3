"""Data-parallel host logic on CPU: world_size-2 gloo run of the sharding + flat-bucket gradient all-reduce
(singa_b200/parallel.py).  The summed gradients of two shards must equal the single-process gradients."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from singa_b200.parallel import GradientBucket, shard_by_edges


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


def test_two_rank_allreduce_matches_single_process(tmp_path):
    out = str(tmp_path / "grads.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    sizes, xs = _data()
    m = _model()
    sum(m(x).pow(2).sum() for x in xs).backward()
    for n, p in m.named_parameters():
        if p.grad is None:
            assert got[n] is None, n
        else:
            assert torch.allclose(got[n], p.grad, rtol=1e-5, atol=1e-6), n
