"""Per-rank phase times of the end-to-end step under torchrun (weak batches): where does e2e lose against the
resident-input step at N > 1?  Each phase is followed by a device synchronise (diagnosis, not a benchmark)."""
import datetime
import os, sys, time
os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import torch.distributed as dist
import bench
from singa_b200 import ops, synthetic as syn
from singa_b200.embedding import DEFAULT_CONFIG, EquivariantEmbedding
from singa_b200.parallel import GradientBuckets

world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=300))
ops.set_precision("bf16")
batch, _ = bench.make_batch(rank, 32, None)
host = {k: v.pin_memory() for k, v in batch.items()}
resident = {k: v.to(dev) for k, v in host.items()}
torch.manual_seed(0)
model = EquivariantEmbedding(DEFAULT_CONFIG, device=dev)
syn.load_portable_parameters(model, 0)
opt = torch.optim.Adam(model.parameters(), lr=1e-4)
gb = GradientBuckets(model.named_parameters())
rs = np.random.RandomState(7)
Rp = torch.from_numpy(rs.normal(size=(resident["z_p"].numel(), 49, 16)).astype(np.float32)).to(dev)
Rl = torch.from_numpy(rs.normal(size=(resident["z_l"].numel(), 49, 16)).astype(np.float32)).to(dev)
sync = torch.cuda.synchronize


def step(g, marks=None, do_sync=False):
    def mark(name):
        if marks is not None:
            if do_sync:
                sync()
            marks.append((name, time.perf_counter()))
    gb.zero_grad(); mark("zero")
    out = model(syn.GraphView(g)); mark("fwd")
    loss = (out[syn.PROTEIN].embedding * Rp).sum() + (out[syn.LIGAND].embedding * Rl).sum()
    loss.backward(); mark("bwd")
    gb.finish(); mark("allreduce")
    opt.step(); mark("opt")
    return loss


def barrier():
    if world > 1:
        dist.barrier()
    sync()


for _ in range(4):
    step(resident)
for name, fresh, item in (("resident/no item", False, False), ("resident/item", False, True), ("fresh/item", True, True),
                          ("fresh/no item", True, False)):
    barrier()
    t0 = time.perf_counter()
    for _ in range(6):
        g = {k: v.to(dev, non_blocking=True) for k, v in host.items()} if fresh else resident
        l = step(g)
        if item:
            float(l.item())
    barrier()
    wall = (time.perf_counter() - t0) / 6 * 1e3
    # host-side (unsynchronised) launch time per phase of one more step
    m = [("start", time.perf_counter())]
    g = {k: v.to(dev, non_blocking=True) for k, v in host.items()} if fresh else resident
    m.append(("h2d", time.perf_counter()))
    step(g, m, do_sync=False)
    sync()
    m.append(("drain", time.perf_counter()))
    ph = {b[0]: round((b[1] - a[1]) * 1e3, 1) for a, b in zip(m[:-1], m[1:])}
    print(f"rank {rank} {name:18s} wall/step {wall:6.1f} ms   host phases {ph}   edges {bench.count_edges(batch)}", flush=True)
if world > 1:
    dist.destroy_process_group()
