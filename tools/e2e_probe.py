"""Where does the end-to-end step (pinned host inputs every step) lose time against the resident-input step?
Wall-clock per phase with a synchronise after each phase (so phases do not overlap: a diagnosis, not a benchmark)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from singa_b200 import ops, synthetic as syn
from singa_b200.embedding import DEFAULT_CONFIG, EquivariantEmbedding
from singa_b200.parallel import GradientBuckets

dev = torch.device("cuda", 0)
ops.set_precision("bf16")
batch, _ = bench.make_batch(0, 32, None)
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


def step(g, phases=None):
    t = [time.perf_counter()]
    def mark():
        if phases is not None:
            sync()
        t.append(time.perf_counter())
    gb.zero_grad(); mark()
    out = model(syn.GraphView(g)); mark()
    loss = (out[syn.PROTEIN].embedding * Rp).sum() + (out[syn.LIGAND].embedding * Rl).sum()
    loss.backward(); mark()
    gb.finish(); opt.step(); mark()
    if phases is not None:
        phases.append([b - a for a, b in zip(t[:-1], t[1:])])
    return loss


for _ in range(4):
    step(resident)
sync()
for name, fresh in (("resident", False), ("fresh host copy", True), ("resident", False), ("fresh host copy", True)):
    ops.STATS["save"] = ops.STATS["recompute"] = 0
    ph, walls = [], []
    for _ in range(4):
        sync(); t0 = time.perf_counter()
        g = {k: v.to(dev, non_blocking=True) for k, v in host.items()} if fresh else resident
        l = step(g)
        float(l.item()); walls.append(time.perf_counter() - t0)
    for _ in range(3):
        g = {k: v.to(dev, non_blocking=True) for k, v in host.items()} if fresh else resident
        sync(); step(g, ph)
    ph = np.mean(ph, axis=0) * 1e3
    free, total = torch.cuda.mem_get_info()
    print(f"{name:16s} wall/step {np.mean(walls)*1e3:7.1f} ms | synced phases: zero {ph[0]:.1f} fwd {ph[1]:.1f} bwd {ph[2]:.1f} "
          f"opt {ph[3]:.1f} | policy {dict(ops.STATS)} | free {free/2**30:.1f} GiB reserved "
          f"{torch.cuda.memory_reserved()/2**30:.1f} GiB allocated {torch.cuda.memory_allocated()/2**30:.1f} GiB")
