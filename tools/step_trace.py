"""Per-step trace from process start: step time (CUDA events), cudaMalloc calls, save/recompute decisions, Python GC
runs.  Looks for transients (steps 6-7 of a run were 330-410 ms in bench traces)."""
import gc, os, sys, time
os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from singa_b200 import ops, synthetic as syn
from singa_b200.embedding import DEFAULT_CONFIG, EquivariantEmbedding
from singa_b200.parallel import GradientBuckets

dev = torch.device("cuda", 0)
ops.set_precision("bf16")
batch, _ = bench.make_batch(0, 32, None)
resident = {k: v.to(dev) for k, v in batch.items()}
torch.manual_seed(0)
model = EquivariantEmbedding(DEFAULT_CONFIG, device=dev)
syn.load_portable_parameters(model, 0)
opt = torch.optim.Adam(model.parameters(), lr=1e-4)
gb = GradientBuckets(model.named_parameters())
Rp = torch.randn(resident["z_p"].numel(), 49, 16, device=dev)
Rl = torch.randn(resident["z_l"].numel(), 49, 16, device=dev)
view = syn.GraphView(resident)
if os.environ.get("RESERVE", "1") == "1":
    print("reserved", ops.reserve_workspace(0.6, dev) / 2 ** 30, "GiB")
gcs = []
gc.callbacks.append(lambda phase, info: gcs.append(info["generation"]) if phase == "start" else None)
sync_every = int(os.environ.get("SYNC_EVERY", "0"))
evs = [torch.cuda.Event(enable_timing=True) for _ in range(31)]
rows = []
evs[0].record()
for i in range(30):
    m0 = torch.cuda.memory_stats(dev).get("num_device_alloc", 0)
    s0 = dict(ops.STATS); g0 = len(gcs); t0 = time.perf_counter()
    gb.zero_grad()
    out = model(view)
    loss = (out[syn.PROTEIN].embedding * Rp).sum() + (out[syn.LIGAND].embedding * Rl).sum()
    loss.backward(); gb.finish(); opt.step()
    evs[i + 1].record()
    if sync_every and (i + 1) % sync_every == 0:
        torch.cuda.synchronize()
    rows.append((time.perf_counter() - t0, torch.cuda.memory_stats(dev).get("num_device_alloc", 0) - m0,
                 ops.STATS["save"] - s0["save"], ops.STATS["recompute"] - s0["recompute"], gcs[g0:],
                 torch.cuda.memory_reserved(dev) / 2 ** 30))
torch.cuda.synchronize()
for i, r in enumerate(rows):
    print(f"step {i:2d} gpu {evs[i].elapsed_time(evs[i+1]):7.1f} ms host {r[0]*1e3:6.1f} ms cudaMalloc {r[1]:3d} save {r[2]} recompute {r[3]} "
          f"gc {r[4]} reserved {r[5]:.1f} GiB")
