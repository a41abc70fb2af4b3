"""torch.profiler view of one bench step (which torch library ops still sit on the step, CPU vs GPU time)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from singa_b200 import ops, synthetic as syn
from singa_b200.embedding import DEFAULT_CONFIG, EquivariantEmbedding
from torch.profiler import profile, ProfilerActivity

dev = torch.device("cuda", 0)
ops.set_precision("bf16")
batch, _ = bench.make_batch(0, int(os.environ.get("BATCH", "32")), None)
g = {k: v.to(dev) for k, v in batch.items()}
torch.manual_seed(0)
model = EquivariantEmbedding(DEFAULT_CONFIG, device=dev)
syn.load_portable_parameters(model, 0)
opt = torch.optim.Adam(model.parameters(), lr=1e-4)

def step():
    opt.zero_grad(set_to_none=True)
    out = model(syn.GraphView(g))
    loss = out[syn.PROTEIN].embedding.sum() + out[syn.LIGAND].embedding.sum()
    loss.backward()
    opt.step()

for _ in range(3):
    step()
torch.cuda.synchronize()
import time
t0 = time.perf_counter(); step(); t_cpu = time.perf_counter() - t0; torch.cuda.synchronize(); t_all = time.perf_counter() - t0
print(f"cpu-side launch time {t_cpu*1e3:.1f} ms, step wall {t_all*1e3:.1f} ms")
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=60))
