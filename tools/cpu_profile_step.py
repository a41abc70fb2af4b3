"""cProfile of one bench step: where the host-side launch time goes (the tiny ligand passes are launch-bound)."""
import cProfile, pstats, io, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from singa_b200 import ops, synthetic as syn
from singa_b200.embedding import DEFAULT_CONFIG, EquivariantEmbedding

dev = torch.device("cuda", 0)
ops.set_precision("bf16")
batch, _ = bench.make_batch(0, int(os.environ.get("BATCH", "32")), None)
g = {k: v.to(dev) for k, v in batch.items()}
torch.manual_seed(0)
model = EquivariantEmbedding(DEFAULT_CONFIG, device=dev)
syn.load_portable_parameters(model, 0)
opt = torch.optim.Adam(model.parameters(), lr=1e-4)
view = syn.GraphView(g)


def step(timing=None):
    t0 = time.perf_counter()
    opt.zero_grad(set_to_none=True)
    out = model(view)
    loss = out[syn.PROTEIN].embedding.sum() + out[syn.LIGAND].embedding.sum()
    t1 = time.perf_counter()
    loss.backward()
    t2 = time.perf_counter()
    opt.step()
    t3 = time.perf_counter()
    if timing is not None:
        timing.append((t1 - t0, t2 - t1, t3 - t2))


for _ in range(3):
    step()
torch.cuda.synchronize()
tm = []
for _ in range(3):
    step(tm)
    torch.cuda.synchronize()
for f, b, o in tm:
    print(f"host time: forward {f*1e3:.1f} ms, backward {b*1e3:.1f} ms, optimizer {o*1e3:.1f} ms")
pr = cProfile.Profile()
pr.enable()
step()
pr.disable()
torch.cuda.synchronize()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(45)
print(s.getvalue()[:9000])
