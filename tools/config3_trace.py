"""Config-3 (full SINGA train step) host/device split at a per-rank batch: step time from CUDA events, host time of
the step call, sync count, and the top host-side ops from torch.profiler.  BATCH=8 is one rank's load at N=8."""
import os, sys, time
os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from singa_b200 import _lib, ops, dataset as ds, synthetic as syn
from singa_b200.gan import SINGA, default_config
from singa_b200.train import Trainer

B = int(os.environ.get("BATCH", "8"))
dev = torch.device("cuda", 0)
ops.set_precision("bf16")
data = ds.ComplexDataset(ds.synthetic_records(range(B), n_protein=None))
torch.manual_seed(0)
model = SINGA(default_config(), device=dev)
syn.load_portable_parameters(model.embedding, 0)
syn.load_portable_parameters_fanin(model.model, 1)
ops.reserve_workspace(0.6, dev)
tr = Trainer(model, data, batch_size_per_rank=B, shuffle=False, prefetch=2, device=dev)
for _ in range(4):
    tr.step()
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(11)]
host = []
ev[0].record()
for i in range(10):
    t0 = time.perf_counter()
    tr.step()
    host.append((time.perf_counter() - t0) * 1e3)
    ev[i + 1].record()
torch.cuda.synchronize()
gpu = [ev[i].elapsed_time(ev[i + 1]) for i in range(10)]
print("batch", B, "gpu ms/step", [round(x, 1) for x in gpu])
print("host ms/step", [round(x, 1) for x in host])
_lib.start_profile()
for _ in range(3):
    tr.step()
prof = _lib.stop_profile()
tot = sum(v["ms"] for v in prof.values()) / 3
print("libefl kernels ms/step", round(tot, 1), "launches/step", sum(v["calls"] for v in prof.values()) / 3)
if os.environ.get("CPROFILE", "0") == "1":
    import cProfile, pstats
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(5):
        tr.step()
    pr.disable()
    torch.cuda.synchronize()
    st = pstats.Stats(pr)
    st.sort_stats("tottime").print_stats(45)
    st.sort_stats("cumulative").print_stats(70)
    sys.exit(0)
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as p:
    for _ in range(3):
        tr.step()
    torch.cuda.synchronize()
print(p.key_averages().table(sort_by="self_cpu_time_total", row_limit=35, max_name_column_width=60))
print(p.key_averages().table(sort_by="self_cuda_time_total", row_limit=30, max_name_column_width=60))
