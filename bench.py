#!/usr/bin/env python
"""bench.py -- EF_layers forward+backward throughput on synthetic CrossDocked-shaped complexes.

One "step" = one pass of the hot path over one batch: EquivariantEmbedding forward (4 passes x 3 TransBlockV2
layers: frames -> Wigner rows -> edge-degree embedding -> attention blocks -> norms) + backward of a fixed linear
loss + gradient all-reduce (N>1) + Adam update.  Workload = BASELINE.json configs[1]: batch 32 complexes of
~450 protein atoms, 8 A radius graph (~28 k edges each), lmax 6, mmax 2, per GPU (weak scaling).
Metric: edges/s = directed edges of the batch (all four edge sets) pushed through the 3-layer stack, fwd+bwd, per
second, summed over ranks.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision fp32|bf16]
                  [--scaling weak|strong] [--global-batch B]

--scaling strong splits a FIXED global batch (default 64 complexes = BASELINE.json config 3's batch) over the ranks
with parallel.shard_by_edges; the default (weak) gives every rank its own 32 complexes.

Prints ONE JSON line on rank 0 (see DESIGN.md "Measurement" for every key).
"""
from __future__ import annotations

import argparse
import json
import datetime
import os
import subprocess
import sys
import threading
import time

# The step allocates and frees tens of GB of chunk workspaces; with the default allocator an unlucky interleaving
# fragments the pool and the step pays cudaFree/cudaMalloc round trips in the timed region (seen as 240-300 ms steps
# on some boxes).  Expandable segments grow in place instead.  Must be set before the first CUDA allocation.
os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "EF_layers fwd+bwd edges/sec"
DOMINANT = "efl_tc_gemm_nt"  # entry point with the largest share of a step (checked against the full profile)
UNIT = "edges/s"
WORKLOAD = ("EF_layers fwd+bwd, batch 32 synthetic CrossDocked-shaped complexes (~450 atoms, 8 A cutoff, lmax=6, mmax=2) "
            "per GPU")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("EFL_PRECISION", "bf16"), choices=["fp32", "bf16", "bf16x3"])
    ap.add_argument("--batch", type=int, default=32, help="complexes per GPU")
    ap.add_argument("--atoms", type=int, default=None, help="fix protein atoms per complex (default U{400..500})")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--config", type=int, default=2, choices=[2, 3, 5],
                    help="BASELINE.json configs index: 2 = EF_layers fwd+bwd (the metric's config, default), 3 = full "
                         "SINGA train step (EquivariantEmbedding + CProMG transformer), global batch 64 over the ranks, "
                         "5 = large-graph stress (8 000-atom complexes, lmax 8, one complex per GPU)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--global-batch", type=int, default=64, help="complexes of the whole job under --scaling strong")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------
LAYERS = 3


def make_batch(rank: int, batch: int, atoms):
    from singa_b200 import synthetic as syn
    graphs = [syn.make_complex(1000 * rank + i, n_protein=atoms) for i in range(batch)]
    return syn.collate(graphs), graphs


def count_edges(g) -> int:
    """All four edge sets (the size of the input)."""
    return int(sum(g[k].shape[1] for k in ("ei_pp", "ei_ll", "ei_lp", "ei_pl")))


def count_edge_layers(g) -> int:
    """Edge-layers actually computed per step: the homogeneous passes push their edges through all LAYERS attention
    blocks; in the two hetero passes only the LAST block's output is observable (Embedding.py:415-428 keeps one), so
    with compute_dead_blocks=False (both arms) their edges pass through ONE block."""
    return int(LAYERS * (g["ei_pp"].shape[1] + g["ei_ll"].shape[1]) + g["ei_lp"].shape[1] + g["ei_pl"].shape[1])


def metric_edges(g) -> float:
    """The metric's unit: directed edges through the 3-layer stack = edge-layers computed / 3 (round 1 counted the
    hetero edges at 3 layers, ~2 % in its own favour: VERDICT r1 weak #10)."""
    return count_edge_layers(g) / float(LAYERS)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines, self.first = index, None, [], 0

    def mark(self):
        """Samples from here on count (the sampler is started before the warm-up)."""
        self.first = len(self.lines)

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines[self.first:]:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        p = json.load(open(path))
        return dict(hbm=p["hbm_gbs"], tensor=p.get("bf16_tflops_sustained", p["bf16_tflops"]), source="measured")
    return dict(hbm=6650.0, tensor=1400.0, source="fallback")


# ----------------------------------------------------------------------------------------------
# reference / CPU arm: the oracle port on the host cores (the reference is pure Python and cannot travel)
# ----------------------------------------------------------------------------------------------
REF_SAMPLE_ATOMS = 200      # the reference arm ALWAYS times this cut (a budget-sized sample moved the ratio 16 %)
PARITY_SAMPLE_ATOMS = 460   # cpu_baseline leg of our arm: one full config-2-sized complex, also the parity check


def cpu_sample_graph(n_protein: int):
    """Complex seed 0 with a fixed number of protein atoms (+ 24 ligand atoms), 8 A radius graph."""
    from singa_b200 import synthetic as syn
    return syn.make_complex(0, n_protein=n_protein, n_ligand=24), n_protein


def oracle_frames(graph, seed: int = 1):
    from oracle import ef_oracle as eo
    from singa_b200 import synthetic as syn
    rnd = syn.frame_randoms(graph, seed)
    frames = {}
    for key, (s, d, ei) in dict(pp=("pos_p", "pos_p", "ei_pp"), ll=("pos_l", "pos_l", "ei_ll"),
                                lp=("pos_l", "pos_p", "ei_lp")).items():
        frames[key] = eo.edge_rot_mat(graph[s][graph[ei][0]] - graph[d][graph[ei][1]], rnd[key])
    return frames


def loss_weights(graph, seed: int = 7):
    rs = np.random.RandomState(seed)
    f = lambda n: torch.from_numpy(rs.normal(size=(n, 49, 16)).astype(np.float32))
    return f(graph["z_p"].numel()), f(graph["z_l"].numel())


def run_cpu_port(graph, steps: int, warmup: int, state=None, keep=False):
    """fwd+bwd of oracle/ef_oracle.py (the CPU port of the reference algorithm) on all host threads, hetero passes
    through the last block only (compute_dead_blocks=False, as the GPU arm).  ``state``: a state_dict to run with
    (default: portable parameters, seed 0).  ``keep``: also return outputs and parameter gradients of the last step."""
    from oracle import ef_oracle as eo
    from singa_b200 import so3_constants as sc
    from singa_b200 import synthetic as syn
    from singa_b200.embedding import DEFAULT_CONFIG, EquivariantEmbedding
    torch.set_num_threads(os.cpu_count() or 1)
    torch.manual_seed(0)
    shapes = EquivariantEmbedding(DEFAULT_CONFIG, device="cpu")
    if state is None:
        syn.load_portable_parameters(shapes, 0)
        state = shapes.state_dict()
    names = set(dict(shapes.named_parameters()))
    p = {k: v.detach().cpu().clone().requires_grad_(v.dtype.is_floating_point and k in names) for k, v in state.items()}
    f32 = lambda a: torch.tensor(a, dtype=torch.float32)
    to62, fr62 = sc.s2_grid_matrices(6, 2)
    to66, fr66 = sc.s2_grid_matrices(6, 6)
    t = eo.Tables(6, 2, [sc.j_matrix(l) for l in range(7)], f32(to62), f32(fr62), f32(to66), f32(fr66))
    frames = oracle_frames(graph)
    Rp, Rl = loss_weights(graph)
    times, out = [], None
    for it in range(warmup + steps):
        for v in p.values():
            v.grad = None
        t0 = time.perf_counter()
        out = eo.embedding_forward(p, graph, frames, t, dict(DEFAULT_CONFIG), compute_dead_blocks=False)
        loss = (out["protein_atoms"] * Rp).sum() + (out["ligand_atoms"] * Rl).sum()
        loss.backward()
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
    sec = float(np.mean(times))
    if keep:
        return sec, {k: v.detach() for k, v in out.items()}, {k: v.grad for k, v in p.items() if v.grad is not None}
    return sec


def parity_against_port(model, graph, outs_ref, grads_ref, dev):
    """The SAME complex through the CUDA path in the precision being benchmarked, same parameters, same frames, same
    loss: relative errors (||d||_inf / ||ref||_inf) of the outputs and of every parameter gradient."""
    from singa_b200 import synthetic as syn
    frames = {k: v.to(dev) for k, v in oracle_frames(graph).items()}
    Rp, Rl = (r.to(dev) for r in loss_weights(graph))
    model.zero_grad(set_to_none=True)
    out = model(syn.GraphView({k: v.to(dev) for k, v in graph.items()}), frames=frames)
    loss = (out[syn.PROTEIN].embedding * Rp).sum() + (out[syn.LIGAND].embedding * Rl).sum()
    loss.backward()

    def rel(a, b):
        a, b = a.detach().double().cpu(), b.detach().double().cpu()
        return float((a - b).abs().max()) / max(float(b.abs().max()), 1e-30)

    e_out = {k: rel(out[k].embedding, outs_ref[k]) for k in outs_ref}
    e_grad = {n: rel(p.grad, grads_ref[n]) for n, p in model.named_parameters() if p.grad is not None and n in grads_ref}
    worst = max(e_grad, key=e_grad.get)
    srt = sorted(e_grad.values())
    model.zero_grad(set_to_none=True)
    return {"out_rel_err": max(e_out.values()), "grad_rel_err": e_grad[worst], "grad_rel_err_median": srt[len(srt) // 2],
            "worst_grad": worst, "outputs": e_out, "n_grads": len(e_grad),
            "against": "oracle/ef_oracle.py (CPU port, fp32) on the cpu_baseline sample, same parameters/frames/loss"}


def run_cpu_reference(graph, steps: int, warmup: int, state=None, keep=False):
    """fwd+bwd of the UNMODIFIED reference ``EquivariantEmbedding`` (model/Embedding.py:205-480 over model/EF_layers.py,
    loaded by oracle/ref_loader.py from /root/reference or the build-time copy in oracle/_ref) on all host threads: the
    reference's stock path, every block of every pass included.  Same parameters (portable seed 0), edge frames
    (the rand_like draws of init_edge_rot_mat are fed from synthetic.frame_randoms seed 1) and loss as run_cpu_port."""
    from oracle import ref_loader as rl
    from singa_b200 import synthetic as syn
    ef, emb = rl.load_reference()
    torch.set_num_threads(os.cpu_count() or 1)
    torch.manual_seed(0)
    ref = emb.EquivariantEmbedding(rl.EasyDict(rl.DEFAULT_EMBEDDING_CONFIG), device="cpu")
    if state is None:
        syn.load_portable_parameters(ref, 0)
    else:   # the B200 modules keep the reference's state_dict surface: strict load
        ref.load_state_dict({k: v.detach().cpu() for k, v in state.items()}, strict=True)
    rnd = syn.frame_randoms(graph, 1)
    Rp, Rl = loss_weights(graph)
    times, out = [], None
    orig = torch.rand_like
    try:
        for it in range(warmup + steps):
            ref.zero_grad(set_to_none=True)
            draws = iter([rnd["pp"], rnd["ll"], rnd["lp"]])
            torch.rand_like = lambda x, **k: next(draws)
            t0 = time.perf_counter()
            out = ref(syn.GraphView(graph))
            loss = (out[syn.PROTEIN].embedding * Rp).sum() + (out[syn.LIGAND].embedding * Rl).sum()
            loss.backward()
            dt = time.perf_counter() - t0
            if it >= warmup:
                times.append(dt)
    finally:
        torch.rand_like = orig
    sec = float(np.mean(times))
    if keep:
        return (sec, {k: v.embedding.detach() for k, v in out.items()},
                {n: p.grad for n, p in ref.named_parameters() if p.grad is not None})
    return sec


def reference_kind() -> str:
    """"reference": the unmodified reference files are on this machine (/root/reference or oracle/_ref); else "port"."""
    from oracle import ref_loader as rl
    return "reference" if rl.reference_available() else "port"


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    graph, n = cpu_sample_graph(REF_SAMPLE_ATOMS)
    edges = metric_edges(graph)
    kind = reference_kind()
    cores = os.cpu_count() or 1
    what = ("UNMODIFIED reference EquivariantEmbedding (model/Embedding.py + model/EF_layers.py via oracle/ref_loader.py, "
            "e3nn / torch_geometric symbols from the loader's shims; all blocks of all passes, the stock path)"
            if kind == "reference" else "CPU port of the reference (oracle/ef_oracle.py; reference files not on this machine)")
    sec = run_cpu_reference(graph, args.steps, args.warmup) if kind == "reference" else run_cpu_port(graph, args.steps, args.warmup)
    val = edges / sec
    sample = (f"FIXED sample: synthetic complex seed 0 cut to {n} protein + 24 ligand atoms, {count_edges(graph)} edges "
              f"({count_edge_layers(graph)} edge-layers), 8 A radius graph; {what}, fp32, {cores} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if kind == "reference":   # once, beside it: the port on the same sample, to show the two agree
        psec = run_cpu_port(graph, 1, 1)
        line["cpu_baseline"]["port_value"] = edges / psec
        line["cpu_baseline"]["port_note"] = ("oracle/ef_oracle.py on the same sample (skips the hetero passes' dead blocks, "
                                             "so it does less work than the stock reference path)")
    emit(line)


# ----------------------------------------------------------------------------------------------
_RESULT_FD = None


def _quiet_stdout():
    """The contract is ONE JSON line on stdout: libraries that write to fd 1 (NCCL prints its version there when
    NCCL_DEBUG is set on the box) are sent to stderr; the result line goes to the saved descriptor."""
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_RESULT_FD, data)


def settle_steps(fn, lo, hi, alloc_count, any_rank=None):
    """``lo`` warm-up steps, then more (at most ``hi``) until a step runs without a cudaMalloc: the timed region starts
    from a steady memory footprint.  ``fn`` contains collectives when N > 1 (the gradient all-reduce), so every rank
    must run the SAME number of steps: the stop decision goes through ``any_rank`` (a MAX all-reduce of the local
    flag, called once per step on every rank).  A rank-local decision deadlocks the job as soon as two ranks settle
    after a different number of steps (seen at N = 8: six ranks in the barrier of the timed region, two still in the
    warm-up's all-reduce)."""
    done = 0
    while done < lo + hi:
        n0 = alloc_count()
        fn()
        done += 1
        grew = alloc_count() != n0
        if any_rank is not None:
            grew = any_rank(grew)
        if done >= lo and not grew:
            break
    return done


def config3(args):
    """BASELINE.json configs[3]: full SINGA train step (EquivariantEmbedding + CProMG encoder/decoder, cross entropy,
    Adam), GLOBAL batch 64, data parallel with the bucketed NCCL gradient all-reduce.  Every step takes its batch from
    the CSR-carrying dataset format in pinned host memory (one H2D copy inside the timed region, loss read back at the
    end of the run), so ``value`` IS end to end.  The dense transformer blocks run on PyTorch library ops (out of scope
    for kernels); the message passing, kNN graph and EF_layers run on libefl.so."""
    import torch.distributed as dist
    from singa_b200 import _lib, ops
    from singa_b200 import dataset as ds
    from singa_b200 import synthetic as syn
    from singa_b200.gan import SINGA, default_config
    from singa_b200.train import Trainer
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=300))
    ops.set_precision(args.precision)
    gb = args.global_batch
    assert gb % world == 0, "global batch must divide over the ranks"
    data = ds.ComplexDataset(ds.synthetic_records(range(gb), n_protein=args.atoms))
    torch.manual_seed(0)
    model = SINGA(default_config(), device=dev)
    syn.load_portable_parameters(model.embedding, 0)
    syn.load_portable_parameters_fanin(model.model, 1)
    ops.reserve_workspace(0.6, dev)
    tr = Trainer(model, data, batch_size_per_rank=gb // world, shuffle=False, prefetch=2, device=dev)
    edges = sum(metric_edges({k: torch.empty(2, len(data[i][f"{k[3:]}_src"])) for k in ("ei_pp", "ei_ll", "ei_lp", "ei_pl")})
                for i in tr.loader.indices)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        tr.step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    n0 = _lib.launch_count
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.steps):
        loss, _ = tr.step()
    b.record()
    barrier()
    launches = _lib.launch_count - n0
    clocks = sampler.stop() if rank == 0 else None
    ms = torch.tensor([a.elapsed_time(b) / args.steps], device=dev)
    tot = torch.tensor([float(edges)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot)
    _lib.start_profile()
    for _ in range(2):
        tr.step()
    full = _lib.stop_profile()
    hb = next(tr.loader)
    if rank == 0:
        ms, tot = float(ms), float(tot)
        emit({
            "metric": METRIC, "value": tot / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
            "config": {"workload": f"BASELINE configs[3]: full SINGA train step (EF_layers embedding + CProMG "
                                   f"encoder/decoder + cross entropy + Adam), GLOBAL batch {gb} synthetic complexes, dp{world}",
                       "complexes_per_gpu": gb // world, "complexes_per_s": gb / (ms * 1e-3), "loss": float(loss),
                       "dense_transformer": "PyTorch library ops (out of scope for kernels, SURVEY.md 2 #5)",
                       "parameters": sum(p.numel() for p in model.parameters()),
                       "edge_count": "EF_layers edges (edge-layers computed / 3) of the global batch",
                       "sharding": "equal complexes per rank, edges balanced within that (parallel.shard_by_edges "
                                   "equal_counts=True); unshuffled, so every step is the same global batch"},
            "e2e": {"value": tot / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": hb.nbytes, "d2h_bytes_per_step": 0,
                    "note": "the timed step itself copies its pinned host batch (one buffer) to the device"},
            "gpu_launches": launches, "clocks": clocks,
            "kernel_ms_per_step": {k: round(v["ms"] / 2, 3) for k, v in sorted(full.items(), key=lambda kv: -kv[1]["ms"])},
        })
    if world > 1:
        dist.destroy_process_group()


def config5(args):
    """BASELINE.json configs[5]: 8 000-atom whole-protein complexes, lmax = 8, 8 A radius graph built on the GPU by the
    cell-list kernel (its CSR feeds the attention path without a sort), one complex per GPU, fwd + bwd + Adam of the
    4-pass stack; under torchrun the batch is sharded by complex and gradients are all-reduced (tools/stress_bench.py)."""
    import types
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import stress_bench
    a = types.SimpleNamespace(atoms=args.atoms or 8000, complexes=1, lmax=8, steps=args.steps,
                              warmup=max(args.warmup, 2), precision=args.precision, profile=True)
    # clocks and throttle reasons DURING the timed region, on rank 0's GPU (round 2's config-5 lines had none)
    rank0 = int(os.environ.get("RANK", "0")) == 0
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", "0")))
    clocks = {}
    if rank0:
        sampler.start()
        a.timed_region_hooks = (sampler.mark, lambda: clocks.update(sampler.stop()))
    r = stress_bench.run(a)
    if r is None:
        return
    ms = r["ms_per_step"]
    emit({"metric": METRIC, "value": r["edges_total"] / (ms * 1e-3), "unit": UNIT, "n_gpus": r["n_gpus"],
          "steps": args.steps, "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
          "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
          "config": {"workload": "BASELINE configs[5]: " + r["workload"], "edges_per_gpu": r["edges_per_gpu"],
                     "nodes_per_gpu": r["nodes_per_gpu"], "mean_degree": r["mean_degree"],
                     "peak_mem_GB": r["peak_mem_GB"], "loss_finite": r["loss_finite"],
                     "edge_count": "all four edge sets, as given (hetero edges counted once)"},
          "e2e": {"value": r["edges_total"] / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                  "note": "graph built on the device by the radius-graph kernel outside the timed region; inputs resident"},
          "gpu_launches": r["gpu_launches_per_step"] * args.steps, "clocks": clocks or None,
          "kernel_ms_per_step": r["kernel_ms_per_step"]})


def main():
    args = parse()
    _quiet_stdout()
    if args.impl == "reference":
        reference_arm(args)
        return
    if args.config == 3:
        config3(args)
        return
    if args.config == 5:
        config5(args)
        return
    import torch.distributed as dist
    from singa_b200 import _lib, ops
    from singa_b200 import synthetic as syn
    from singa_b200.embedding import DEFAULT_CONFIG, EquivariantEmbedding
    from singa_b200.parallel import GradientBuckets, shard_by_edges

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=300))
    ops.set_precision(args.precision)

    if args.scaling == "strong":
        # fixed global batch (seeds 0..B-1), whole complexes dealt to the ranks by edge count (parallel.shard_by_edges);
        # the per-complex cost proxy is the protein atom count make_complex(seed) draws first (edges ~ 63 per atom)
        sizes = [int(np.random.RandomState(i).randint(400, 501)) if args.atoms is None else args.atoms
                 for i in range(args.global_batch)]
        mine = shard_by_edges(sizes, world)[rank]
        graphs = [syn.make_complex(i, n_protein=args.atoms) for i in mine]
        n_local = len(mine)
    else:
        _, graphs = make_batch(rank, args.batch, args.atoms)
        n_local = args.batch
    batch = syn.collate(graphs)
    edges = metric_edges(batch)
    host = {k: v.pin_memory() for k, v in batch.items()}
    resident = {k: v.to(dev) for k, v in host.items()}
    h2d_bytes = int(sum(v.numel() * v.element_size() for v in host.values()))

    torch.manual_seed(0)
    model = EquivariantEmbedding(DEFAULT_CONFIG, device=dev)
    syn.load_portable_parameters(model, 0)
    opt = torch.optim.Adam(model.parameters(), lr=1e-4)
    buckets = GradientBuckets(model.named_parameters())   # gradients live in per-block flat buffers (views)
    rs = np.random.RandomState(7)
    Rp = torch.from_numpy(rs.normal(size=(resident["z_p"].numel(), 49, 16)).astype(np.float32)).to(dev)
    Rl = torch.from_numpy(rs.normal(size=(resident["z_l"].numel(), 49, 16)).astype(np.float32)).to(dev)

    def step(g):
        buckets.zero_grad()
        out = model(syn.GraphView(g))
        loss = (out[syn.PROTEIN].embedding * Rp).sum() + (out[syn.LIGAND].embedding * Rl).sum()
        loss.backward()        # each block's bucket starts its all-reduce as backward completes it (N > 1)
        buckets.finish()
        opt.step()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    per_step = {}

    def timed(fn, n, tag=None):
        """K steps between two events on the launching stream (the reported time); with ``tag`` an event after every
        step as well, so that the spread of the individual steps is on record (transients show up there)."""
        barrier()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
        evs[0].record()
        for i in range(n):
            fn()
            evs[i + 1].record()
        barrier()
        ms = torch.tensor([evs[0].elapsed_time(evs[n])], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if tag:
            per_step[tag] = [round(evs[i].elapsed_time(evs[i + 1]), 2) for i in range(n)]
        return float(ms) / n

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()   # before the warm-up: nvidia-smi's own start-up (NVML init) must not land in the timed region
    ops.reserve_workspace(0.6, dev)   # pool sized once, so that no step extends it (see ops.reserve_workspace)

    def any_rank(flag):
        """True on every rank when ``flag`` is true on at least one: keeps rank-local decisions collective."""
        if world == 1:
            return flag
        t = torch.tensor([1.0 if flag else 0.0], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return bool(t.item() > 0)

    def settle(fn, lo, hi=12):
        return settle_steps(fn, lo, hi, lambda: torch.cuda.memory_stats(dev).get("num_device_alloc", 0), any_rank)

    warm_done = settle(lambda: step(resident), max(args.warmup, 3))
    if rank == 0:
        sampler.mark()
    launches0 = _lib.launch_count
    mem0 = torch.cuda.memory_stats(dev)
    # Event pairs in the timed region only around the dominant entry point (the roofline's kernel): pairs around all
    # calls of a step cost enough host time to make the step host-bound on a slow box.  The per-entry-point
    # breakdown comes from a second, fully profiled pass of the same K steps below (outside the timed region).
    _lib.start_profile(only={DOMINANT})
    ms_dev = timed(lambda: step(resident), args.steps, "device")
    prof = _lib.stop_profile()
    launches = _lib.launch_count - launches0
    mem1 = torch.cuda.memory_stats(dev)
    allocator = {"cudaMalloc_calls_in_timed_region": int(mem1.get("num_device_alloc", 0) - mem0.get("num_device_alloc", 0)),
                 "cudaFree_calls_in_timed_region": int(mem1.get("num_device_free", 0) - mem0.get("num_device_free", 0)),
                 "alloc_retries": int(mem1.get("num_alloc_retries", 0) - mem0.get("num_alloc_retries", 0)),
                 "reserved_GB": round(torch.cuda.memory_reserved(dev) / 2 ** 30, 1),
                 "conf": os.environ.get("PYTORCH_CUDA_ALLOC_CONF", "")}

    def e2e_step():
        g = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
        return float(step(g).item())

    settle(e2e_step, 5)   # fresh input tensors every step: the caching allocator and the CSR cache reach steady state
    ms_e2e = timed(e2e_step, args.steps, "e2e")
    clocks = sampler.stop() if rank == 0 else None
    _lib.start_profile()
    ms_profiled = timed(lambda: step(resident), args.steps)
    full = _lib.stop_profile()

    tot = torch.tensor([float(edges), float(n_local), float(count_edges(batch))], device=dev)
    if world > 1:
        dist.all_reduce(tot)
    tot_edges, tot_complexes, tot_input_edges = (float(v) for v in tot)

    if rank == 0:
        pk = peaks()
        top = max(prof.items(), key=lambda kv: kv[1]["ms"]) if prof else (None, None)
        if full and max(full.items(), key=lambda kv: kv[1]["ms"])[0] != DOMINANT:  # another entry point took over
            top = max(full.items(), key=lambda kv: kv[1]["ms"])
        roofline = None
        if top[0] is not None:
            name, d = top
            tensor = "gemm" in name
            per_launch_work = d["work"] / max(d["calls"], 1)
            per_launch_s = d["ms"] * 1e-3 / max(d["calls"], 1)
            achieved = per_launch_work / per_launch_s / (1e12 if tensor else 1e9)
            peak = pk["tensor"] if tensor else pk["hbm"]
            traffic, traffic_note = None, None
            tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
            if os.path.isfile(tpath):
                tj = json.load(open(tpath))
                ent = tj.get(name)
                if isinstance(ent, dict):
                    traffic, traffic_note = ent.get("dram_bytes_per_launch"), ent.get("note")
                else:
                    traffic, traffic_note = ent, tj.get("_note")
            roofline = {"kernel": name, "bound": "tensor" if tensor else "hbm", "achieved": achieved, "peak": peak,
                        "unit": "TFLOP/s" if tensor else "GB/s", "frac": achieved / peak, "traffic": traffic,
                        "traffic_unit": traffic_note,
                        "algorithmic_per_launch": per_launch_work,
                        "algorithmic_unit": "FLOP (2*M*N*K)" if tensor else "bytes",
                        "algorithmic_bytes_per_launch": (d.get("bytes", 0.0) / max(d["calls"], 1)) if tensor
                        else per_launch_work,
                        "ms_per_launch": per_launch_s * 1e3,
                        "peak_source": pk["source"], "share_of_step": d["ms"] / (ms_dev * args.steps),
                        "launches": d["calls"]}
        breakdown = {k: round(v["ms"] / args.steps, 3) for k, v in sorted(full.items(), key=lambda kv: -kv[1]["ms"])}
        # whole-step tensor fraction: 36.0 MFLOP per edge-layer fwd+bwd over the step time
        step_tflops = 36.0e6 * tot_edges * LAYERS / world / (ms_dev * 1e-3) / 1e12
        line = {
            "metric": METRIC, "value": tot_edges / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "warmup_done": warm_done, "ms_per_step": ms_dev, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": {"bf16": "bf16", "bf16x3": "bf16x3 (error-compensated bf16 operands, fp32 storage)"}.get(
                args.precision, "f32"), "data": "synthetic",
            "config": {"workload": WORKLOAD if args.scaling == "weak" else
                       f"EF_layers fwd+bwd, GLOBAL batch {args.global_batch} synthetic CrossDocked-shaped complexes "
                       f"(~450 atoms, 8 A cutoff, lmax=6, mmax=2) split over the ranks by edge count",
                       "complexes_per_gpu": n_local, "complexes": tot_complexes,
                       "input_edges": tot_input_edges, "edge_layers_per_step": tot_edges * LAYERS,
                       "edge_count": "value counts edge-layers actually computed / 3: pp and ll edges through 3 blocks, "
                                     "lp and pl edges through the one live block",
                       "layers": LAYERS, "passes": 4, "parallelism": f"dp{world}",
                       "gemm_precision": args.precision, "l2": "inputs and per-chunk activations >> 126 MB L2",
                       "complexes_per_s": tot_complexes / (ms_dev * 1e-3)},
            "e2e": {"value": tot_edges / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e},
            "gpu_launches": launches, "roofline": roofline, "clocks": clocks, "allocator": allocator,
            "step_tensor_tflops": step_tflops, "step_tensor_frac": step_tflops / pk["tensor"],
            "step_ms_rank0": per_step,
            "kernel_ms_per_step": breakdown,
            "kernel_ms_per_step_source": f"second pass of {args.steps} steps with event pairs around every entry point "
                                         f"({ms_profiled:.1f} ms/step under that profile)",
        }
        if world == 1 and not args.no_cpu_baseline:
            # CPU port on ONE full config-2-sized complex, with the model's current parameters; the same complex then
            # goes through the CUDA path in the benchmarked precision: measured parity at config-2 size in every run
            graph, n = cpu_sample_graph(PARITY_SAMPLE_ATOMS)
            kind = reference_kind()
            run = run_cpu_reference if kind == "reference" else run_cpu_port
            sec, outs_ref, grads_ref = run(graph, 1, 0, state=model.state_dict(), keep=True)
            e = metric_edges(graph)
            line["cpu_baseline"] = {
                "value": e / sec, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": kind,
                "sample": f"one synthetic complex (seed 0) with {n} protein + 24 ligand atoms, {count_edges(graph)} "
                          f"edges ({count_edge_layers(graph)} edge-layers), fwd+bwd once, {sec:.1f} s"}
            line["parity"] = parity_against_port(model, graph, outs_ref, grads_ref, dev)
            line["parity"]["precision"] = args.precision
            if kind == "reference":
                line["parity"]["against"] = ("the UNMODIFIED reference EquivariantEmbedding on the host (oracle/ref_loader.py), "
                                             "fp32, on the cpu_baseline sample, same parameters/frames/loss")
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
