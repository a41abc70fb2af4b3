"""BASELINE.json config 5: large-graph stress -- whole-protein complexes of ~8 000 atoms, lmax = 8 (L = 81, M = 39),
8 A radius graph built on the GPU by the cell-list kernel (singa_b200.radius), forward + backward + Adam of the full
EquivariantEmbedding stack (4 passes x 3 layers) on ONE GPU.  Under torchrun every rank runs its own complexes (the
batch is sharded by whole complex) and gradients are all-reduced; rank 0 prints one JSON line.

  python tools/stress_bench.py [--atoms 8000] [--complexes 1] [--lmax 8] [--steps 3] [--warmup 2]
"""
import argparse
import json
import datetime
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from singa_b200 import _lib, ops
from singa_b200 import synthetic as syn
from singa_b200.embedding import DEFAULT_CONFIG, EquivariantEmbedding
from singa_b200.parallel import GradientBuckets
from singa_b200.radius import edge_csr_from_radius, radius_graph_csr


def ball(rs, n, radius):
    """n points uniform in a ball (no minimum separation: the stress case only needs the density)."""
    pts = np.zeros((0, 3))
    while len(pts) < n:
        c = rs.uniform(-radius, radius, size=(2 * n, 3))
        pts = np.concatenate((pts, c[np.linalg.norm(c, axis=1) <= radius]))
    return pts[:n].astype(np.float32)


def make_batch(rank, complexes, atoms, dev, cutoff=8.0, contact=4.0, density=0.045):
    rs = np.random.RandomState(5000 + rank)
    radius = (3.0 * atoms / (4.0 * np.pi * density)) ** (1.0 / 3.0)
    pos_p, pos_l, bp, bl = [], [], [], []
    for c in range(complexes):
        n_l = int(rs.randint(20, 46))
        off = np.array([1000.0 * c, 0, 0], dtype=np.float32)  # complexes far apart; batch ids keep them disjoint anyway
        pos_p.append(ball(rs, atoms, radius) + off)
        pos_l.append(ball(rs, n_l, 5.0) + off)
        bp.append(np.full(atoms, c))
        bl.append(np.full(n_l, c))
    pos_p, pos_l = np.concatenate(pos_p), np.concatenate(pos_l)
    n_p, n_l = len(pos_p), len(pos_l)

    def feats(n):
        x = np.zeros((n, 59), dtype=np.float32)
        x[:, -15:] = rs.randint(0, 2, size=(n, 15))
        x[:, -8] = rs.uniform(-0.5, 0.5, size=n)
        return x

    g = {"z_p": rs.choice([6, 7, 8, 16], size=n_p, p=[0.6, 0.17, 0.2, 0.03]).astype(np.int64),
         "z_l": rs.choice([6, 7, 8, 16], size=n_l, p=[0.6, 0.17, 0.2, 0.03]).astype(np.int64),
         "x_p": feats(n_p), "x_l": feats(n_l), "pos_p": pos_p, "pos_l": pos_l}
    g = {k: torch.from_numpy(v).to(dev) for k, v in g.items()}
    bpt = torch.from_numpy(np.concatenate(bp)).to(dev)
    blt = torch.from_numpy(np.concatenate(bl)).to(dev)
    # the cell-list kernel's destination-sorted CSR feeds the attention path directly (no sort on the step)
    for key, pos, bt in (("ei_pp", g["pos_p"], bpt), ("ei_ll", g["pos_l"], blt)):
        rc = radius_graph_csr(pos, cutoff, bt, max_num_neighbors=1 << 30)
        edge_csr_from_radius(rc, pos.shape[0])
        g[key] = rc.edge_index
    d = torch.cdist(g["pos_l"], g["pos_p"])
    li, pi = torch.nonzero((d <= contact) & (blt[:, None] == bpt[None, :]), as_tuple=True)
    g["ei_lp"] = torch.stack((li, pi))
    g["ei_pl"] = torch.stack((pi, li))
    return g


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--atoms", type=int, default=8000)
    ap.add_argument("--complexes", type=int, default=1, help="complexes per GPU")
    ap.add_argument("--lmax", type=int, default=8)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--precision", default="bf16", choices=["fp32", "bf16"])
    ap.add_argument("--profile", action="store_true", help="CUDA-event time per entry point (rank 0)")
    args = ap.parse_args()
    line = run(args)
    if line is not None:
        print(json.dumps(line), flush=True)


def run(args):
    """The measurement itself (also behind ``bench.py --config 5``); returns the result dict on rank 0."""
    import torch.distributed as dist
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=300))
    ops.set_precision(args.precision)
    g = make_batch(rank, args.complexes, args.atoms, dev)
    edges = int(sum(g[k].shape[1] for k in ("ei_pp", "ei_ll", "ei_lp", "ei_pl")))
    cfg = dict(DEFAULT_CONFIG)
    cfg["lmax_list"] = [args.lmax]
    torch.manual_seed(0)
    model = EquivariantEmbedding(cfg, device=dev)
    syn.load_portable_parameters(model, 0)
    opt = torch.optim.Adam(model.parameters(), lr=1e-4)
    bucket = GradientBuckets(model.named_parameters())
    view = syn.GraphView(g)

    def step():
        bucket.zero_grad()
        out = model(view)
        loss = out[syn.PROTEIN].embedding.square().mean() + out[syn.LIGAND].embedding.square().mean()
        loss.backward()
        bucket.finish()
        opt.step()
        return loss

    for _ in range(max(args.warmup, 1)):
        loss = step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = _lib.launch_count
    if args.profile:
        _lib.start_profile()
    hooks = getattr(args, "timed_region_hooks", None)   # (on_start, on_stop): bench.py's nvidia-smi clock sampler
    if hooks:
        hooks[0]()
    a.record()
    for _ in range(args.steps):
        loss = step()
    b.record()
    torch.cuda.synchronize()
    if hooks:
        hooks[1]()
    prof = _lib.stop_profile() if args.profile else {}
    ms = torch.tensor([a.elapsed_time(b) / args.steps], device=dev)
    tot = torch.tensor([float(edges)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot)
    finite = bool(torch.isfinite(loss))
    line = None
    if rank == 0:
        line = ({
            "workload": f"config 5 stress: {args.complexes} x {args.atoms}-atom complexes per GPU, lmax {args.lmax}, "
                        f"mmax 2, 8 A radius graph from the cell-list kernel, fwd+bwd+Adam, {args.precision}",
            "n_gpus": world, "edges_total": int(tot), "edges_per_gpu": edges, "nodes_per_gpu": int(g["z_p"].numel()),
            "mean_degree": round(g["ei_pp"].shape[1] / g["z_p"].numel(), 2), "ms_per_step": round(float(ms), 3),
            "edges_per_s": round(float(tot) / (float(ms) * 1e-3)), "loss_finite": finite,
            "gpu_launches_per_step": (_lib.launch_count - n0) // args.steps,
            "peak_mem_GB": round(torch.cuda.max_memory_allocated() / 2 ** 30, 2),
            "kernel_ms_per_step": {k: round(v["ms"] / args.steps, 3)
                                   for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])}})
    if world > 1:
        dist.destroy_process_group()
    return line


if __name__ == "__main__":
    main()
