"""Micro-benchmark of the edge path of one SO2EquivariantGraphAttention layer (forward + backward) on a batch of
synthetic radius-graph complexes: per entry point CUDA-event time, algorithmic bytes / FLOPs and the achieved
GB/s or TFLOP/s (the same accounting bench.py uses for its roofline entry).

  python tools/edge_bench.py [--batch 8] [--iters 3] [--profile]     (--profile: one pass, no warm-up, for ncu)
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from singa_b200 import _lib, ops
from singa_b200 import synthetic as syn
from singa_b200 import EF_layers as ef
from singa_b200.embedding import DEFAULT_CONFIG, EquivariantEmbedding


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--profile", action="store_true")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    ops.set_precision(args.precision)
    graphs = [syn.make_complex(i) for i in range(args.batch)]
    g = {k: v.to(dev) for k, v in syn.collate(graphs).items()}
    torch.manual_seed(0)
    model = EquivariantEmbedding(DEFAULT_CONFIG, device=dev)
    syn.load_portable_parameters(model, 0)
    ga = model.blocks[0].ga
    ei = g["ei_pp"]
    pos = g["pos_p"]
    N, E = pos.shape[0], ei.shape[1]
    vec = pos[ei[0]] - pos[ei[1]]
    rot = ef.init_edge_rot_mat(vec, device=dev)
    model.SO3_rotation[0].set_wigner(rot)
    x = ef.SO3_Embedding(0, [6], 16, dtype=torch.float32, device=dev)
    x.set_embedding(torch.randn(N, 49, 16, device=dev).requires_grad_(True))
    x_edge = torch.randn(E, 48, device=dev).requires_grad_(True)
    z = g["z_p"]
    R = torch.randn(N, 49, 16, device=dev)

    def once():
        for p in model.parameters():
            p.grad = None
        out = ga(x, z, x_edge, ei, hetero=False)
        (out.embedding * R).sum().backward()

    warm = 0 if args.profile else 2
    iters = 1 if args.profile else args.iters
    for _ in range(warm):
        once()
    torch.cuda.synchronize()
    _lib.start_profile()
    for _ in range(iters):
        once()
    prof = _lib.stop_profile()
    peaks = {"hbm": 6557.4, "tensor": 1356.3}
    pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.isfile(pk):
        p = json.load(open(pk))
        peaks = {"hbm": p["hbm_gbs"], "tensor": p.get("bf16_tflops_sustained", p["bf16_tflops"])}
    print(json.dumps({"nodes": N, "edges": E, "iters": iters, "precision": args.precision}))
    tot = 0.0
    for name, d in sorted(prof.items(), key=lambda kv: -kv[1]["ms"]):
        ms = d["ms"] / iters
        tot += ms
        line = {"entry": name, "ms": round(ms, 4), "calls": d["calls"] // iters}
        if d["work"] > 0:
            tensor = "gemm" in name or "s2act_tc" in name
            rate = d["work"] / iters / (ms * 1e-3) / (1e12 if tensor else 1e9)
            line.update({"achieved": round(rate, 1), "unit": "TFLOP/s" if tensor else "GB/s",
                         "frac_of_measured_peak": round(rate / (peaks["tensor"] if tensor else peaks["hbm"]), 3)})
            if name in ("efl_s2act_tc_fwd", "efl_s2act_tc_bwd"):
                # 70 FLOP/B, below the ridge (tensor peak / HBM peak ~ 207 FLOP/B): HBM is the binding roofline.
                # bf16 rows in and out: fwd 2 x [M,H] x 2 B, bwd 3 x; work = {4,6} G M H per edge, so bytes = work / G
                gbs = d["work"] / iters / 70.0 / (ms * 1e-3) / 1e9
                line.update({"hbm_achieved": round(gbs, 1), "hbm_frac_of_measured_peak": round(gbs / peaks["hbm"], 3)})
        print(json.dumps(line))
    print(json.dumps({"summary": "entry points, one layer fwd+bwd", "ms": round(tot, 3),
                      "edges_per_s": round(E / (tot * 1e-3))}))


if __name__ == "__main__":
    main()
