"""BASELINE.json config 4: radius_graph + Wigner-D micro-benchmark on a 1 M-atom synthetic point cloud
(uniform in a cube at 0.045 atoms / A^3), r in {5, 8, 10} A.  CUDA-event times per stage, achieved GB/s against the
algorithmic bytes of SURVEY.md 8(d): radius graph = 12 B/point read + 16 B/edge written (int64 edge_index);
Wigner = 36 B read + WSZ*4 B written per edge.

  python tools/radius_bench.py [--points 1000000] [--iters 3] [--profile]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from singa_b200 import EF_layers as ef
from singa_b200 import _lib, ops
from singa_b200.radius import radius_graph_csr


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=1_000_000)
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--radii", default="5,8,10")
    ap.add_argument("--profile", action="store_true")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    n = args.points
    side = (n / 0.045) ** (1.0 / 3.0)
    gen = torch.Generator(device=dev).manual_seed(0)
    pos = torch.rand(n, 3, device=dev, generator=gen) * side
    hbm = 6557.4
    pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.isfile(pk):
        hbm = json.load(open(pk))["hbm_gbs"]
    for r in [float(v) for v in args.radii.split(",")]:
        iters = 1 if args.profile else args.iters
        if not args.profile:
            radius_graph_csr(pos, r, max_num_neighbors=1 << 30)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        _lib.start_profile()
        a.record()
        for _ in range(iters):
            g = radius_graph_csr(pos, r, max_num_neighbors=1 << 30)
        b.record()
        prof = _lib.stop_profile()
        ms = a.elapsed_time(b) / iters
        E = g.edge_index.shape[1]
        alg = n * 12.0 + E * 16.0
        line = {"op": "radius_graph", "points": n, "r": r, "edges": E, "mean_degree": round(E / n, 2),
                "max_degree": g.max_degree, "ms_total_incl_host_syncs": round(ms, 3),
                "edges_per_s": round(E / (ms * 1e-3)), "algorithmic_GBs": round(alg / (ms * 1e-3) / 1e9, 1),
                "frac_of_measured_hbm": round(alg / (ms * 1e-3) / 1e9 / hbm, 4),
                "stage_ms": {k: round(v["ms"] / iters, 3) for k, v in prof.items()}}
        print(json.dumps(line), flush=True)
        # Wigner-D rows of every edge (lmax 6, mmax 2), in slabs that fit comfortably
        ei = g.edge_index
        slab = 4_000_000
        tot_ms, done = 0.0, 0
        wsz = ops.wigner_size(6, 2)
        for s0 in range(0, min(E, 3 * slab), slab):
            sl = slice(s0, min(E, s0 + slab))
            vec = pos[ei[0, sl]] - pos[ei[1, sl]]
            torch.manual_seed(0)
            rot = ef.init_edge_rot_mat(vec, device=dev)
            ops.wigner_from_rotmat(rot, 6, 2)
            torch.cuda.synchronize()
            a.record()
            w = ops.wigner_from_rotmat(rot, 6, 2)
            b.record()
            torch.cuda.synchronize()
            tot_ms += a.elapsed_time(b)
            done += rot.shape[0]
            del w, rot, vec
        if done:
            byt = done * (36.0 + wsz * 4.0)
            print(json.dumps({"op": "wigner_from_rotmat", "r": r, "edges": done, "ms": round(tot_ms, 3),
                              "edges_per_s": round(done / (tot_ms * 1e-3)),
                              "algorithmic_GBs": round(byt / (tot_ms * 1e-3) / 1e9, 1),
                              "frac_of_measured_hbm": round(byt / (tot_ms * 1e-3) / 1e9 / hbm, 4)}), flush=True)
        del g, ei


if __name__ == "__main__":
    main()
