"""Per-kernel micro-benchmarks at the chunk size of the bench workload (262 144 edges): every SO(2) GEMM shape of
the edge path (forward, dgrad, wgrad) on the tcgen05 engine, timed with CUDA events on the launching stream after
warm-up; operands are far larger than the 126 MB L2.  Prints one JSON line per shape and a summary.

  python tools/kernel_bench.py [--ec 262144] [--iters 5] [--only nt|tn]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from singa_b200 import tc_gemm


WARM = 2


def timeit(fn, iters):
    for _ in range(WARM):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ec", type=int, default=262144)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--only", default="")
    ap.add_argument("--profile", action="store_true", help="one launch per shape, no warm-up (for ncu)")
    args = ap.parse_args()
    if args.profile:
        global WARM
        WARM, args.iters = 0, 1
    dev = torch.device("cuda", 0)
    Ec = args.ec
    peak = 1356.3
    pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.isfile(pk):
        peak = json.load(open(pk)).get("bf16_tflops_sustained", peak)
    rnd = lambda m, n: (torch.randn(m, n, device=dev) * 0.1).to(torch.bfloat16)
    # (name, epilogue, M, N, K): round-2 shapes, every SO(2) linear as a real GEMM over M = edges ([x_re | x_im] rows)
    nt = [("conv1.m0 fwd", 1, Ec, 1248, 240), ("conv1.m1 fwd", 0, Ec, 1536, 384), ("conv1.m2 fwd", 0, Ec, 1280, 320),
          ("conv2.m0 fwd", 0, Ec, 784, 896), ("conv2.m1 fwd", 0, Ec, 1344, 1536), ("conv2.m2 fwd", 0, Ec, 1120, 1280),
          ("conv2.m0 dgrad", 0, Ec, 896, 784), ("conv2.m1 dgrad", 0, Ec, 1536, 1344), ("conv2.m2 dgrad", 0, Ec, 1280, 1120),
          ("conv1.m0 dgrad", 0, Ec, 240, 1248), ("conv1.m1 dgrad", 0, Ec, 384, 1536), ("conv1.m2 dgrad", 0, Ec, 320, 1280)]
    tn = [("conv2.m0 wgrad", Ec, 784, 896), ("conv2.m1 wgrad", Ec, 1344, 1536), ("conv2.m2 wgrad", Ec, 1120, 1280),
          ("conv1.m0 wgrad", Ec, 1248, 240), ("conv1.m1 wgrad", Ec, 1536, 384), ("conv1.m2 wgrad", Ec, 1280, 320)]
    tot_ms = tot_fl = 0.0
    if args.only in ("", "nt"):
        for name, epi, M, N, K in nt:
            A, B = rnd(M, K), rnd(N, K)
            bias = torch.randn(N, device=dev) if name == "conv2.m0 fwd" else None
            ns = 352 if epi == 1 else 0
            nout = N // 2 if epi == 2 else N - ns
            C = torch.empty(M, nout, dtype=torch.bfloat16, device=dev)
            C0 = torch.empty(M, ns, dtype=torch.float32, device=dev) if epi == 1 else None
            ms = timeit(lambda: tc_gemm.gemm_nt(epi, A, B, bias, C, nout, C0, ns, ns, M, N, K), args.iters)
            fl = 2.0 * M * N * K
            byt = (M * K + N * K + M * nout) * 2 + M * ns * 4
            tot_ms += ms
            tot_fl += fl
            print(json.dumps({"kernel": "efl_tc_gemm_nt", "shape": name, "M": M, "N": N, "K": K, "epi": epi,
                              "ms": round(ms, 4), "tflops": round(fl / ms / 1e9, 1), "frac_of_measured_peak": round(fl / ms / 1e9 / peak, 3),
                              "algorithmic_GBs": round(byt / ms / 1e6, 1)}))
            del A, B, C, C0
    if args.only in ("", "tn"):
        for name, Mred, N1, N2 in tn:
            A, B = rnd(Mred, N1), rnd(Mred, N2)
            C = torch.zeros(N1, N2, dtype=torch.float32, device=dev)
            ms = timeit(lambda: tc_gemm.gemm_tn(A, B, C, True, Mred, N1, N2), args.iters)
            fl = 2.0 * Mred * N1 * N2
            tot_ms += ms
            tot_fl += fl
            print(json.dumps({"kernel": "efl_tc_gemm_tn", "shape": name, "Mred": Mred, "N1": N1, "N2": N2, "ms": round(ms, 4),
                              "tflops": round(fl / ms / 1e9, 1), "frac_of_measured_peak": round(fl / ms / 1e9 / peak, 3),
                              "algorithmic_GBs": round((Mred * (N1 + N2)) * 2 / ms / 1e6, 1)}))
            del A, B, C
    if args.only == "tnx":
        # wgrad orientation experiment: C = A^T B as is, or C^T = B^T A (swapped), single-CTA or CTA-pair tiles
        from singa_b200 import _lib
        for name, Mred, N1, N2 in tn:
            for swapped in (0, 1):
                n1, n2 = (N2, N1) if swapped else (N1, N2)
                A, B = rnd(Mred, n1), rnd(Mred, n2)
                C = torch.zeros(n1, n2, dtype=torch.float32, device=dev)
                for pm in (0, 2):
                    prev = _lib.LIB.efl_tc_gemm_pair_mode(pm)
                    ms = timeit(lambda: tc_gemm.gemm_tn(A, B, C, True, Mred, n1, n2), args.iters)
                    _lib.LIB.efl_tc_gemm_pair_mode(prev)
                    print(json.dumps({"shape": name, "swapped": swapped, "pair": pm, "N1": n1, "N2": n2, "ms": round(ms, 4),
                                      "tflops": round(2.0 * Mred * n1 * n2 / ms / 1e9, 1)}))
                del A, B, C
        return
    print(json.dumps({"summary": "all shapes", "ms": round(tot_ms, 3), "tflops": round(tot_fl / tot_ms / 1e9, 1),
                      "peak_tflops_measured_sustained": peak}))


if __name__ == "__main__":
    main()
