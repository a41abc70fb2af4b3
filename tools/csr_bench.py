"""Wall-clock of ops.build_csr (efl_csr_from_edge_index) on the bench batch's four edge lists, beside the torch
sort/bincount formulation it replaced (tool only)."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from singa_b200 import ops  # noqa: E402


def torch_csr(edge_index, num_src, num_dst):
    src, dst = edge_index[0], edge_index[1]
    dst_s, perm = torch.sort(dst, stable=True)
    src_s = src[perm]
    rowptr = torch.zeros(num_dst + 1, dtype=torch.int64, device=dst.device)
    rowptr[1:] = torch.cumsum(torch.bincount(dst, minlength=num_dst), 0)
    _, src_idx = torch.sort(src_s, stable=True)
    src_ptr = torch.zeros(num_src + 1, dtype=torch.int64, device=dst.device)
    src_ptr[1:] = torch.cumsum(torch.bincount(src, minlength=num_src), 0)
    out = [t.to(torch.int32) for t in (perm, rowptr, src_s, dst_s, dst, src_ptr, src_idx)]
    rowptr.cpu()
    return out


def wall(fn, n=10):
    fn()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t) / n * 1e3


def main():
    dev = torch.device("cuda:0")
    for E, N in ((900_000, 14_400), (30_000, 15_500), (2_500, 1_100)):
        g = torch.Generator().manual_seed(E)
        ei = torch.stack((torch.randint(0, N, (E,), generator=g), torch.randint(0, N, (E,), generator=g))).to(dev)
        print(json.dumps({"E": E, "N": N, "cuda_ms": round(wall(lambda: ops.build_csr(ei, N, N)), 3),
                          "torch_ms": round(wall(lambda: torch_csr(ei, N, N)), 3)}))


if __name__ == "__main__":
    main()
