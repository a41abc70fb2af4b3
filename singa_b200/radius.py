"""radius_graph on the GPU: cell-list neighbour search emitting a destination-sorted CSR edge list
(csrc/radius_graph.cu).  Same call signature as torch_cluster.radius_graph (torch-cluster 1.6.3, the reference's
un-vendored dependency: gen.py:17), plus ``radius_graph_csr`` which also returns the CSR the
attention kernels consume.  No CPU path: tensors must live on a CUDA device.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import torch

from ._lib import LIB, call, check_cuda, ptr

CELL_MARGIN = 1.001  # cell edge = CELL_MARGIN * r: points closer than r always sit in adjacent cells (fp32 safe)


@dataclass
class RadiusCSR:
    edge_index: torch.Tensor   # [2, E] int64, row 0 sources, row 1 targets (targets ascending, sources ascending)
    rowptr: torch.Tensor       # [N+1] int32 over targets
    col: torch.Tensor          # [E] int32 sources
    max_degree: int            # largest degree before the max_num_neighbors cap
    truncated: bool            # True when the neighbour cap removed edges


def radius_graph_csr(x: torch.Tensor, r: float, batch: Optional[torch.Tensor] = None, loop: bool = False,
                     max_num_neighbors: int = 32, flow: str = 'source_to_target') -> RadiusCSR:
    assert flow in ('source_to_target', 'target_to_source')
    check_cuda(x, batch)
    assert x.dim() == 2 and x.shape[1] == 3, "radius_graph kernel handles 3-D positions"
    dev = x.device
    N = x.shape[0]
    if N == 0:
        z = torch.zeros(1, dtype=torch.int32, device=dev)
        return RadiusCSR(torch.empty(2, 0, dtype=torch.int64, device=dev), z, z[:0], 0, False)
    pos = x.detach().contiguous().float()
    if batch is not None:
        b32 = batch.to(torch.int32).contiguous()
        nbatch = int(b32[-1].item()) + 1  # torch_cluster requires a sorted batch vector
    else:
        b32, nbatch = None, 1
    r = float(r)
    # 1. per-batch bounding boxes -> grid dimensions (one small D2H read)
    box_ws = torch.empty(nbatch * 6, dtype=torch.int32, device=dev)
    box = torch.empty(nbatch, 6, dtype=torch.float32, device=dev)
    call("efl_radius_graph_bbox", ptr(pos), ptr(b32), N, nbatch, ptr(box_ws), ptr(box))
    boxh = box.cpu()
    if b32 is not None:
        assert bool(torch.isfinite(boxh).all()), "batch vector must be sorted and cover 0..B-1 without gaps"
    cs = r * CELL_MARGIN
    ext = (boxh[:, 3:] - boxh[:, :3]).amax(dim=0)
    nx, ny, nz = (max(1, int(math.floor(float(e) / cs)) + 1) for e in ext)
    ncells = nbatch * nx * ny * nz
    origin = box[:, :3].contiguous()
    # 2-5. bin, scan, scatter, count
    cell_of = torch.empty(N, dtype=torch.int32, device=dev)
    cell_start = torch.empty(ncells + 1, dtype=torch.int32, device=dev)
    scratch = torch.empty(2 * (ncells + 1), dtype=torch.int32, device=dev)
    sorted_pts = torch.empty(N, 4, dtype=torch.float32, device=dev)
    deg = torch.empty(2 * (N + 1), dtype=torch.int32, device=dev)
    rowptr_full = torch.empty(N + 1, dtype=torch.int32, device=dev)
    rowptr_cap = torch.empty(N + 1, dtype=torch.int32, device=dev)
    tbytes = int(LIB.efl_radius_graph_temp_bytes(max(N, ncells) + 1))
    temp = torch.empty(max(tbytes, 16), dtype=torch.uint8, device=dev)
    cap = min(int(max_num_neighbors), 1 << 30)  # the kernel forms cap + 1 (upstream searches with one extra slot for self)
    call("efl_radius_graph_count", ptr(pos), ptr(b32), N, nbatch, ptr(origin), cs, nx, ny, nz, r, int(loop), cap,
         ptr(cell_of), ptr(cell_start), ptr(scratch), ptr(sorted_pts), ptr(deg), ptr(rowptr_full), ptr(rowptr_cap),
         ptr(temp), tbytes, work=N * 16.0 * 27)
    totals = torch.stack((rowptr_full[N], rowptr_cap[N], deg[:N].max())).cpu()
    e_full, e_cap, max_deg = (int(v) for v in totals)
    # 6-7. fill + per-row sort
    col_full = torch.empty(max(e_full, 1), dtype=torch.int32, device=dev)
    call("efl_radius_graph_fill", N, nbatch, cs, nx, ny, nz, r, int(loop), ptr(cell_of), ptr(cell_start),
         ptr(sorted_pts), ptr(rowptr_full), ptr(col_full), max_deg, work=N * 16.0 * 27 + e_full * 12.0)
    # 8. emit int64 edge_index with the neighbour cap applied
    edge_index = torch.empty(2, e_cap, dtype=torch.int64, device=dev)
    call("efl_radius_graph_emit", ptr(rowptr_full), ptr(rowptr_cap), ptr(col_full), N, e_cap, ptr(edge_index),
         int(flow == 'target_to_source'), work=e_cap * 20.0)
    src_row = 1 if flow == 'target_to_source' else 0
    col = col_full[:e_cap] if e_cap == e_full else edge_index[src_row].to(torch.int32)
    return RadiusCSR(edge_index, rowptr_cap, col, max_deg, e_cap != e_full)


def radius_graph(x: torch.Tensor, r: float, batch: Optional[torch.Tensor] = None, loop: bool = False,
                 max_num_neighbors: int = 32, flow: str = 'source_to_target', num_workers: int = 1) -> torch.Tensor:
    """Drop-in for torch_cluster.radius_graph: [2, E] int64 edge_index (``num_workers`` is accepted and ignored)."""
    return radius_graph_csr(x, r, batch, loop, max_num_neighbors, flow).edge_index
