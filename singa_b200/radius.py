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


def edge_csr_from_radius(rc: "RadiusCSR", num_nodes: int, register: bool = True):
    """The attention path's ``ops.EdgeCSR`` straight from a ``RadiusCSR`` (source_to_target flow): the kernel already
    emits edges grouped by destination with ascending sources, so ``perm`` is the identity and the destination sort
    of ``efl_csr_from_edge_index`` is skipped; only the source-side grouping (backward of the gather) is a stable sort
    of the source column.  ``register`` attaches it to ``rc.edge_index`` (EF_layers.register_edge_csr), so the edge
    modules pick it up when handed that tensor."""
    from . import EF_layers as ef
    from . import ops
    E = rc.edge_index.shape[1]
    dev = rc.edge_index.device
    src32 = rc.col.contiguous()
    dst32 = rc.edge_index[1].to(torch.int32)
    perm = torch.arange(E, dtype=torch.int32, device=dev)
    _, src_idx = torch.sort(src32, stable=True)
    src_ptr = torch.zeros(num_nodes + 1, dtype=torch.int32, device=dev)
    src_ptr[1:] = torch.cumsum(torch.bincount(src32, minlength=num_nodes), 0).to(torch.int32)
    csr = ops.EdgeCSR(num_nodes, num_nodes, E, perm, rc.rowptr, src32, dst32, dst32, src_ptr, src_idx.to(torch.int32))
    csr.chunks = ops.plan_chunks(rc.rowptr.cpu().numpy(), ops.MAX_CHUNK_EDGES)
    if register:
        ef.register_edge_csr(rc.edge_index, num_nodes, num_nodes, csr)
    return csr


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


def knn_graph(x: torch.Tensor, k: int, batch: Optional[torch.Tensor] = None, loop: bool = False,
              flow: str = 'source_to_target', cosine: bool = False, num_workers: int = 1) -> torch.Tensor:
    """Drop-in for torch_cluster.knn_graph (reference call sites model/CProMG.py:293, 330, with
    ``flow='target_to_source'``: row 0 = centre, row 1 = neighbour): for every point its ``k`` nearest OTHER points of
    the same batch item, grouped by centre (ascending) with neighbours in order of increasing distance, ties broken by
    the smaller index; fewer than ``k`` where the batch item is smaller.  Cell-list search (csrc/radius_graph.cu:
    knn_kernel), no CPU path.  torch-cluster is un-vendored and no reference vector exists: "parity unpinned", checked
    against the brute-force restatement oracle/cpromg_oracle.knn_graph_bruteforce (bit-exact edge list)."""
    assert flow in ('source_to_target', 'target_to_source')
    assert not cosine, "cosine distance is not used by the reference and not built"
    check_cuda(x, batch)
    assert x.dim() == 2 and x.shape[1] == 3, "knn_graph kernel handles 3-D positions"
    assert 1 <= k <= 64, "knn_graph kernel keeps up to 64 neighbours per point"
    dev = x.device
    N = x.shape[0]
    if N == 0:
        return torch.empty(2, 0, dtype=torch.int64, device=dev)
    pos = x.detach().contiguous().float()
    if batch is not None:
        b32 = batch.to(torch.int32).contiguous()
        nbatch = int(b32[-1].item()) + 1
    else:
        b32, nbatch = None, 1
    box_ws = torch.empty(nbatch * 6, dtype=torch.int32, device=dev)
    box = torch.empty(nbatch, 6, dtype=torch.float32, device=dev)
    call("efl_radius_graph_bbox", ptr(pos), ptr(b32), N, nbatch, ptr(box_ws), ptr(box))
    boxh = box.cpu()
    if b32 is not None:
        assert bool(torch.isfinite(boxh).all()), "batch vector must be sorted and cover 0..B-1 without gaps"
    size = (boxh[:, 3:] - boxh[:, :3]).clamp_min(1e-3)
    ext = size.amax(dim=0)
    # cell edge: the 27-cell stencil should hold about 1.5 k points at the mean density of the batch items
    density = N / float(size.prod(dim=1).sum())
    cs = max((1.5 * k / (27.0 * density)) ** (1.0 / 3.0), float(ext.max()) / 256.0, 1e-3)
    nx, ny, nz = (max(1, int(math.floor(float(e) / cs)) + 1) for e in ext)
    ncells = nbatch * nx * ny * nz
    origin = box[:, :3].contiguous()
    i32 = lambda n: torch.empty(n, dtype=torch.int32, device=dev)
    cell_of, cell_start, scratch = i32(N), i32(ncells + 1), i32(2 * (ncells + 1))
    sorted_pts = torch.empty(N, 4, dtype=torch.float32, device=dev)
    nbr, cnt, rowptr_full, rowptr = i32(N * k), i32(N + 1), i32(N + 1), i32(N + 1)
    tbytes = int(LIB.efl_radius_graph_temp_bytes(max(N, ncells) + 1))
    temp = torch.empty(max(tbytes, 16), dtype=torch.uint8, device=dev)
    call("efl_knn_graph", ptr(pos), ptr(b32), N, nbatch, ptr(origin), cs, nx, ny, nz, int(k), int(loop), ptr(cell_of),
         ptr(cell_start), ptr(scratch), ptr(sorted_pts), ptr(nbr), ptr(cnt), ptr(rowptr_full), ptr(rowptr), ptr(temp),
         tbytes, work=N * (16.0 * 27 + 4.0 * k))
    E = int(rowptr[N].item())
    edge_index = torch.empty(2, E, dtype=torch.int64, device=dev)
    # emit: row of the centres and row of the neighbours; target_to_source puts the centre in row 0
    call("efl_radius_graph_emit", ptr(rowptr_full), ptr(rowptr), ptr(nbr), N, E, ptr(edge_index),
         int(flow == 'target_to_source'), work=E * 20.0)
    return edge_index
