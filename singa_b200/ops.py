"""Host-side orchestration of the EF_layers CUDA kernels (autograd wrappers).

Everything numerical happens in libefl.so through the C ABI (``_lib.call``); this
file only owns buffers, the destination-sorted CSR, edge chunking and the
recompute-in-backward policy (SURVEY.md 8a "saved-for-backward note"): the edge
activations of one attention layer at config 2 (0.9 M edges) would be ~40 GB, so
forward keeps nothing per edge except logits/alpha ([E, heads]) and backward
re-runs the forward kernels chunk by chunk before the gradient kernels.
"""
from __future__ import annotations

import ctypes
import os
import weakref
from dataclasses import dataclass, field

import numpy as np
import torch

from . import so3_constants as sc
from ._lib import LIB, call, check_cuda, ptr

# GEMM operand precision of the SO(2) linear maps, set through set_precision():
#   "fp32"    SIMT fp32 (parity mode, 1e-5 / 1e-4)
#   "bf16"    tcgen05 tensor cores, bf16 operands and edge activations, fp32 accumulate (the benchmarked mode)
#   "bf16x3"  tcgen05 tensor cores on error-compensated operands: fp32 activations, every GEMM operand split into
#             bf16 (hi, lo) and a*b taken as a_hi b_hi + a_hi b_lo + a_lo b_hi in ONE launch with a 3x longer
#             reduction (csrc/split3.cu); everything else as in "fp32".  The accurate tensor-core mode of SURVEY.md 7.1.
_PRECISION = "fp32"
# edges per chunk of the edge path (whole destinations): 524 288 measured 2 % faster per step than 262 144 (fewer
# launches, longer persistent-GEMM runs; tools/ab_knobs.sh), 1 048 576 no better and twice the backward workspace
MAX_CHUNK_EDGES = int(os.environ.get("EFL_MAX_CHUNK_EDGES", "524288"))
# Saved-for-backward policy (SURVEY.md 8a note): keep the chunk's edge activations (A, extra, hidden, S2 output,
# values: ~25 KB/edge in bf16) when they fit a fraction of the free HBM, otherwise recompute them in backward.
# "auto" | "save" | "recompute"
_SAVE_POLICY = os.environ.get("EFL_SAVE_POLICY", "auto")
_SAVE_FREE_FRACTION = float(os.environ.get("EFL_SAVE_FREE_FRACTION", "0.45"))


STATS = {"save": 0, "recompute": 0}   # edge-attention forwards that kept / will recompute their edge activations


def reserve_workspace(fraction: float = 0.6, device=None) -> int:
    """Pre-size the caching allocator's pool: allocate ``fraction`` of the currently free device memory once and
    release it to the pool.  The chunk workspaces of a step (tens of GB at config 2) are then carved out of memory
    the process already owns, instead of the pool being extended while a step is running (each extension is a
    device-wide host stall of 0.5-0.9 s on the B200 boxes: tools/step_trace.py).  Returns the bytes reserved."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    free, _ = torch.cuda.mem_get_info(device)
    want = int(free * fraction) - (torch.cuda.memory_reserved(device) - torch.cuda.memory_allocated(device))
    if want <= 0:
        return 0
    block = torch.empty(want, dtype=torch.uint8, device=device)
    del block
    return want


def set_save_policy(mode: str) -> None:
    global _SAVE_POLICY
    if mode not in ("auto", "save", "recompute"):
        raise ValueError("save policy must be 'auto', 'save' or 'recompute'")
    _SAVE_POLICY = mode


_SAVED_BYTES = {}   # device index -> bytes of edge activations currently kept for backward
_MEMINFO = {}       # device index -> (timestamp, free, total) of the last cudaMemGetInfo


class _SavedChunks(list):
    """The kept workspaces of one layer; the running total follows the object's lifetime (released by backward, or
    with the autograd graph when backward never runs)."""

    def __init__(self, dev_index: int, nbytes: int):
        super().__init__()
        self.dev_index, self.nbytes = dev_index, nbytes
        _SAVED_BYTES[dev_index] = _SAVED_BYTES.get(dev_index, 0) + nbytes

    def __del__(self):
        _SAVED_BYTES[self.dev_index] = max(0, _SAVED_BYTES.get(self.dev_index, 0) - self.nbytes)


def _should_save(nbytes: int, device) -> bool:
    """Keep this layer's edge activations?  Two limits: the layer must fit a fraction of what is free NOW, and
    everything kept so far in this step plus this layer must fit the same fraction of the DEVICE (layers decide one
    after the other: without the running total, batch 64 kept saving until the last layers had no memory left)."""
    if _SAVE_POLICY != "auto":
        return _SAVE_POLICY == "save"
    # cudaMemGetInfo costs ~0.6 ms of host time per call (5 ms per step over the 24 layer passes): ask the driver at
    # most every 100 ms, the allocator's own counters (cheap) are read every time
    import time
    now = time.monotonic()
    idx = torch.device(device).index or 0
    hit = _MEMINFO.get(idx)
    if hit is None or now - hit[0] > 0.1:
        hit = _MEMINFO[idx] = (now,) + tuple(torch.cuda.mem_get_info(device))
    free, total = hit[1], hit[2]
    cached = torch.cuda.memory_reserved(device) - torch.cuda.memory_allocated(device)
    kept = _SAVED_BYTES.get(torch.device(device).index or 0, 0)
    return nbytes < _SAVE_FREE_FRACTION * (free + cached) and kept + nbytes < _SAVE_FREE_FRACTION * total


# backward of rotate-in: "nodes" = node-owned accumulation (no per-edge gradient tensor), "edges" = per-edge d x
# followed by two segment sums.  Both deterministic; chosen by measurement (tools/edge_bench.py).
ROTATE_IN_BWD = os.environ.get("EFL_ROTIN_BWD", "nodes")

def set_precision(mode: str) -> None:
    global _PRECISION
    if mode not in ("fp32", "bf16", "bf16x3"):
        raise ValueError("precision must be 'fp32', 'bf16' or 'bf16x3'")
    _PRECISION = mode


def get_precision() -> str:
    return _PRECISION


# ---------------------------------------------------------------------------
# constants on device
# ---------------------------------------------------------------------------
_CONST_CACHE: dict = {}


def jd_device(lmax: int, device) -> torch.Tensor:
    """Concatenated J_l blocks (row-major, l = 0..lmax) as fp32 on ``device``."""
    key = ("jd", lmax, str(device))
    if key not in _CONST_CACHE:
        flat = np.concatenate([sc.j_matrix(l).reshape(-1) for l in range(lmax + 1)])
        _CONST_CACHE[key] = torch.tensor(flat, dtype=torch.float32, device=device)
    return _CONST_CACHE[key]


def wigner_size(lmax: int, mmax: int) -> int:
    """Floats per edge of the Wigner store: kept rows |m'| <= mmax of each D^l, rows padded to 4 floats
    (csrc/efl_common.cuh Tab::WSZ)."""
    return sum(min(2 * l + 1, 2 * mmax + 1) * ((2 * l + 1 + 3) // 4 * 4) for l in range(lmax + 1))


def wigner_from_rotmat(rot: torch.Tensor, lmax: int, mmax: int) -> torch.Tensor:
    """[E,3,3] edge frames -> kept Wigner rows [E, WSZ] (set_wigner, EF_layers.py:485-528)."""
    check_cuda(rot)
    rot = rot.detach().to(torch.float32).contiguous()
    E = rot.shape[0]
    wig = torch.empty(E, wigner_size(lmax, mmax), dtype=torch.float32, device=rot.device)
    call("efl_wigner_from_rotmat", lmax, mmax, ptr(rot), ptr(jd_device(lmax, rot.device)), ptr(wig), E,
         work=E * (36.0 + wig.shape[1] * 4))
    return wig


# ---------------------------------------------------------------------------
# destination-sorted CSR
# ---------------------------------------------------------------------------
@dataclass
class EdgeCSR:
    num_src: int
    num_dst: int
    num_edges: int
    perm: torch.Tensor        # int32 [E]  sorted position -> original edge id (stable in edge id)
    rowptr: torch.Tensor      # int32 [num_dst + 1]
    src_sorted: torch.Tensor  # int32 [E]
    dst_sorted: torch.Tensor  # int32 [E]
    dst_orig: torch.Tensor    # int32 [E]  original order
    src_ptr: torch.Tensor     # int32 [num_src + 1]
    src_idx: torch.Tensor     # int32 [E]  sorted positions grouped by source (stable)
    chunks: list = field(default_factory=list)  # (n0, n1, e0, e1), edges of whole destinations
    aux: dict = field(default_factory=dict)     # lazily derived index vectors (int64 copies, composite groupings)
    dst_only: bool = False    # True: src_ptr / src_idx are placeholders (segment sums over destinations only)

    def index(self, name: str) -> torch.Tensor:
        """int64 index vectors for torch gathers: 'perm', 'dst_orig', 'src_orig' (original edge order) and
        'src_group' (original edge ids grouped by source, aligned with src_ptr)."""
        hit = self.aux.get(name)
        if hit is None:
            if name == "perm":
                hit = self.perm.long()
            elif name == "dst_orig":
                hit = self.dst_orig.long()
            elif name == "src_orig":
                hit = torch.empty_like(self.index("perm"))
                hit[self.index("perm")] = self.src_sorted.long()
            elif name == "src_group":
                if self.dst_only:
                    raise ValueError("this EdgeCSR carries no source-side grouping (built for segment sums only)")
                hit = self.perm.index_select(0, self.src_idx.long()).contiguous()   # int32
            else:
                raise KeyError(name)
            self.aux[name] = hit
        return hit


def _launch_csr(edge_index: torch.Tensor, num_src: int, num_dst: int):
    """Queue the CSR kernels of one edge list; returns (csr without chunk plan, device status word)."""
    check_cuda(edge_index)
    assert edge_index.dim() == 2 and edge_index.shape[0] == 2, "edge_index must be [2, E]"
    E = edge_index.shape[1]
    dev = edge_index.device
    ei = edge_index.detach().to(torch.int64)
    src, dst = ei[0].contiguous(), ei[1].contiguous()
    i32 = lambda n: torch.empty(n, dtype=torch.int32, device=dev)
    perm, src_s, dst_s, dst_o, src_idx = (i32(E) for _ in range(5))
    rowptr, src_ptr = i32(num_dst + 1), i32(num_src + 1)
    scratch = i32(3 * max(E, 1))
    tbytes = int(LIB.efl_csr_temp_bytes(E, max(num_src, num_dst)))
    temp = torch.empty(tbytes, dtype=torch.uint8, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    call("efl_csr_from_edge_index", ptr(src), ptr(dst), E, num_src, num_dst, ptr(perm), ptr(rowptr), ptr(src_s),
         ptr(dst_s), ptr(dst_o), ptr(src_ptr), ptr(src_idx), ptr(scratch), ptr(temp), tbytes, ptr(status),
         work=E * 60.0)
    return EdgeCSR(num_src, num_dst, E, perm, rowptr, src_s, dst_s, dst_o, src_ptr, src_idx), status


def build_csr_many(specs, max_chunk_edges: int = None) -> list:
    """``build_csr`` for several edge lists ``[(edge_index, num_src, num_dst), ...]`` with ONE device->host read
    (status words + row pointers of all of them) instead of two per list: the chunk plan needs the row pointers on
    the host, and every read drains the launch queue."""
    if not specs:
        return []
    launched = [_launch_csr(*spec) for spec in specs]
    host = torch.cat([t for csr, st in launched for t in (st, csr.rowptr)]).cpu().numpy()
    out, o = [], 0
    for (csr, _), (_, num_src, num_dst) in zip(launched, specs):
        status, rowptr = int(host[o]), host[o + 1:o + 2 + num_dst]
        o += 2 + num_dst
        if status != 0:
            raise IndexError(f"edge_index holds node ids outside [0, {num_src}) x [0, {num_dst})")
        csr.chunks = plan_chunks(rowptr, max_chunk_edges or MAX_CHUNK_EDGES)
        out.append(csr)
    return out


def build_csr(edge_index: torch.Tensor, num_src: int, num_dst: int, max_chunk_edges: int = None) -> EdgeCSR:
    """Stable sort of the edge list by destination (edge_index[1]) plus the source-side
    grouping used by the backward of the gather.  Row 0 = source, row 1 = target
    (aggregation is over row 1: EF_layers.py:143-145, 1180, 1197-1199)."""
    return build_csr_many([(edge_index, num_src, num_dst)], max_chunk_edges)[0]


def plan_chunks(rowptr: np.ndarray, max_edges: int) -> list:
    """Greedy split of the destination range into chunks of whole destinations with at most
    ``max_edges`` edges (a single destination with more edges gets a chunk of its own)."""
    n = len(rowptr) - 1
    chunks = []
    n0 = 0
    while n0 < n:
        limit = rowptr[n0] + max_edges
        n1 = int(np.searchsorted(rowptr, limit, side="right")) - 1
        n1 = max(n1, n0 + 1)
        n1 = min(n1, n)
        chunks.append((n0, n1, int(rowptr[n0]), int(rowptr[n1])))
        n0 = n1
    return chunks


# ---------------------------------------------------------------------------
# helpers
# ---------------------------------------------------------------------------
class _Ptrs:
    """Keeps ctypes arrays alive for the row-map arguments of the SIMT S2 kernels."""

    def __init__(self, bases, strides, first):
        n = len(bases)
        self.n = n
        self.b = (ctypes.c_void_p * n)(*bases)
        self.s = (ctypes.c_longlong * n)(*strides)
        self.f = (ctypes.c_int * (n + 1))(*first)

    def args(self):
        return (self.n, ctypes.cast(self.b, ctypes.c_void_p), ctypes.cast(self.s, ctypes.c_void_p),
                ctypes.cast(self.f, ctypes.c_void_p))


def _split3(x, rows, cols, stack: bool, second: bool):
    """fp32 [rows, cols] (any row pitch) -> bf16 (hi, hi, lo) / (hi, lo, hi) copies for the bf16x3 GEMMs."""
    out = torch.empty((3 * rows, cols) if stack else (rows, 3 * cols), dtype=torch.bfloat16, device=x.device)
    call("efl_split3", ptr(x), x.stride(0), rows, cols, ptr(out), out.stride(0), int(stack), int(second),
         work=float(rows) * cols * 10.0)
    return out


def _x3_ok(*dims) -> bool:
    return _PRECISION == "bf16x3" and all(v % 16 == 0 for v in dims)


def _weight_x3(B, N, K):
    """(hi | lo | hi) copy of a derived weight, cached on the tensor object (the _Weights cache keeps it alive)."""
    hit = getattr(B, "_efl_x3", None)
    if hit is None:
        hit = _split3(B, N, K, False, True)
        try:
            B._efl_x3 = hit
        except AttributeError:
            pass
    return hit


def _gemm_nt(epi, A, B, bias, C, ldc, C0, ldc0, nsplit, M, N, K):
    if A.dtype == torch.float32:
        if _x3_ok(N, K) and epi in (0, 1) and M > 0:   # accurate tensor-core mode: fp32 in, fp32 out, 3x K
            from . import tc_gemm
            A3, B3 = _split3(A, M, K, False, False), _weight_x3(B, N, K)
            if epi == 0:
                tc_gemm.gemm_nt(1, A3, B3, bias, C, 8, C, ldc, N, M, N, 3 * K)
            else:   # split epilogue: both parts are fp32 here, two launches over the two column ranges
                tc_gemm.gemm_nt(1, A3, B3[:nsplit], None if bias is None else bias[:nsplit], C0, 8, C0, ldc0, nsplit, M,
                                nsplit, 3 * K)
                if N > nsplit:
                    tc_gemm.gemm_nt(1, A3, B3[nsplit:], None if bias is None else bias[nsplit:], C, 8, C, ldc,
                                    N - nsplit, M, N - nsplit, 3 * K)
            return
        call("efl_gemm_nt_f32", epi, ptr(A), A.stride(0), ptr(B), B.stride(0), ptr(bias), ptr(C), ldc, ptr(C0), ldc0,
             nsplit, M, N, K, work=2.0 * M * N * K)
    else:
        from . import tc_gemm
        tc_gemm.gemm_nt(epi, A, B, bias, C, ldc, C0, ldc0, nsplit, M, N, K)


def _gemm_tn(A, B, C, accumulate, Mred, N1, N2):
    if A.dtype == torch.float32:
        if _x3_ok(N1, N2) and Mred > 0 and C.stride(0) == N2:
            from . import tc_gemm
            tc_gemm.gemm_tn(_split3(A, Mred, N1, True, False), _split3(B, Mred, N2, True, True), C, accumulate, 3 * Mred,
                            N1, N2)
            return
        splits = LIB.efl_gemm_tn_splits(Mred)
        ws = torch.empty(splits * N1 * N2, dtype=torch.float32, device=A.device)
        call("efl_gemm_tn_f32", ptr(A), A.stride(0), ptr(B), B.stride(0), ptr(C), C.stride(0), int(accumulate), ptr(ws),
             Mred, N1, N2, work=2.0 * Mred * N1 * N2)
    else:
        from . import tc_gemm
        tc_gemm.gemm_tn(A, B, C, accumulate, Mred, N1, N2)


def _colsum(X, rows, cols, out, accumulate):
    ws = torch.empty(((rows + 511) // 512 + 1) * cols, dtype=torch.float32, device=X.device)
    call("efl_colsum", int(X.dtype == torch.bfloat16), ptr(X), X.stride(0), rows, cols, ptr(out), int(accumulate),
         ptr(ws), work=float(rows) * cols * X.element_size())


# ---------------------------------------------------------------------------
# the fused edge path of SO2EquivariantGraphAttention
# ---------------------------------------------------------------------------
@dataclass
class AttnDims:
    lmax: int
    mmax: int
    C: int          # sphere channels
    H: int          # attn hidden channels
    heads: int
    alpha_ch: int
    value_ch: int

    @property
    def ms(self):  # rows per order
        return [self.lmax + 1 - m for m in range(self.mmax + 1)]

    @property
    def M(self):
        return self.ms[0] + 2 * sum(self.ms[1:])

    @property
    def first(self):  # m-primary row ranges: m0 | +1 | -1 | +2 | -2
        f = [0, self.ms[0]]
        for m in range(1, self.mmax + 1):
            f.append(f[-1] + self.ms[m])
            f.append(f[-1] + self.ms[m])
        return f


def _pick_bn(n: int) -> int:
    """Largest multiple of 16 that divides n and is <= 256 (csrc/tc_gemm.cu pick_bn): the MMA N of an n-wide output."""
    return next((bn for bn in range(256, 15, -16) if n % bn == 0), 0)


class _Layout:
    """Column offsets of the per-edge rows (include/efl.h "edge rows"): inside a row the orders follow each other,
    m = 0 | (+1 | -1) | (+2 | -2); the A row is preceded by the ones pad of the m = 0 GEMM's bias column."""

    def __init__(self, d: AttnDims):
        C2, H, VH = 2 * d.C, d.H, d.heads * d.value_ch
        k0 = d.ms[0] * C2
        # smallest pad that leaves the padded m = 0 K extent a decent MMA N for the dgrad / wgrad GEMMs
        self.pad = next((p for p in (16, 32, 48, 64) if _pick_bn(k0 + p) >= 128), 16)
        self.extra_w = d.heads * d.alpha_ch + H

        def offs(CH, lead):
            o, w = [lead], lead + d.ms[0] * CH
            for m in range(1, d.mmax + 1):
                o.append(w)
                w += 2 * d.ms[m] * CH
            return o, w
        self.offA, self.wA = offs(C2, self.pad)         # A / dA rows: [pad | m0 | m1 | m2]
        self.offH, self.wH = offs(H, 0)                 # hidden (Hb, B, dB) rows
        self.offV, self.wV = offs(VH, 0)                # value (V, dY2) rows
        self.wY1 = self.extra_w + self.wH               # dY1 rows: [d alpha_in | d gate | d hidden]
        self.k0e = self.pad + k0                        # K of the so2_conv_1 m = 0 GEMM incl. the bias column
        self.blkA = [d.ms[0] * C2] + [2 * d.ms[m] * C2 for m in range(1, d.mmax + 1)]
        self.blkH = [d.ms[0] * H] + [2 * d.ms[m] * H for m in range(1, d.mmax + 1)]
        self.blkV = [d.ms[0] * VH] + [2 * d.ms[m] * VH for m in range(1, d.mmax + 1)]


class _Workspace:
    """Chunk-sized edge buffers, one row per edge.  ``forward`` buffers: A (rotate-in output), extra (alpha/gate
    columns, fp32), Hb (so2_conv_1 hidden), B (S2 output), V (so2_conv_2 values); ``backward`` buffers: the gradients."""

    def __init__(self, d: AttnDims, Ec: int, dtype, device, forward: bool = True, backward: bool = False):
        L = self.L = _Layout(d)
        e = lambda c, dt=dtype: torch.empty(max(Ec, 1), c, dtype=dt, device=device)
        if forward:
            self.A, self.extra, self.Hb, self.B, self.V = e(L.wA), e(L.extra_w, torch.float32), e(L.wH), e(L.wH), e(L.wV)
        if backward:
            self.dY2, self.dB, self.dY1, self.dA = e(L.wV), e(L.wH), e(L.wY1), e(L.wA)

    @staticmethod
    def forward_bytes_per_edge(d: AttnDims, es: int) -> int:
        L = _Layout(d)
        return (L.wA + 2 * L.wH + L.wV) * es + L.extra_w * 4


def _blk_ptrs(buf, offs):
    """Base pointers of the m = 0 / 1 / 2 blocks of an edge-row buffer."""
    es = buf.element_size()
    return [buf.data_ptr() + o * es for o in offs]


class _Weights:
    """GEMM-ready derived copies of one SO2_Convolution's weights (never the stored parameters), cached on the
    parameters' versions: the four passes of a step and its backward share them.

    m = 0: ``w0`` [N0, K0 (+pad)]; with ``pad`` the bias is column 0 of the pad (the A operand carries a ones column).
    m > 0: SO2_m_Convolution as ONE real GEMM over [x_re | x_im] rows (EF_layers.py:721-729):
        [out_re | out_im] = [x_re | x_im] @ Wf^T,   Wf = [[Wr, -Wi], [Wi, Wr]]      (Wr / Wi = the fc weight's two halves)
    ``t`` variants are the transposes the dgrad GEMMs take as their B operand."""

    _CACHE: dict = {}
    BUILDS = 0  # derived-weight constructions so far (tests count cache misses with it)

    def __init__(self, conv_w, dtype, pad: int):
        W0, b0 = conv_w[0].detach(), conv_w[1].detach()
        if pad:
            w0 = torch.zeros(W0.shape[0], pad + W0.shape[1], dtype=torch.float32, device=W0.device)
            w0[:, 0], w0[:, pad:] = b0, W0
            self.b0 = None
        else:
            w0 = W0
            self.b0 = b0.to(torch.float32).contiguous()
        self.w0 = w0.to(dtype).contiguous()
        self.wf = []
        for w in conv_w[2:]:
            h = w.shape[0] // 2
            wr, wi = w.detach()[:h], w.detach()[h:]
            self.wf.append(torch.cat((torch.cat((wr, -wi), 1), torch.cat((wi, wr), 1)), 0).to(dtype).contiguous())
        self._t = None

    def transposed(self):
        if self._t is None:
            self._t = (self.w0.t().contiguous(), [w.t().contiguous() for w in self.wf])
        return self._t

    @classmethod
    def get(cls, conv_w, dtype, pad: int):
        # one slot per (parameter storage, dtype, pad); a new parameter version REPLACES the slot, so the cache (and
        # with it the step's memory footprint) does not grow from step to step: a footprint that creeps makes the
        # caching allocator extend its pool in the middle of a step, a 0.5-0.9 s host stall on this box
        # (tools/step_trace.py)
        #
        # A slot is only valid for the very tensor objects it was derived from: it keeps weak references to them and a
        # hit needs `ref() is w` for every one (plus equal version counters).  Address + version alone is NOT an
        # identity: when a model is freed and the next one is built, the caching allocator hands the same blocks back
        # and the same constructor sequence gives the same versions, so the new model would silently run with the old
        # model's SO(2) weights (round-2 review; tests/test_host_logic.py::test_derived_weight_cache_*).  The reference
        # reads its weights fresh on every call (EF_layers.py:807-875).  Writes through `p.data` do not bump
        # `_version`; call ops.clear_derived_caches() after such a write.
        key = tuple(w.data_ptr() for w in conv_w) + (dtype, pad)
        ver = tuple(w._version for w in conv_w)
        hit = cls._CACHE.get(key)
        if hit is None or hit[0] != ver or len(hit[2]) != len(conv_w) or any(r() is not w for r, w in zip(hit[2], conv_w)):
            if len(cls._CACHE) > 64:
                cls._CACHE.clear()
            cls.BUILDS += 1
            hit = cls._CACHE[key] = (ver, cls(conv_w, dtype, pad), tuple(weakref.ref(w) for w in conv_w))
        return hit[1]


def clear_derived_caches():
    """Drop every derived-operand cache (GEMM-ready SO(2) weights, bf16 grid matrices).  Needed only after parameters
    were overwritten behind autograd's back (``p.data.copy_``), which leaves the version counters unchanged."""
    _Weights._CACHE.clear()
    _GRID_OPERANDS.clear()


def _combine_complex_wgrad(P):
    """wgrad of the real-GEMM form back to the fc weight's layout: P = d Wf [2h, 2K] -> [d Wr ; d Wi] [2h, K]."""
    h, K = P.shape[0] // 2, P.shape[1] // 2
    return torch.cat((P[:h, :K] + P[h:, K:], P[h:, :K] - P[:h, K:]), 0)


def _chunk_forward(d: AttnDims, ws: _Workspace, w1: _Weights, w2: _Weights, x_src, x_dst, rad, csr: EdgeCSR, wig,
                   ln_w, ln_b, adot, logits, alpha, grids, chunk, out):
    n0, n1, e0, e1 = chunk
    Ec = e1 - e0
    if Ec == 0:
        if out is not None:
            out[n0:n1].zero_()
        return
    L = ws.L
    bf16 = int(ws.A.dtype == torch.bfloat16)
    pa = _blk_ptrs(ws.A, L.offA)
    call("efl_rotate_in_fwd", d.lmax, d.mmax, d.C, bf16, ptr(x_src), ptr(x_dst), ptr(csr.src_sorted),
         ptr(csr.dst_sorted), ptr(csr.perm), ptr(wig), ptr(rad), rad.shape[1], pa[0], pa[1], pa[2], L.wA, L.pad, e0, Ec,
         work=Ec * _bytes_rotate_in(d, ws.A.element_size(), wig.shape[1]))
    # so2_conv_1 (EF_layers.py:807-875): m = 0 with the bias column and the alpha/gate columns split off in fp32
    _gemm_nt(1, ws.A[:, :L.k0e], w1.w0, None, ws.Hb, L.wH, ws.extra, L.extra_w, L.extra_w, Ec,
             L.extra_w + L.blkH[0], L.k0e)
    for m in range(1, d.mmax + 1):
        _gemm_nt(0, ws.A[:, L.offA[m]:L.offA[m] + L.blkA[m]], w1.wf[m - 1], None, ws.Hb[:, L.offH[m]:], L.wH, None, 0, 0,
                 Ec, L.blkH[m], L.blkA[m])
    call("efl_alpha_logits_fwd", ptr(ws.extra), L.extra_w, d.heads, d.alpha_ch, ptr(ln_w), ptr(ln_b), ptr(adot),
         logits.data_ptr() + e0 * d.heads * 4, Ec, work=Ec * d.heads * (d.alpha_ch + 1) * 4.0)
    call("efl_segment_softmax_fwd", ptr(logits), d.heads, ptr(csr.rowptr), ptr(alpha), n0, n1)
    to_g, from_g, G = grids
    gate_ptr = ws.extra.data_ptr() + d.heads * d.alpha_ch * 4
    if bf16:
        call("efl_s2act_tc", 0, d.lmax, d.mmax, G, ptr(ws.Hb), None, ptr(ws.B), L.wH, None, gate_ptr, L.extra_w,
             ptr(to_g), ptr(from_g), Ec, work=Ec * 4.0 * G * d.M * d.H, tag="efl_s2act_tc_fwd")
    else:
        rin = _Ptrs([ws.Hb.data_ptr()], [L.wH], [0, d.M])
        rout = _Ptrs([ws.B.data_ptr()], [L.wH], [0, d.M])
        call("efl_s2act_fwd", d.M, G, d.H, 0, 0, *rin.args(), *rout.args(), gate_ptr, L.extra_w, ptr(to_g),
             ptr(from_g), Ec)
    # so2_conv_2
    _gemm_nt(0, ws.B[:, :L.blkH[0]], w2.w0, w2.b0, ws.V, L.wV, None, 0, 0, Ec, L.blkV[0], L.blkH[0])
    for m in range(1, d.mmax + 1):
        _gemm_nt(0, ws.B[:, L.offH[m]:L.offH[m] + L.blkH[m]], w2.wf[m - 1], None, ws.V[:, L.offV[m]:], L.wV, None, 0, 0,
                 Ec, L.blkV[m], L.blkH[m])
    if out is not None:
        pv = _blk_ptrs(ws.V, L.offV)
        call("efl_rotate_back_aggregate_fwd", d.lmax, d.mmax, d.heads, d.value_ch, bf16, pv[0], pv[1], pv[2], L.wV,
             ptr(alpha), ptr(csr.rowptr), ptr(csr.perm), ptr(wig), ptr(out), n0, n1, e0,
             work=Ec * _bytes_rotate_back(d, ws.V.element_size(), wig.shape[1]) + (n1 - n0) * out[0].numel() * 4)


def _bytes_rotate_in(d: AttnDims, es: int, wsz: int) -> float:
    """Algorithmic HBM bytes per edge of rotate-in: two node rows, Wigner rows, radial scale,
    indices; writes the m-primary operands."""
    L = (d.lmax + 1) ** 2
    return 2 * L * d.C * 4 + wsz * 4 + d.M * 2 * d.C * 4 + 12 + d.M * 2 * d.C * es


def _bytes_rotate_back(d: AttnDims, es: int, wsz: int) -> float:
    """Algorithmic HBM bytes per edge of rotate-back+aggregate: value rows, alpha, Wigner rows, index."""
    return d.M * d.heads * d.value_ch * es + d.heads * 4 + wsz * 4 + 4


def _bytes_rotate_back_bwd(d: AttnDims, es: int, wsz: int) -> float:
    """Backward: value rows and Wigner rows read, d value rows written (same layout), d alpha partials."""
    VH = d.heads * d.value_ch
    return 2 * d.M * VH * es + wsz * 4 + d.heads * 4 * 5 + 4


def _check_edge_inputs(who, csr, wig, lmax, mmax, rad=None, m0=None, x_src=None, x_dst=None):
    """The kernels index the Wigner store and the per-edge scalars by ORIGINAL edge id (csr.perm): a store computed
    for another edge list (a forgotten set_wigner, or pass 4 reusing pass 3's rows with |PL| != |LP|) would be read out
    of bounds.  The reference fails with a bmm shape error in that case (EF_layers.py:494-505); so do we."""
    E = csr.num_edges
    if wig.dim() != 2 or wig.shape[0] != E or wig.shape[1] != wigner_size(lmax, mmax):
        raise RuntimeError(
            f"{who}: Wigner rows {tuple(wig.shape)} do not match the edge list ({E} edges, {wigner_size(lmax, mmax)} "
            f"floats per edge for lmax={lmax}, mmax={mmax}); call SO3_Rotation.set_wigner for this edge_index first")
    for name, t in (("radial weights", rad), ("m=0 coefficients", m0)):
        if t is not None and t.shape[0] != E:
            raise RuntimeError(f"{who}: {name} have {t.shape[0]} rows for {E} edges")
    if x_src is not None and x_src.shape[0] != csr.num_src:
        raise RuntimeError(f"{who}: {x_src.shape[0]} source rows for a CSR over {csr.num_src} source nodes")
    if x_dst is not None and x_dst.shape[0] != csr.num_dst:
        raise RuntimeError(f"{who}: {x_dst.shape[0]} target rows for a CSR over {csr.num_dst} target nodes")


class EdgeAttentionFunction(torch.autograd.Function):
    """x_src, x_dst, rad, conv1 (W0,b0,W1,W2), alpha LN (w,b), alpha_dot, conv2 (W0,b0,W1,W2)
    -> aggregated messages [num_dst, L, heads*value] (input of ``proj``)."""

    @staticmethod
    def forward(ctx, x_src, x_dst, rad, c1w0, c1b0, c1w1, c1w2, ln_w, ln_b, adot, c2w0, c2b0, c2w1, c2w2, csr, wig,
                grids, dims):
        d: AttnDims = dims
        check_cuda(x_src, x_dst, rad, wig)
        assert d.mmax == 2, "the edge path is built for mmax == 2"
        _check_edge_inputs("SO2EquivariantGraphAttention", csr, wig, d.lmax, d.mmax, rad=rad,
                           x_src=x_src, x_dst=x_dst)
        x_src = x_src.contiguous()
        x_dst = x_dst.contiguous()
        rad = rad.contiguous()
        dev = x_dst.device
        dtype = torch.bfloat16 if _PRECISION == "bf16" else torch.float32
        E = csr.num_edges
        Lc = (d.lmax + 1) ** 2
        out = torch.empty(csr.num_dst, Lc, d.heads * d.value_ch, dtype=torch.float32, device=dev)
        logits = torch.empty(max(E, 1), d.heads, dtype=torch.float32, device=dev)
        alpha = torch.empty(max(E, 1), d.heads, dtype=torch.float32, device=dev)
        maxE = max((c[3] - c[2] for c in csr.chunks), default=0)
        es = 4 if dtype == torch.float32 else 2
        save = (any(ctx.needs_input_grad) and E > 0
                and _should_save(E * _Workspace.forward_bytes_per_edge(d, es), dev))
        STATS["save" if save else "recompute"] += 1
        ws = None if save else _Workspace(d, maxE, dtype, dev)
        pad = _Layout(d).pad
        w1 = _Weights.get([c1w0, c1b0, c1w1, c1w2], dtype, pad)
        w2 = _Weights.get([c2w0, c2b0, c2w1, c2w2], dtype, 0)
        lw, lb, ad = ln_w.detach().contiguous(), ln_b.detach().contiguous(), adot.detach().contiguous()
        saved = _SavedChunks(dev.index or 0, E * _Workspace.forward_bytes_per_edge(d, es)) if save else None
        for chunk in csr.chunks:
            if save:
                ws = _Workspace(d, chunk[3] - chunk[2], dtype, dev)
                saved.append(ws)
            _chunk_forward(d, ws, w1, w2, x_src, x_dst, rad, csr, wig, lw, lb, ad, logits, alpha, grids, chunk, out)
        ctx.save_for_backward(x_src, x_dst, rad, c1w0, c1b0, c1w1, c1w2, ln_w, ln_b, adot, c2w0, c2b0, c2w1, c2w2,
                              logits, alpha)
        ctx.misc = (csr, wig, grids, d, dtype)
        ctx.saved_ws = saved if save else None
        return out

    @staticmethod
    def backward(ctx, dout):
        (x_src, x_dst, rad, c1w0, c1b0, c1w1, c1w2, ln_w, ln_b, adot, c2w0, c2b0, c2w1, c2w2, logits,
         alpha) = ctx.saved_tensors
        csr, wig, grids, d, dtype = ctx.misc
        dev = x_dst.device
        dout = dout.contiguous().to(torch.float32)
        E = csr.num_edges
        Lc = (d.lmax + 1) ** 2
        C2 = 2 * d.C
        bf16 = int(dtype == torch.bfloat16)
        maxE = max((c[3] - c[2] for c in csr.chunks), default=0)
        saved_ws = ctx.saved_ws
        ctx.saved_ws = None
        wsr = _Workspace(d, maxE, dtype, dev, forward=saved_ws is None, backward=True)
        L = wsr.L
        w1 = _Weights.get([c1w0, c1b0, c1w1, c1w2], dtype, L.pad)
        w2 = _Weights.get([c2w0, c2b0, c2w1, c2w2], dtype, 0)
        w1_0t, w1_ft = w1.transposed()
        w2_0t, w2_ft = w2.transposed()
        lw, lb, ad = ln_w.detach().contiguous(), ln_b.detach().contiguous(), adot.detach().contiguous()
        f32 = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)
        n01 = L.extra_w + L.blkH[0]
        g_c1w0e = f32(n01, L.k0e)                                         # [d W0 | pad | bias column]
        P1 = [f32(L.blkH[m], L.blkA[m]) for m in range(1, d.mmax + 1)]    # d Wf of so2_conv_1, m > 0
        g_c2w0, g_c2b0 = f32(*c2w0.shape), f32(*c2b0.shape)
        P2 = [f32(L.blkV[m], L.blkH[m]) for m in range(1, d.mmax + 1)]
        drad = torch.empty_like(rad)   # every column of every edge is written by the rotate-in backward
        dx_dst = torch.zeros_like(x_dst)
        dx_src = torch.zeros_like(x_src)
        dxe = None
        nparts = LIB.efl_rotate_back_parts(d.lmax, d.mmax)
        dalpha = torch.empty(nparts, max(E, 1), d.heads, dtype=torch.float32, device=dev)
        dlogits = torch.empty(max(E, 1), d.heads, dtype=torch.float32, device=dev)
        nblk = 148 * 4
        stride = d.heads * d.alpha_ch + 2 * d.alpha_ch
        part_all = f32(stride)
        part = torch.empty(nblk, stride, dtype=torch.float32, device=dev)
        to_g, from_g, G = grids
        ha = d.heads * d.alpha_ch
        first = True
        for ci, chunk in enumerate(csr.chunks):
            n0, n1, e0, e1 = chunk
            Ec = e1 - e0
            if Ec == 0:
                continue
            if saved_ws is None:
                ws = wsr
                _chunk_forward(d, ws, w1, w2, x_src, x_dst, rad, csr, wig, lw, lb, ad, logits, alpha, grids, chunk,
                               None)
            else:
                ws = saved_ws[ci]
                saved_ws[ci] = None     # release the chunk's activations as soon as its gradients are done
                ws.dY2, ws.dB, ws.dY1, ws.dA = wsr.dY2, wsr.dB, wsr.dY1, wsr.dA
            es = ws.dY1.element_size()
            # ---- rotate-back + aggregate backward -> dV (edge rows), d alpha
            pv, pdy = _blk_ptrs(ws.V, L.offV), _blk_ptrs(ws.dY2, L.offV)
            call("efl_rotate_back_aggregate_bwd", d.lmax, d.mmax, d.heads, d.value_ch, bf16, pv[0], pv[1], pv[2], L.wV,
                 pdy[0], pdy[1], pdy[2], L.wV, ptr(alpha), ptr(csr.rowptr), ptr(csr.perm), ptr(wig), ptr(dout),
                 ptr(dalpha), dalpha.stride(0), n0, n1, e0,
                 work=Ec * _bytes_rotate_back_bwd(d, ws.V.element_size(), wig.shape[1]))
            # ---- so2_conv_2 backward (plain GEMMs over M = edges: wgrad TN, dgrad NT)
            acc = not first
            dy20 = ws.dY2[:Ec, :L.blkV[0]]
            _gemm_tn(dy20, ws.B[:, :L.blkH[0]], g_c2w0, acc, Ec, L.blkV[0], L.blkH[0])
            _colsum(dy20, Ec, L.blkV[0], g_c2b0, acc)
            _gemm_nt(0, dy20, w2_0t, None, ws.dB, L.wH, None, 0, 0, Ec, L.blkH[0], L.blkV[0])
            for m in range(1, d.mmax + 1):
                dy2m = ws.dY2[:Ec, L.offV[m]:L.offV[m] + L.blkV[m]]
                _gemm_tn(dy2m, ws.B[:, L.offH[m]:L.offH[m] + L.blkH[m]], P2[m - 1], acc, Ec, L.blkV[m], L.blkH[m])
                _gemm_nt(0, dy2m, w2_ft[m - 1], None, ws.dB[:, L.offH[m]:], L.wH, None, 0, 0, Ec, L.blkH[m], L.blkV[m])
            # ---- S2 activation backward -> dY1 row = [d alpha_in | d gate | d hidden rows]
            gate_ptr = ws.extra.data_ptr() + ha * 4
            dgate_ptr = ws.dY1.data_ptr() + ha * es
            dhid_ptr = ws.dY1.data_ptr() + L.extra_w * es
            if bf16:
                call("efl_s2act_tc", 1, d.lmax, d.mmax, G, ptr(ws.Hb), ptr(ws.dB), dhid_ptr, L.wY1, dgate_ptr, gate_ptr,
                     L.extra_w, ptr(to_g), ptr(from_g), Ec, work=Ec * 6.0 * G * d.M * d.H, tag="efl_s2act_tc_bwd")
            else:
                rin = _Ptrs([ws.Hb.data_ptr()], [L.wH], [0, d.M])
                rdin = _Ptrs([ws.dB.data_ptr()], [L.wH], [0, d.M])
                rdout = _Ptrs([dhid_ptr], [L.wY1], [0, d.M])
                call("efl_s2act_bwd", d.M, G, d.H, 0, 0, *rin.args(), *rdin.args(), *rdout.args(), 0, gate_ptr,
                     L.extra_w, dgate_ptr, L.wY1, ptr(to_g), ptr(from_g), Ec)
            # ---- softmax + alpha-logit backward -> first heads*alpha columns of the dY1 rows
            call("efl_segment_softmax_bwd", ptr(alpha), ptr(dalpha), nparts, dalpha.stride(0), d.heads, ptr(csr.rowptr),
                 ptr(dlogits), n0, n1)
            call("efl_alpha_logits_bwd", ptr(ws.extra), L.extra_w, d.heads, d.alpha_ch, ptr(lw), ptr(lb), ptr(ad),
                 dlogits.data_ptr() + e0 * d.heads * 4, ptr(ws.dY1), L.wY1, bf16, ptr(part), nblk, Ec,
                 work=Ec * d.heads * ((d.alpha_ch + 1) * 4.0 + d.alpha_ch * es))
            part_all += part.sum(dim=0)
            # ---- so2_conv_1 backward; the bias gradient is the ones-column row of the m = 0 wgrad product
            dy10 = ws.dY1[:Ec, :n01]
            _gemm_tn(dy10, ws.A[:, :L.k0e], g_c1w0e, acc, Ec, n01, L.k0e)
            _gemm_nt(0, dy10, w1_0t, None, ws.dA, L.wA, None, 0, 0, Ec, L.k0e, n01)
            for m in range(1, d.mmax + 1):
                o = L.extra_w + L.offH[m]
                dy1m = ws.dY1[:Ec, o:o + L.blkH[m]]
                _gemm_tn(dy1m, ws.A[:, L.offA[m]:L.offA[m] + L.blkA[m]], P1[m - 1], acc, Ec, L.blkH[m], L.blkA[m])
                _gemm_nt(0, dy1m, w1_ft[m - 1], None, ws.dA[:, L.offA[m]:], L.wA, None, 0, 0, Ec, L.blkA[m], L.blkH[m])
            # ---- rotate-in backward -> d rad and d x
            pda = _blk_ptrs(ws.dA, L.offA)
            if ROTATE_IN_BWD == "nodes":  # node-owned: no per-edge gradient tensor
                wrk = Ec * (_bytes_rotate_in(d, ws.dA.element_size(), wig.shape[1]) + d.M * C2 * 4) / 2
                call("efl_rotate_in_bwd_nodes", d.lmax, d.mmax, d.C, bf16, 0, ptr(x_dst), ptr(csr.rowptr), None,
                     ptr(csr.perm), ptr(wig), ptr(rad), rad.shape[1], pda[0], pda[1], pda[2], L.wA,
                     ptr(drad), ptr(dx_dst), e0, Ec, n0, n1, 0, work=wrk, tag="efl_rotate_in_bwd_dst")
                call("efl_rotate_in_bwd_nodes", d.lmax, d.mmax, d.C, bf16, 1, ptr(x_src), ptr(csr.src_ptr),
                     ptr(csr.src_idx), ptr(csr.perm), ptr(wig), ptr(rad), rad.shape[1], pda[0], pda[1], pda[2], L.wA,
                     ptr(drad), ptr(dx_src), e0, Ec, 0, csr.num_src, 1, work=wrk, tag="efl_rotate_in_bwd_src")
            else:  # edge-owned: per-edge d x [E, L, 2C] + two deterministic segment sums after the loop
                if dxe is None:
                    dxe = torch.empty(max(E, 1), Lc, C2, dtype=torch.float32, device=dev)
                call("efl_rotate_in_bwd", d.lmax, d.mmax, d.C, bf16, ptr(x_src), ptr(x_dst), ptr(csr.src_sorted),
                     ptr(csr.dst_sorted), ptr(csr.perm), ptr(wig), ptr(rad), rad.shape[1], pda[0], pda[1], pda[2],
                     L.wA, ptr(drad), ptr(dxe), e0, Ec,
                     work=Ec * (_bytes_rotate_in(d, ws.dA.element_size(), wig.shape[1]) + d.M * C2 * 4 + Lc * C2 * 4))
            first = False
        if dxe is not None:
            call("efl_segment_sum_rows", ptr(dxe), Lc, C2, d.C, d.C, ptr(csr.rowptr), None, ptr(dx_dst), csr.num_dst, 0,
                 work=(E + csr.num_dst) * Lc * d.C * 4.0)
            call("efl_segment_sum_rows", ptr(dxe), Lc, C2, 0, d.C, ptr(csr.src_ptr), ptr(csr.src_idx), ptr(dx_src),
                 csr.num_src, 0, work=(E + csr.num_src) * Lc * d.C * 4.0 + E * 4.0)
        g_adot = part_all[:ha].view(d.heads, d.alpha_ch)
        g_lnw = part_all[ha:ha + d.alpha_ch]
        g_lnb = part_all[ha + d.alpha_ch:]
        return (dx_src, dx_dst, drad, g_c1w0e[:, L.pad:].contiguous(), g_c1w0e[:, 0].contiguous(),
                _combine_complex_wgrad(P1[0]), _combine_complex_wgrad(P1[1]), g_lnw, g_lnb, g_adot, g_c2w0, g_c2b0,
                _combine_complex_wgrad(P2[0]), _combine_complex_wgrad(P2[1]), None, None, None, None)


# ---------------------------------------------------------------------------
# EdgeDegreeEmbedding tail
# ---------------------------------------------------------------------------
class EdgeDegreeFunction(torch.autograd.Function):
    """m0 [E, (lmax+1)*C] (radial MLP output, original edge order) -> [num_dst, L, C]."""

    @staticmethod
    def forward(ctx, m0, csr, wig, lmax, mmax, C, rescale_factor):
        check_cuda(m0, wig)
        _check_edge_inputs("EdgeDegreeEmbedding", csr, wig, lmax, mmax, m0=m0)
        m0 = m0.contiguous()
        L = (lmax + 1) ** 2
        out = torch.empty(csr.num_dst, L, C, dtype=torch.float32, device=m0.device)
        call("efl_edge_degree_fwd", lmax, mmax, C, ptr(m0), ptr(csr.rowptr), ptr(csr.perm), ptr(wig),
             float(rescale_factor), ptr(out), csr.num_dst)
        ctx.misc = (csr, wig, lmax, mmax, C, rescale_factor, m0.shape)
        return out

    @staticmethod
    def backward(ctx, dout):
        csr, wig, lmax, mmax, C, rf, shape = ctx.misc
        dout = dout.contiguous()
        dm0 = torch.empty(shape, dtype=torch.float32, device=dout.device)
        call("efl_edge_degree_bwd", lmax, mmax, C, ptr(dout), ptr(csr.dst_orig), ptr(wig), float(rf), ptr(dm0),
             csr.num_edges)
        return dm0, None, None, None, None, None, None


# ---------------------------------------------------------------------------
# node-level separable S2 activation (FeedForwardNetwork)
# ---------------------------------------------------------------------------
# Node-level S2 activation on grids the fused tcgen05 kernel is not built for (lmax 8: 81 coefficients, 342 grid
# points): the two projections are plain tensor-core GEMMs over all (node, channel) rows with the small grid matrix as
# the B operand, and the [N*CH, G] grid values make one bf16 round trip through HBM (2.9 GB per layer at 8 000 nodes,
# ~1 ms) instead of 90 ms in the fp32 SIMT kernel.  EFL_NODE_S2_GEMM=0 falls back to the SIMT kernel.
_NODE_S2_GEMM = os.environ.get("EFL_NODE_S2_GEMM", "1") != "0"
_GRID_OPERANDS: dict = {}


def _grid_operands(to_g, from_g):
    """bf16 B operands of the four GEMMs, zero-padded to multiples of 16: T [GP, LP], T^T [LP, GP], F [GP, LP],
    F^T [LP, GP] (to_g / from_g are [G, L] fp32)."""
    key = (to_g.data_ptr(), to_g._version, from_g.data_ptr(), from_g._version, tuple(to_g.shape))
    hit = _GRID_OPERANDS.get(key)
    if hit is not None and (hit[4]() is not to_g or hit[5]() is not from_g):
        hit = None  # same address and version, another tensor: a freed model's block handed out again (see _Weights.get)
    if hit is None:
        G, L = to_g.shape
        GP, LP = _pad(G, 16), _pad(L, 16)

        def padded(m, r, c):
            out = torch.zeros(r, c, dtype=torch.bfloat16, device=m.device)
            out[:m.shape[0], :m.shape[1]] = m
            return out
        hit = (padded(to_g, GP, LP), padded(to_g.t(), LP, GP), padded(from_g, GP, LP), padded(from_g.t(), LP, GP),
               weakref.ref(to_g), weakref.ref(from_g))
        if len(_GRID_OPERANDS) > 32:
            _GRID_OPERANDS.clear()
        _GRID_OPERANDS[key] = hit
    return hit


def _rows_from_nodes(x, LP, zero_row0=False):
    """[N, L, CH] fp32 -> bf16 [N*CH, LP] (coefficients contiguous per (node, channel) row, zero padded)."""
    N, L, CH = x.shape
    xt = torch.empty(N * CH, LP, dtype=torch.bfloat16, device=x.device)
    call("efl_node_rows_pack", ptr(x), N, L, CH, LP, int(zero_row0), ptr(xt), work=N * L * CH * 4.0 + N * CH * LP * 2.0)
    return xt


def _nodes_from_rows(yt, N, L, CH, gate=None):
    """bf16 [N*CH, LP] -> fp32 [N, L, CH]; with ``gate`` the l = 0 row is SiLU(gate) (EF_layers.py:1765-1773)."""
    y = torch.empty(N, L, CH, dtype=torch.float32, device=yt.device)
    call("efl_node_rows_unpack", ptr(yt), N, L, CH, yt.shape[1], ptr(gate), ptr(y),
         work=N * L * CH * 4.0 + N * CH * yt.shape[1] * 2.0)
    return y


def _bf16_gemm(A, B, M, N, K):
    out = torch.empty(M, N, dtype=torch.bfloat16, device=A.device)
    _gemm_nt(0, A, B, None, out, N, None, 0, 0, M, N, K)
    return out


def _node_s2_gemm_forward(gate, x, to_g, from_g):
    N, L, CH = x.shape
    Tb, _, _, FbT = _grid_operands(to_g, from_g)[:4]
    GP, LP = Tb.shape
    R = N * CH
    xt = _rows_from_nodes(x.contiguous().float(), LP)
    grid = torch.empty(R, GP, dtype=torch.bfloat16, device=x.device)
    _gemm_nt(3, xt, Tb, None, grid, GP, None, 0, 0, R, GP, LP)   # to_grid with SiLU in the GEMM epilogue
    yt = _bf16_gemm(grid, FbT, R, LP, GP)                        # from_grid
    del grid
    return _nodes_from_rows(yt, N, L, CH, gate), xt


def _node_s2_gemm_backward(gate, xt, to_g, from_g, dy, shape):
    N, L, CH = shape
    Tb, TbT, Fb, _ = _grid_operands(to_g, from_g)[:4]
    GP, LP = Tb.shape
    R = N * CH
    dy = dy.contiguous().float()
    dgate = torch.ops.aten.silu_backward(dy[:, 0, :].contiguous(), gate)
    dyt = _rows_from_nodes(dy, LP, zero_row0=True)               # the l = 0 output row does not come from the grid
    grid = _bf16_gemm(xt, Tb, R, GP, LP)                         # recomputed, not saved
    dact = _bf16_gemm(dyt, Fb, R, GP, LP)
    call("efl_bf16_silu", ptr(dact), ptr(grid), dact.numel(), 1, work=dact.numel() * 6.0)   # d grid, in place
    del grid
    dxt = _bf16_gemm(dact, TbT, R, LP, GP)
    return dgate, _nodes_from_rows(dxt, N, L, CH)


class NodeS2ActFunction(torch.autograd.Function):
    """(gate [N, CH], x [N, L, CH]) -> SeparableS2Activation output [N, L, CH] on the (lmax, lmax) grid.
    bf16 mode: tcgen05 kernel on bf16 copies (fp32 accumulation); fp32 mode: SIMT kernel."""

    @staticmethod
    def forward(ctx, gate, x, to_g, from_g):
        check_cuda(gate, x)
        gate = gate.contiguous().float()
        N, Lc, CH = x.shape
        G = to_g.shape[0]
        tc = _PRECISION == "bf16" and CH % 128 == 0 and Lc <= 64 and G <= 224
        ctx.tc = tc
        ctx.gemm = (not tc) and _PRECISION == "bf16" and CH % 8 == 0 and N > 0 and _NODE_S2_GEMM
        if ctx.gemm:
            y, xt = _node_s2_gemm_forward(gate, x, to_g, from_g)
            ctx.save_for_backward(gate, xt, to_g, from_g)
            ctx.shape = (N, Lc, CH)
            return y
        if tc:
            xb = x.to(torch.bfloat16).contiguous()
            yb = torch.empty_like(xb)
            call("efl_s2act_tc_nodes", 0, Lc, G, CH, ptr(xb), None, ptr(yb), None, ptr(gate), CH, ptr(to_g), ptr(from_g),
                 N, work=N * 4.0 * G * Lc * CH, tag="efl_s2act_tc_nodes_fwd")
            ctx.save_for_backward(gate, xb, to_g, from_g)
            return yb.float()
        x = x.contiguous()
        y = torch.empty_like(x)
        pin = _Ptrs([x.data_ptr()], [Lc * CH], [0, Lc])
        pout = _Ptrs([y.data_ptr()], [Lc * CH], [0, Lc])
        call("efl_s2act_fwd", Lc, G, CH, 0, 0, *pin.args(), *pout.args(), ptr(gate), CH, ptr(to_g), ptr(from_g), N)
        ctx.save_for_backward(gate, x, to_g, from_g)
        return y

    @staticmethod
    def backward(ctx, dy):
        gate, x, to_g, from_g = ctx.saved_tensors
        if ctx.gemm:
            dgate, dx = _node_s2_gemm_backward(gate, x, to_g, from_g, dy, ctx.shape)
            return dgate, dx, None, None
        N, Lc, CH = x.shape
        G = to_g.shape[0]
        if ctx.tc:
            dyb = dy.to(torch.bfloat16).contiguous()
            dxb = torch.empty_like(x)
            dgb = torch.empty(N, CH, dtype=torch.bfloat16, device=x.device)
            call("efl_s2act_tc_nodes", 1, Lc, G, CH, ptr(x), ptr(dyb), ptr(dxb), ptr(dgb), ptr(gate), CH, ptr(to_g),
                 ptr(from_g), N, work=N * 6.0 * G * Lc * CH, tag="efl_s2act_tc_nodes_bwd")
            return dgb.float(), dxb.float(), None, None
        dy = dy.contiguous()
        dx = torch.empty_like(x)
        dgate = torch.empty_like(gate)
        pin = _Ptrs([x.data_ptr()], [Lc * CH], [0, Lc])
        pdin = _Ptrs([dy.data_ptr()], [Lc * CH], [0, Lc])
        pdout = _Ptrs([dx.data_ptr()], [Lc * CH], [0, Lc])
        call("efl_s2act_bwd", Lc, G, CH, 0, 0, *pin.args(), *pdin.args(), *pdout.args(), 0, ptr(gate), CH, ptr(dgate),
             CH, ptr(to_g), ptr(from_g), N)
        return dgate, dx, None, None


# ---------------------------------------------------------------------------
# element-embedding lookup of the edge scalars
# ---------------------------------------------------------------------------
class EmbeddingLookupFunction(torch.autograd.Function):
    """weight[ids] with a deterministic backward (csrc/attention.cu: efl_embedding_grad) instead of the float atomics
    of torch's embedding_dense_backward.  weight [rows, C] with C | 256 and a small table (source/target element
    embeddings, Embedding.py:97-99)."""

    @staticmethod
    def eligible(weight) -> bool:
        rows, C = weight.shape
        return weight.is_cuda and 256 % C == 0 and rows * 256 * 4 <= 200 * 1024

    @staticmethod
    def forward(ctx, weight, ids):
        check_cuda(weight, ids)
        ids = ids.contiguous().long()
        ctx.save_for_backward(ids)
        ctx.shape = tuple(weight.shape)
        return weight.detach().index_select(0, ids)

    @staticmethod
    def backward(ctx, dout):
        (ids,) = ctx.saved_tensors
        rows, C = ctx.shape
        dout = dout if dout.stride(-1) == 1 and dout.dtype == torch.float32 else dout.contiguous().float()
        E = ids.shape[0]
        dW = torch.empty(rows, C, dtype=torch.float32, device=dout.device)
        nblk = max(1, min(2 * 148, (E + 1023) // 1024))
        ws = torch.empty(nblk * rows * C, dtype=torch.float32, device=dout.device)
        call("efl_embedding_grad", ptr(ids), ptr(dout), dout.stride(0), E, rows, C, ptr(dW), ptr(ws), nblk,
             work=E * (C * 4.0 + 8.0))
        return dW, None


# ---------------------------------------------------------------------------
# RadialFunction
# ---------------------------------------------------------------------------
class RadialMLPFunction(torch.autograd.Function):
    """RadialFunction (EF_layers.py:1634-1657) with hidden width 16 as one kernel forward and one backward
    (csrc/radial.cu); only x is kept for backward.  params: W1,b1,g1,be1,W2,b2,g2,be2,W3,b3."""

    @staticmethod
    def eligible(net) -> bool:
        if len(net) != 7:
            return False
        l1, n1, _, l2, n2, _, l3 = net
        return (isinstance(l1, torch.nn.Linear) and isinstance(l2, torch.nn.Linear) and isinstance(l3, torch.nn.Linear)
                and isinstance(n1, torch.nn.LayerNorm) and isinstance(n2, torch.nn.LayerNorm)
                and l1.out_features == 16 and l2.in_features == 16 and l2.out_features == 16 and l3.in_features == 16
                and l1.in_features in (16, 32, 48, 64) and l3.out_features % 4 == 0 and l3.out_features <= 768
                and abs(n1.eps - 1e-5) < 1e-12 and abs(n2.eps - 1e-5) < 1e-12)

    @staticmethod
    def forward(ctx, x, *params):
        check_cuda(x, *params)
        x = x.contiguous().float()
        ps = [q.detach().contiguous().float() for q in params]
        E, K0 = x.shape
        NO = ps[8].shape[0]
        ctx.tc = _PRECISION == "bf16" and NO % 16 == 0 and E > 0
        if ctx.tc:
            # hidden layers SIMT -> a2 bf16; last layer = K=16 tcgen05 GEMM with the bias in the epilogue
            a2 = torch.empty(E, 16, dtype=torch.bfloat16, device=x.device)
            call("efl_radial_hidden_fwd", ptr(x), K0, *[ptr(q) for q in ps[:8]], ptr(a2), E, work=E * (4.0 * K0 + 32))
            w3b = ps[8].to(torch.bfloat16)
            out = torch.empty(E, NO, dtype=torch.float32, device=x.device)
            _gemm_nt(1, a2, w3b, ps[9], out, 8, out, NO, NO, E, NO, 16)
            ctx.save_for_backward(x, a2, w3b, *ps[:8])
            return out
        out = torch.empty(E, NO, dtype=torch.float32, device=x.device)
        call("efl_radial_mlp_fwd", ptr(x), K0, *[ptr(q) for q in ps], NO, ptr(out), E, work=E * 4.0 * (K0 + NO))
        ctx.save_for_backward(x, *ps)
        return out

    @staticmethod
    def _split(t, K0, NO):
        o = 0

        def take(*shape):
            nonlocal o
            n = int(np.prod(shape))
            v = t[o:o + n].view(*shape)
            o += n
            return v

        if NO:
            dW3, db3 = take(NO, 16), take(NO)
        else:
            dW3 = db3 = None
        dW2, db2, dg2, dbe2 = take(16, 16), take(16), take(16), take(16)
        dW1, db1, dg1, dbe1 = take(16, K0), take(16), take(16), take(16)
        return dW1, db1, dg1, dbe1, dW2, db2, dg2, dbe2, dW3, db3

    @staticmethod
    def backward(ctx, dout):
        if ctx.tc:
            x, a2, w3b, *ps = ctx.saved_tensors
            E, K0 = x.shape
            NO = w3b.shape[0]
            dob = dout.contiguous().to(torch.bfloat16)
            da2 = _f32_from_gemm(dob, w3b.t().contiguous(), E, 16, NO)
            dW3 = torch.empty(NO, 16, dtype=torch.float32, device=x.device)
            _gemm_tn(dob, a2, dW3, False, E, NO, 16)
            db3 = torch.empty(NO, dtype=torch.float32, device=x.device)
            _colsum(dob, E, NO, db3, False)
            dx = torch.empty_like(x)
            psz = LIB.efl_radial_hidden_part_size(K0)
            nblk = max(1, min(4 * 148, (E + 31) // 32))
            part = torch.empty(nblk, psz, dtype=torch.float32, device=x.device)
            call("efl_radial_hidden_bwd", ptr(x), K0, *[ptr(q) for q in ps], ptr(da2), ptr(dx), ptr(part), nblk, E,
                 work=E * (8.0 * K0 + 64))
            g = RadialMLPFunction._split(part.sum(dim=0), K0, 0)
            return (dx, *g[:8], dW3, db3)
        x, *ps = ctx.saved_tensors
        E, K0 = x.shape
        NO = ps[8].shape[0]
        dout = dout.contiguous().float()
        dx = torch.empty_like(x)
        psz = LIB.efl_radial_mlp_part_size(K0, NO)
        nblk = max(1, min(2 * 148, (E + 31) // 32))
        part = torch.empty(nblk, psz, dtype=torch.float32, device=x.device)
        call("efl_radial_mlp_bwd", ptr(x), K0, *[ptr(q) for q in ps], NO, ptr(dout), ptr(dx), ptr(part), nblk, E,
             work=E * 4.0 * (2 * K0 + NO))
        return (dx, *RadialMLPFunction._split(part.sum(dim=0), K0, NO))


# ---------------------------------------------------------------------------
# node-level SO3_LinearV2 and RMS norm
# ---------------------------------------------------------------------------
def _pad(n: int, m: int) -> int:
    return (n + m - 1) // m * m


def _so3_expand(x, lmax: int, KP: int, expand: bool, bias_col: int = -1):
    """x [N, L, C] fp32 -> bf16 GEMM operand [N*L, KP] (csrc/so3lin.cu): degree-block expanded, or a plain cast."""
    N, L, C = x.shape
    xe = torch.empty(N * L, KP, dtype=torch.bfloat16, device=x.device)
    call("efl_so3_expand", ptr(x), ptr(xe), N * L, lmax, C, KP, int(expand), bias_col,
         work=N * L * (C * 4.0 + KP * 2.0))
    return xe


def _so3_select(ye, bias, N: int, lmax: int, C: int):
    """ye [N*L, >= (lmax+1)*C] fp32 -> y [N, L, C]: column block l(i) of each row (+ bias on i = 0)."""
    L = (lmax + 1) ** 2
    y = torch.empty(N, L, C, dtype=torch.float32, device=ye.device)
    call("efl_so3_select", ptr(ye), ye.stride(0), ptr(bias), ptr(y), N * L, lmax, C, work=N * L * C * 8.0)
    return y


def _so3_w_expand_in(w, bias, KP: int):
    """[lmax+1, Cout, Cin] (+ bias) -> B operand [Cout, KP] bf16 with the W[l] side by side and the bias column."""
    nl, Cout, Cin = w.shape
    we = torch.zeros(Cout, KP, dtype=torch.float32, device=w.device)
    we[:, :nl * Cin] = w.permute(1, 0, 2).reshape(Cout, nl * Cin)
    if bias is not None:
        we[:, nl * Cin] = bias
    return we.to(torch.bfloat16)


def _so3_w_stack(w, NP: int):
    """[lmax+1, Cout, Cin] -> B operand [NP, Cin] bf16, the W[l] stacked (zero rows up to NP)."""
    nl, Cout, Cin = w.shape
    wa = torch.zeros(NP, Cin, dtype=torch.float32, device=w.device)
    wa[:nl * Cout] = w.reshape(nl * Cout, Cin)
    return wa.to(torch.bfloat16)


def _f32_from_gemm(A, B, M, N, K):
    """fp32 [M, N] = A[M, K] B[N, K]^T on the tensor cores (split epilogue with every column on the fp32 side)."""
    out = torch.empty(M, N, dtype=torch.float32, device=A.device)
    _gemm_nt(1, A, B, None, out, 8, out, N, N, M, N, K)
    return out


def _so3_tc_ok(Cin: int, Cout: int) -> bool:
    return _PRECISION == "bf16" and Cin % 16 == 0 and Cout % 16 == 0


class SO3LinearFunction(torch.autograd.Function):
    """x [N, L, Cin], weight [lmax+1, Cout, Cin], bias [Cout] -> [N, L, Cout] (EF_layers.py:655-671).
    bf16 mode: one tcgen05 GEMM with the degree selection folded into an operand (csrc/so3lin.cu);
    fp32 mode: batched SIMT GEMM."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        check_cuda(x, weight, bias)
        x = x.contiguous().float()
        w = weight.detach().contiguous().float()
        N, L, Cin = x.shape
        lmax, Cout = w.shape[0] - 1, w.shape[1]
        ctx.tc = _so3_tc_ok(Cin, Cout) and N > 0
        if ctx.tc:
            nl, R = lmax + 1, N * L
            b = bias.detach().float()
            if Cin < Cout:  # expand-in
                KP = _pad(nl * Cin + 1, 64)
                we = _so3_w_expand_in(w, b, KP)
                xe = _so3_expand(x, lmax, KP, True, nl * Cin)
                out = _f32_from_gemm(xe, we, R, Cout, KP).view(N, L, Cout)
                ctx.save_for_backward(x, we)
            else:  # select-out
                NP = _pad(nl * Cout, 16)
                wa = _so3_w_stack(w, NP)
                xb = _so3_expand(x, lmax, Cin, False)
                out = _so3_select(_f32_from_gemm(xb, wa, R, NP, Cin), b.contiguous(), N, lmax, Cout)
                ctx.save_for_backward(xb, wa)
            ctx.dims = (N, L, Cin, Cout, lmax)
            return out
        out = torch.empty(N, L, Cout, dtype=torch.float32, device=x.device)
        call("efl_so3_linear_fwd", ptr(x), ptr(w), ptr(bias.detach().contiguous().float()), ptr(out), N, lmax, Cin, Cout,
             work=2.0 * N * L * Cin * Cout)
        ctx.save_for_backward(x, w)
        return out

    @staticmethod
    def backward(ctx, dout):
        dout = dout.contiguous().float()
        if ctx.tc:
            return SO3LinearFunction._backward_tc(ctx, dout)
        x, w = ctx.saved_tensors
        N, L, Cin = x.shape
        lmax, Cout = w.shape[0] - 1, w.shape[1]
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            wt = w.transpose(1, 2).contiguous()
            dx = torch.empty_like(x)
            call("efl_so3_linear_bwd_data", ptr(dout), ptr(wt), ptr(dx), N, lmax, Cin, Cout, work=2.0 * N * L * Cin * Cout)
        if ctx.needs_input_grad[1]:
            dw = torch.empty_like(w)
            ws = torch.empty(L * Cout * Cin, dtype=torch.float32, device=x.device)
            call("efl_so3_linear_bwd_weight", ptr(dout), ptr(x), ptr(dw), ptr(ws), N, lmax, Cin, Cout,
                 work=2.0 * N * L * Cin * Cout)
        if ctx.needs_input_grad[2]:
            db = dout[:, 0, :].sum(dim=0)
        return dx, dw, db


    @staticmethod
    def _backward_tc(ctx, dout):
        N, L, Cin, Cout, lmax = ctx.dims
        nl, R = lmax + 1, N * L
        dx = dw = db = None
        if Cin < Cout:
            x, we = ctx.saved_tensors
            KP = we.shape[1]
            dob = _so3_expand(dout, lmax, Cout, False)
            if ctx.needs_input_grad[0]:
                KN = _pad(nl * Cin, 16)
                wet = torch.zeros(KN, Cout, dtype=torch.bfloat16, device=we.device)
                wet[:nl * Cin] = we[:, :nl * Cin].t()
                dx = _so3_select(_f32_from_gemm(dob, wet, R, KN, Cout), None, N, lmax, Cin)
            if ctx.needs_input_grad[1] or ctx.needs_input_grad[2]:
                xe = _so3_expand(x, lmax, KP, True, nl * Cin)
                dwe = torch.empty(Cout, KP, dtype=torch.float32, device=x.device)
                _gemm_tn(dob, xe, dwe, False, R, Cout, KP)
                dw = dwe[:, :nl * Cin].reshape(Cout, nl, Cin).permute(1, 0, 2).contiguous()
                db = dwe[:, nl * Cin].contiguous()
        else:
            xb, wa = ctx.saved_tensors
            NP = wa.shape[0]
            dye = _so3_expand(dout, lmax, NP, True)
            if ctx.needs_input_grad[0]:
                dx = _f32_from_gemm(dye, wa.t().contiguous(), R, Cin, NP).view(N, L, Cin)
            if ctx.needs_input_grad[1]:
                dwa = torch.empty(NP, Cin, dtype=torch.float32, device=xb.device)
                _gemm_tn(dye, xb, dwa, False, R, NP, Cin)
                dw = dwa[:nl * Cout].view(nl, Cout, Cin)
            if ctx.needs_input_grad[2]:
                db = dout[:, 0, :].sum(dim=0)
        return dx, dw, db


class FFNFunction(torch.autograd.Function):
    """FeedForwardNetwork core (EF_layers.py:234-270) in bf16 mode: so3_linear_1 -> SeparableS2Activation on the
    (lmax, lmax) grid -> so3_linear_2 with the [N, L, hidden] activations kept in bf16 end to end: two tcgen05 GEMMs
    around the tcgen05 S2 kernel, no fp32 copy of the hidden tensor ever written.
    (x [N, L, C], w1 [nl, CH, C], b1 [CH], gate [N, CH], w2 [nl, Co, CH], b2 [Co]) -> [N, L, Co] fp32."""

    @staticmethod
    def eligible(C, CH, Co, L, G) -> bool:
        return (_PRECISION == "bf16" and C % 16 == 0 and Co % 16 == 0 and CH % 128 == 0 and L <= 64 and G <= 224)

    @staticmethod
    def forward(ctx, x, w1, b1, gate, w2, b2, to_g, from_g):
        check_cuda(x, w1, gate, w2)
        x = x.contiguous().float()
        gate = gate.contiguous().float()
        N, L, C = x.shape
        nl, CH, Co = w1.shape[0], w1.shape[1], w2.shape[1]
        lmax, R, G = nl - 1, N * L, to_g.shape[0]
        KP = _pad(nl * C + 1, 64)
        NP = _pad(nl * Co, 16)
        w1e = _so3_w_expand_in(w1.detach().float(), b1.detach().float(), KP)
        w2a = _so3_w_stack(w2.detach().float(), NP)
        xe = _so3_expand(x, lmax, KP, True, nl * C)
        h = torch.empty(R, CH, dtype=torch.bfloat16, device=x.device)
        _gemm_nt(0, xe, w1e, None, h, CH, None, 0, 0, R, CH, KP)
        del xe
        yb = torch.empty_like(h)
        call("efl_s2act_tc_nodes", 0, L, G, CH, ptr(h), None, ptr(yb), None, ptr(gate), CH, ptr(to_g), ptr(from_g), N,
             work=N * 4.0 * G * L * CH, tag="efl_s2act_tc_nodes_fwd")
        out = _so3_select(_f32_from_gemm(yb, w2a, R, NP, CH), b2.detach().float().contiguous(), N, lmax, Co)
        ctx.save_for_backward(x, gate, h, yb, w1e, w2a, to_g, from_g)
        ctx.dims = (N, L, C, CH, Co, lmax, G)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, gate, h, yb, w1e, w2a, to_g, from_g = ctx.saved_tensors
        N, L, C, CH, Co, lmax, G = ctx.dims
        nl, R = lmax + 1, N * L
        KP, NP = w1e.shape[1], w2a.shape[0]
        dout = dout.contiguous().float()
        dye = _so3_expand(dout, lmax, NP, True)
        dyb = torch.empty(R, CH, dtype=torch.bfloat16, device=x.device)
        _gemm_nt(0, dye, w2a.t().contiguous(), None, dyb, CH, None, 0, 0, R, CH, NP)
        dw2a = torch.empty(NP, CH, dtype=torch.float32, device=x.device)
        _gemm_tn(dye, yb, dw2a, False, R, NP, CH)
        dw2 = dw2a[:nl * Co].view(nl, Co, CH)
        db2 = dout[:, 0, :].sum(dim=0)
        del dye
        dh = torch.empty_like(h)
        dgb = torch.empty(N, CH, dtype=torch.bfloat16, device=x.device)
        call("efl_s2act_tc_nodes", 1, L, G, CH, ptr(h), ptr(dyb), ptr(dh), ptr(dgb), ptr(gate), CH, ptr(to_g),
             ptr(from_g), N, work=N * 6.0 * G * L * CH, tag="efl_s2act_tc_nodes_bwd")
        del dyb
        KN = _pad(nl * C, 16)
        w1t = torch.zeros(KN, CH, dtype=torch.bfloat16, device=x.device)
        w1t[:nl * C] = w1e[:, :nl * C].t()
        dx = _so3_select(_f32_from_gemm(dh, w1t, R, KN, CH), None, N, lmax, C)
        xe = _so3_expand(x, lmax, KP, True, nl * C)
        dw1e = torch.empty(CH, KP, dtype=torch.float32, device=x.device)
        _gemm_tn(dh, xe, dw1e, False, R, CH, KP)
        dw1 = dw1e[:, :nl * C].reshape(CH, nl, C).permute(1, 0, 2).contiguous()
        db1 = dw1e[:, nl * C].contiguous()
        return dx, dw1, db1, dgb.float(), dw2, db2, None, None


class RMSNormSHFunction(torch.autograd.Function):
    """Centred, degree-balanced RMS norm with per-degree gain and l=0 bias (EF_layers.py:2099-2192)."""

    @staticmethod
    def forward(ctx, x, gamma, beta, lmax, eps):
        check_cuda(x, gamma, beta)
        x = x.contiguous().float()
        N, L, C = x.shape
        y = torch.empty_like(x)
        g = gamma.detach().contiguous().float()
        call("efl_rmsnorm_sh_fwd", lmax, C, ptr(x), ptr(g), ptr(beta.detach().contiguous().float()), float(eps), ptr(y), N,
             work=2.0 * x.numel() * 4)
        ctx.save_for_backward(x, g)
        ctx.misc = (lmax, eps)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, g = ctx.saved_tensors
        lmax, eps = ctx.misc
        dy = dy.contiguous().float()
        N, L, C = x.shape
        dx = torch.empty_like(x)
        nblk = max(1, min(296, (N + 7) // 8))
        part = torch.empty(nblk, (lmax + 2) * C, dtype=torch.float32, device=x.device)
        call("efl_rmsnorm_sh_bwd", lmax, C, ptr(x), ptr(g), float(eps), ptr(dy), ptr(dx), ptr(part), nblk, N,
             work=3.0 * x.numel() * 4)
        tot = part.sum(dim=0)
        return dx, tot[:(lmax + 1) * C].view(lmax + 1, C), tot[(lmax + 1) * C:], None, None


# ---------------------------------------------------------------------------
# generic deterministic segment ops over an EdgeCSR (CProMG message passing)
# ---------------------------------------------------------------------------
class SegmentSoftmaxFunction(torch.autograd.Function):
    """torch_scatter.scatter_softmax(src, index, dim=0) with index = the CSR's destination of every edge
    (reference model/CProMG.py:69).  logits [E, H] in ORIGINAL edge order; one warp per destination over the sorted
    CSR, no atomics (csrc/attention.cu).  The kernel's 1e-16 in the denominator (pyg's softmax, EF_layers.py:1180)
    is below one ulp of a sum that is always >= 1, so the result equals scatter_softmax bit for bit."""

    @staticmethod
    def forward(ctx, logits, csr):
        check_cuda(logits)
        E, H = logits.shape
        assert E == csr.num_edges, "logits rows must match the edge list"
        perm = csr.index("perm")
        ls = logits.detach().float().index_select(0, perm).contiguous()
        alpha_s = torch.empty_like(ls)
        if E > 0:
            call("efl_segment_softmax_fwd", ptr(ls), H, ptr(csr.rowptr), ptr(alpha_s), 0, csr.num_dst,
                 work=E * H * 8.0)
        ctx.save_for_backward(alpha_s)
        ctx.csr = csr
        out = torch.empty_like(alpha_s)
        out.index_copy_(0, perm, alpha_s)
        return out

    @staticmethod
    def backward(ctx, dalpha):
        (alpha_s,) = ctx.saved_tensors
        csr = ctx.csr
        E, H = alpha_s.shape
        perm = csr.index("perm")
        da_s = dalpha.float().index_select(0, perm).contiguous()
        dl_s = torch.empty_like(da_s)
        if E > 0:
            call("efl_segment_softmax_bwd", ptr(alpha_s), ptr(da_s), 1, 0, H, ptr(csr.rowptr), ptr(dl_s), 0, csr.num_dst,
                 work=E * H * 12.0)
        out = torch.empty_like(dl_s)
        out.index_copy_(0, perm, dl_s)
        return out, None


class SegmentSumFunction(torch.autograd.Function):
    """torch_scatter.scatter_sum(src, index, dim=0, dim_size=N) (CProMG.py:76): values [E, D] in original edge order
    -> [num_dst, D], fixed summation order per destination (efl_segment_sum_rows), zero rows for empty segments."""

    @staticmethod
    def forward(ctx, values, csr):
        check_cuda(values)
        values = values.contiguous().float()
        E, D = values.shape
        assert E == csr.num_edges, "value rows must match the edge list"
        out = torch.empty(csr.num_dst, D, dtype=torch.float32, device=values.device)
        call("efl_segment_sum_rows", ptr(values), 1, D, 0, D, ptr(csr.rowptr), ptr(csr.perm), ptr(out), csr.num_dst, 0,
             work=(E + csr.num_dst) * D * 4.0)
        ctx.csr = csr
        return out

    @staticmethod
    def backward(ctx, dout):
        return dout.index_select(0, ctx.csr.index("dst_orig")), None


class GatherRowsFunction(torch.autograd.Function):
    """h[index of every edge] with side = 'src' | 'dst' (CProMG.py:64-65: h_keys[col], h_queries[row]).  The backward
    is the deterministic segment sum over that side's grouping instead of torch's atomic index_add_."""

    @staticmethod
    def forward(ctx, h, csr, side):
        check_cuda(h)
        assert side in ("src", "dst")
        ctx.csr, ctx.side, ctx.rows = csr, side, h.shape[0]
        return h.index_select(0, csr.index("src_orig" if side == "src" else "dst_orig"))

    @staticmethod
    def backward(ctx, dout):
        csr, side = ctx.csr, ctx.side
        dout = dout.contiguous().float()
        E, D = dout.shape
        dh = torch.empty(ctx.rows, D, dtype=torch.float32, device=dout.device)
        if side == "dst":
            seg, idx = csr.rowptr, csr.perm
        else:
            seg, idx = csr.src_ptr, csr.index("src_group")
        call("efl_segment_sum_rows", ptr(dout), 1, D, 0, D, ptr(seg), ptr(idx), ptr(dh), ctx.rows, 0,
             work=(E + ctx.rows) * D * 4.0)
        return dh, None, None


class LinearFunction(torch.autograd.Function):
    """y = x W^T + b on the library's GEMMs: fp32 SIMT in accuracy mode, tcgen05 (bf16 operands, fp32 accumulate and
    output) in bf16 mode; wgrad through the deterministic split-K TN GEMM.  x [M, K], W [N, K], b [N] or None."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        check_cuda(x, weight)
        x = x.contiguous().float()
        M, K = x.shape
        N = weight.shape[0]
        tc = _PRECISION == "bf16" and K % 8 == 0 and N % 16 == 0 and M > 0
        dt = torch.bfloat16 if tc else torch.float32
        xa = x.to(dt)
        w = weight.detach().to(dt).contiguous()
        b = None if bias is None else bias.detach().float().contiguous()
        out = torch.empty(M, N, dtype=torch.float32, device=x.device)
        if M > 0:
            if tc:
                _gemm_nt(1, xa, w, b, out, 8, out, N, N, M, N, K)
            else:
                _gemm_nt(0, xa, w, b, out, N, None, 0, 0, M, N, K)
        ctx.save_for_backward(xa, w)
        ctx.has_bias = bias is not None
        return out

    @staticmethod
    def backward(ctx, dout):
        xa, w = ctx.saved_tensors
        M, K = xa.shape
        N = w.shape[0]
        tc = xa.dtype == torch.bfloat16
        dy = dout.contiguous().to(xa.dtype)
        dx = dw = db = None
        if M == 0:
            return (torch.zeros(M, K, device=xa.device), torch.zeros(N, K, device=xa.device),
                    torch.zeros(N, device=xa.device) if ctx.has_bias else None)
        if ctx.needs_input_grad[0]:
            dx = torch.empty(M, K, dtype=torch.float32, device=xa.device)
            wt = w.t().contiguous()
            if tc and K % 16 == 0 and N % 8 == 0:
                _gemm_nt(1, dy, wt, None, dx, 8, dx, K, K, M, K, N)
            elif tc:
                _gemm_nt(0, dy.float(), wt.float(), None, dx, K, None, 0, 0, M, K, N)
            else:
                _gemm_nt(0, dy, wt, None, dx, K, None, 0, 0, M, K, N)
        if ctx.needs_input_grad[1]:
            dw = torch.empty(N, K, dtype=torch.float32, device=xa.device)
            if tc and K % 16 == 0:
                _gemm_tn(dy, xa, dw, False, M, N, K)
            else:
                _gemm_tn(dy.float(), xa.float(), dw, False, M, N, K)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = torch.empty(N, dtype=torch.float32, device=xa.device)
            _colsum(dy, M, N, db, False)
        return dx, dw, db
