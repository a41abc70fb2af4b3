"""ctypes binding of the C-ABI CUDA library (include/efl.h).

The library is built in-tree by ``__graft_entry__.build()`` as
``singa_b200/libefl.so``.  There is no CPU fallback: importing this module without
the library raises, and every op checks that its tensors live on a CUDA device.
"""
from __future__ import annotations

import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libefl.so")

_P, _I, _L, _F = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float
_CODES = {"p": _P, "i": _I, "l": _L, "f": _F}

# name -> argument type codes (p pointer, i int, l long long, f float); every entry point returns int
SIGNATURES = {
    "efl_wigner_from_rotmat": "iipppip",
    "efl_rotate_in_fwd": "iiiipppppppippp" "liiip",
    "efl_rotate_in_bwd": "iiiipppppppippp" "lpp" "iip",
    "efl_rotate_in_bwd_nodes": "iiiiipppppp" "ippp" "lpp" "iiiiip",
    "efl_segment_sum_rows": "piiiipppiip",
    "efl_rotate_back_aggregate_fwd": "iiiiippp" "lppppp" "iiip",
    "efl_rotate_back_aggregate_bwd": "iiiiippp" "lppp" "lpppppp" "liiip",
    "efl_edge_degree_fwd": "iiippppfpip",
    "efl_edge_degree_bwd": "iiipppfpip",
    "efl_alpha_logits_fwd": "piiippppip",
    "efl_alpha_logits_bwd": "piiippppp" "iipiip",
    "efl_segment_softmax_fwd": "pippiip",
    "efl_segment_softmax_bwd": "ppilippiip",
    "efl_s2act_fwd": "iiiii" "ippp" "ippp" "plpplp",
    "efl_s2act_bwd": "iiiii" "ippp" "ippp" "ippp" "i" "plplpplp",
    "efl_gemm_nt_f32": "iplplpplpliiiip",
    "efl_gemm_tn_f32": "plplplipiiip",
    "efl_colsum": "ipliipipp",
    "efl_colsum_static": "ipliipip",
    "efl_split3": "plliplii" "p",
    "efl_node_rows_pack": "piiiiipp",
    "efl_node_rows_unpack": "piiiippp",
    "efl_bf16_silu": "pplip",
    "efl_tc_gemm_nt": "iplplpplpliiiip",
    "efl_tc_gemm_tn": "plplplipiiip",
    "efl_s2act_tc": "iiii" "ppp" "l" "pp" "l" "pp" "i" "p",
    "efl_s2act_tc_nodes": "iiii" "ppppp" "l" "pp" "i" "p",
    "efl_so3_linear_fwd": "ppppiiiip",
    "efl_so3_linear_bwd_data": "pppiiiip",
    "efl_so3_linear_bwd_weight": "ppppiiiip",
    "efl_so3_expand": "ppliiiiip",
    "efl_so3_select": "plppliip",
    "efl_radial_mlp_fwd": "pi" + "p" * 10 + "ipip",
    "efl_radial_mlp_bwd": "pi" + "p" * 10 + "ipppiip",
    "efl_radial_hidden_fwd": "pi" + "p" * 8 + "pip",
    "efl_radial_hidden_bwd": "pi" + "p" * 8 + "pppiip",
    "efl_embedding_grad": "ppliiippip",
    "efl_radius_graph_bbox": "ppiippp",
    "efl_radius_graph_count": "ppiipfiiifiippppppppl" "p",
    "efl_radius_graph_fill": "iifiiifippppp" "ip",
    "efl_radius_graph_emit": "pppilpip",
    "efl_knn_graph": "ppiipfiiiiipppppppppl" "p",
    "efl_csr_from_edge_index": "ppiiipppppppp" "plpp",
    "efl_rmsnorm_sh_fwd": "iipppfpip",
    "efl_rmsnorm_sh_bwd": "iippfpppiip",
}


class EflError(RuntimeError):
    pass


def _load():
    if not os.path.isfile(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build the CUDA library first (python -c 'import __graft_entry__ as g; g.build()'). "
            "singa_b200 has no CPU or PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, sig in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.argtypes = [_CODES[c] for c in sig]
        fn.restype = _I
    lib.efl_last_error_string.restype = ctypes.c_char_p
    lib.efl_version.restype = _I
    lib.efl_gemm_tn_splits.argtypes = [_I]
    lib.efl_gemm_tn_splits.restype = _I
    lib.efl_tc_gemm_tn_splits.argtypes = [_I, _I, _I]
    lib.efl_tc_gemm_tn_splits.restype = _I
    lib.efl_tc_gemm_pair_mode.argtypes = [_I]
    lib.efl_tc_gemm_pair_mode.restype = _I
    lib.efl_tc_gemm_ring_kb.argtypes = [_I]
    lib.efl_tc_gemm_ring_kb.restype = _I
    lib.efl_radial_mlp_part_size.argtypes = [_I, _I]
    lib.efl_radial_mlp_part_size.restype = _I
    lib.efl_radial_hidden_part_size.argtypes = [_I]
    lib.efl_radial_hidden_part_size.restype = _I
    lib.efl_radius_graph_temp_bytes.argtypes = [_L]
    lib.efl_radius_graph_temp_bytes.restype = _L
    lib.efl_csr_temp_bytes.argtypes = [_L, _L]
    lib.efl_csr_temp_bytes.restype = _L
    lib.efl_rotate_back_parts.argtypes = [_I, _I]
    lib.efl_rotate_back_parts.restype = _I
    return lib


LIB = _load()


def ptr(t):
    """Device pointer of a tensor (or None -> NULL)."""
    if t is None:
        return None
    if isinstance(t, int):
        return t
    assert t.is_cuda, "singa_b200 ops need CUDA tensors (no CPU fallback)"
    return t.data_ptr()


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_ptr():
    """Raw handle of torch's current stream on the current device.  ``torch.cuda.current_stream()`` builds a Stream
    object per call (16 us: 13 ms of host time per step over ~830 entry-point calls); the raw getter is ~50x cheaper."""
    if _raw_stream is not None:
        return _raw_stream(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream


# kernels launched per entry point (for the launch counter bench.py reports)
_LAUNCHES = {"efl_csr_from_edge_index": 12, "efl_embedding_grad": 2, "efl_gemm_tn_f32": 2, "efl_colsum": 2, "efl_colsum_static": 2, "efl_tc_gemm_tn": 2, "efl_so3_linear_bwd_weight": 2}
launch_count = 0
_profile = None  # None, or a list receiving (name, work, start_event, end_event)
_profile_only = None  # None = every entry point, or the set of names (tags) that get event pairs


def start_profile(only=None):
    """Record a CUDA-event pair around entry-point calls (on the launching stream) until stop_profile().  ``only``
    restricts the pairs to a set of entry-point names / tags: two event records per call are host time, and with
    ~5 000 calls per step the full profile can make a GPU-bound step host-bound."""
    global _profile, _profile_only
    _profile = []
    _profile_only = set(only) if only else None


def stop_profile():
    """-> {name: dict(ms=total ms, calls=n, work=sum of algorithmic work units, bytes=sum of algorithmic bytes)}"""
    global _profile, _profile_only
    rec, _profile, _profile_only = _profile or [], None, None
    torch.cuda.synchronize()
    out = {}
    for name, work, nbytes, a, b in rec:
        d = out.setdefault(name, dict(ms=0.0, calls=0, work=0.0, bytes=0.0))
        d["ms"] += a.elapsed_time(b)
        d["calls"] += 1
        d["work"] += work
        d["bytes"] += nbytes
    return out


def call(name: str, *args, work: float = 0.0, nbytes: float = 0.0, tag: str = None):
    """Invoke an entry point; the trailing stream argument is appended here.  ``work`` is the algorithmic
    FLOPs (GEMMs) or bytes (HBM-bound kernels) of this call and ``nbytes`` the algorithmic operand + result bytes of a
    GEMM (so its measured DRAM traffic has something to be set against); both used only by the profiler."""
    global launch_count
    launch_count += _LAUNCHES.get(name, 1)
    if _profile is not None and (_profile_only is None or (tag or name) in _profile_only):
        a = torch.cuda.Event(enable_timing=True)
        b = torch.cuda.Event(enable_timing=True)
        a.record()
        rc = getattr(LIB, name)(*args, stream_ptr())
        b.record()
        _profile.append((tag or name, work, nbytes, a, b))
    else:
        rc = getattr(LIB, name)(*args, stream_ptr())
    if rc != 0:
        raise EflError(f"{name} failed (code {rc}): {LIB.efl_last_error_string().decode()}")


def check_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise EflError("singa_b200 kernels only run on CUDA tensors; there is no CPU path")
