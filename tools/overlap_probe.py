"""Can an HBM-bound kernel of one stream run in the shadow of the persistent tcgen05 GEMM of another stream?
Times (CUDA events, whole-device) a GEMM alone, an HBM-bound kernel alone (rotate-in forward of the same chunk) and
both launched back to back on two streams, for the default operand ring (192 KB: the GEMM CTA owns the SM's shared
memory, nothing co-resides) and a reduced ring (144 KB).  Serial sum vs concurrent time is the overlap potential."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from singa_b200 import _lib, ops, tc_gemm, synthetic as syn
from singa_b200 import EF_layers as ef

dev = torch.device("cuda", 0)
ops.set_precision("bf16")
graphs = [syn.make_complex(i) for i in range(8)]
g = {k: v.to(dev) for k, v in syn.collate(graphs).items()}
ei, pos = g["ei_pp"], g["pos_p"]
N, E = pos.shape[0], ei.shape[1]
rot = ef.init_edge_rot_mat(pos[ei[0]] - pos[ei[1]], device=dev)
wig = ops.wigner_from_rotmat(rot, 6, 2)
csr = ops.build_csr(ei, N, N)
d = ops.AttnDims(6, 2, 16, 128, 7, 32, 16)
L = ops._Layout(d)
x = torch.randn(N, 49, 16, device=dev)
rad = torch.randn(E, 576, device=dev)
A = torch.empty(E, L.wA, dtype=torch.bfloat16, device=dev)
V = (torch.randn(E, L.wV, device=dev) * 0.1).to(torch.bfloat16)
alpha = torch.rand(E, 7, device=dev)
out = torch.empty(N, 49, 112, device=dev)
M, Ng, K = E, 1344, 1536
Ag = (torch.randn(M, K, device=dev) * 0.1).to(torch.bfloat16)
Bg = (torch.randn(Ng, K, device=dev) * 0.1).to(torch.bfloat16)
Cg = torch.empty(M, Ng, dtype=torch.bfloat16, device=dev)
pa = ops._blk_ptrs(A, L.offA)
pv = ops._blk_ptrs(V, L.offV)
call, ptr = _lib.call, _lib.ptr


def gemm():
    tc_gemm.gemm_nt(0, Ag, Bg, None, Cg, Ng, None, 0, 0, M, Ng, K)


def rot_in():
    call("efl_rotate_in_fwd", 6, 2, 16, 1, ptr(x), ptr(x), ptr(csr.src_sorted), ptr(csr.dst_sorted), ptr(csr.perm), ptr(wig),
         ptr(rad), 576, pa[0], pa[1], pa[2], L.wA, L.pad, 0, E)


def rot_back():
    call("efl_rotate_back_aggregate_fwd", 6, 2, 7, 16, 1, pv[0], pv[1], pv[2], L.wV, ptr(alpha), ptr(csr.rowptr),
         ptr(csr.perm), ptr(wig), ptr(out), 0, N, 0)


s1, s2 = torch.cuda.Stream(priority=-1), torch.cuda.Stream()


def timed(fn_a, fn_b, reps=5):
    """fn_a on the high-priority stream, fn_b (may be None) on the other; device time from first launch to last end."""
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s1):
            fn_a()
        if fn_b is not None:
            with torch.cuda.stream(s2):
                fn_b()
        torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


for ring in (192, 144, 96):
    _lib.LIB.efl_tc_gemm_ring_kb(ring)
    for _ in range(2):
        gemm(); rot_in(); rot_back()
    tg, ti, tb = timed(gemm, None), timed(rot_in, None), timed(rot_back, None)
    print(json.dumps({"ring_kb": ring, "edges": E, "gemm_ms": round(tg, 3), "rotate_in_ms": round(ti, 3),
                      "rotate_back_ms": round(tb, 3),
                      "gemm+rotate_in concurrent": round(timed(gemm, rot_in), 3), "serial": round(tg + ti, 3),
                      "gemm+rotate_back concurrent": round(timed(gemm, rot_back), 3), "serial_back": round(tg + tb, 3),
                      "rotate_in first then gemm": round(timed(rot_in, gemm), 3)}))
