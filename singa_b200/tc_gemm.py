"""bf16 tcgen05 GEMM entry points (libefl.so: efl_tc_gemm_nt / efl_tc_gemm_tn)."""
from __future__ import annotations

import torch

from ._lib import LIB, call, ptr


def gemm_nt(epi, A, B, bias, C, ldc, C0, ldc0, nsplit, M, N, K):
    """C[M, N] = A[M, K] @ B[N, K]^T with the fused epilogue ``epi`` (0 plain, 1 split, 2 complex combine)."""
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16
    nout = N // 2 if epi == 2 else N
    call("efl_tc_gemm_nt", epi, ptr(A), A.stride(0), ptr(B), B.stride(0), ptr(bias), ptr(C), ldc, ptr(C0), ldc0, nsplit,
         M, N, K, work=2.0 * M * N * K,
         nbytes=2.0 * (M * K + N * K) + (4.0 * nsplit + 2.0 * (nout - nsplit) if epi == 1 else 2.0 * nout) * M)


def gemm_tn(A, B, C, accumulate, Mred, N1, N2):
    """C[N1, N2] (+)= A[Mred, N1]^T @ B[Mred, N2] (fp32 output, deterministic split reduction)."""
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16 and C.dtype == torch.float32
    splits = LIB.efl_tc_gemm_tn_splits(Mred, N1, N2)
    ws = torch.empty(splits * N1 * N2, dtype=torch.float32, device=A.device)
    call("efl_tc_gemm_tn", ptr(A), A.stride(0), ptr(B), B.stride(0), ptr(C), C.stride(0), int(accumulate), ptr(ws), Mred,
         N1, N2, work=2.0 * Mred * N1 * N2, nbytes=2.0 * Mred * (N1 + N2) + 4.0 * N1 * N2)
