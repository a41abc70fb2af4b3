"""tcgen05 GEMM engine against torch.matmul on the exact shapes of the SO(2) linear maps (bf16 inputs, so the
comparison is bf16-rounding exact up to fp32 accumulation order)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

NT_SHAPES = [  # (M, N, K)
    (1000, 1248, 224), (2002, 1536, 192), (2002, 1280, 160), (1000, 784, 896), (2002, 1344, 768),
    (2002, 1120, 640), (1000, 224, 1248), (2002, 192, 1536), (2002, 160, 1280), (1000, 896, 784),
    (2002, 768, 1344), (2002, 640, 1120), (128, 256, 64), (70, 16, 40), (40000, 1248, 224),
]


def _rand(m, n, scale=1.0):
    return (torch.randn(m, n, device=DEV) * scale).to(torch.bfloat16)


@pytest.fixture(params=[0, 1, 2], ids=["single_cta", "auto", "cta_pair"])
def pair_mode(request):
    """Every GEMM test runs on the single-CTA tiles, on the cta_group::2 pair tiles (256-row tiles, two SMs) and in
    the default mode, which picks tile mode and wgrad orientation per shape."""
    from singa_b200 import _lib
    prev = _lib.LIB.efl_tc_gemm_pair_mode(request.param)
    yield request.param
    _lib.LIB.efl_tc_gemm_pair_mode(prev)


@pytest.mark.parametrize("M,N,K", NT_SHAPES)
def test_nt_plain(M, N, K, pair_mode):
    from singa_b200 import tc_gemm
    torch.manual_seed(M + N + K)
    A, B = _rand(M, K), _rand(N, K, 0.1)
    bias = torch.randn(N, device=DEV)
    C = torch.zeros(M, N, dtype=torch.bfloat16, device=DEV)
    tc_gemm.gemm_nt(0, A, B, bias, C, N, None, 0, 0, M, N, K)
    ref = A.float() @ B.float().t() + bias
    err = float((C.float() - ref).abs().max()) / float(ref.abs().max())
    assert err < 1e-2, err


def test_nt_split_epilogue(pair_mode):
    from singa_b200 import tc_gemm
    M, N, K, ns = 777, 1248, 224, 352
    A, B = _rand(M, K), _rand(N, K, 0.1)
    bias = torch.randn(N, device=DEV)
    C = torch.zeros(M, N - ns, dtype=torch.bfloat16, device=DEV)
    C0 = torch.zeros(M, ns, dtype=torch.float32, device=DEV)
    tc_gemm.gemm_nt(1, A, B, bias, C, N - ns, C0, ns, ns, M, N, K)
    ref = A.float() @ B.float().t() + bias
    assert float((C0 - ref[:, :ns]).abs().max()) < 2e-3 * float(ref.abs().max())
    assert float((C.float() - ref[:, ns:]).abs().max()) < 1e-2 * float(ref.abs().max())


def test_nt_complex_epilogue(pair_mode):
    from singa_b200 import ops, tc_gemm
    E, K, half = 501, 192, 768
    A = _rand(2 * E, K)
    W = _rand(2 * half, K, 0.1)
    Wi = ops._interleave_rows(W)
    C = torch.zeros(2 * E, half, dtype=torch.bfloat16, device=DEV)
    tc_gemm.gemm_nt(2, A, Wi, None, C, half, None, 0, 0, 2 * E, 2 * half, K)
    y = (A.float() @ W.float().t()).view(E, 2, 2 * half)
    yr, yi = y[:, :, :half], y[:, :, half:]
    ref = torch.stack((yr[:, 0] - yi[:, 1], yr[:, 1] + yi[:, 0]), dim=1).reshape(2 * E, half)
    assert float((C.float() - ref).abs().max()) < 1e-2 * float(ref.abs().max())


@pytest.mark.parametrize("Mred,N1,N2", [(5000, 1248, 224), (9999, 1536, 192), (4097, 2560, 160), (3000, 784, 896),
                                         (6000, 2688, 768), (64, 128, 64), (130, 144, 16), (131072, 2240, 640),
                                         (20000, 1120, 640), (7000, 1344, 768), (300, 48, 1120)])
def test_tn(Mred, N1, N2, pair_mode):
    from singa_b200 import tc_gemm
    torch.manual_seed(Mred)
    A, B = _rand(Mred, N1, 0.3), _rand(Mred, N2, 0.3)
    C = torch.ones(N1, N2, dtype=torch.float32, device=DEV)
    tc_gemm.gemm_tn(A, B, C, True, Mred, N1, N2)
    ref = A.float().t() @ B.float() + 1.0
    err = float((C - ref).abs().max()) / float(ref.abs().max())
    assert err < 2e-3, err
    C2 = torch.full((N1, N2), 7.0, dtype=torch.float32, device=DEV)
    tc_gemm.gemm_tn(A, B, C2, False, Mred, N1, N2)
    assert float((C2 - (ref - 1.0)).abs().max()) / float(ref.abs().max()) < 2e-3
    C3 = torch.empty_like(C2)
    tc_gemm.gemm_tn(A, B, C3, False, Mred, N1, N2)
    assert torch.equal(C2, C3)  # deterministic
