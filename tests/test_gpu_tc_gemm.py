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
    # round 2: the SO(2) linears as real GEMMs over M = edges ([x_re | x_im] rows, ops._Weights): forward ...
    (1001, 1536, 384), (1001, 1280, 320), (1001, 1344, 1536), (1001, 1120, 1280),
    # ... and dgrad, plus the m = 0 GEMM with its bias column (K = 16 + 224) and its dgrad
    (1001, 384, 1536), (1001, 320, 1280), (1001, 1536, 1344), (1001, 1280, 1120), (1001, 240, 1248),
    # lmax = 8 (config 5): K0 = 32 + 288
    (1001, 1504, 320), (1001, 320, 1504), (1001, 2048, 512), (1001, 1792, 2048), (1001, 2048, 1792),
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
    Wi = torch.stack((W[:half], W[half:]), dim=1).reshape(2 * half, K).contiguous()   # rows (real_q, imag_q)
    C = torch.zeros(2 * E, half, dtype=torch.bfloat16, device=DEV)
    tc_gemm.gemm_nt(2, A, Wi, None, C, half, None, 0, 0, 2 * E, 2 * half, K)
    y = (A.float() @ W.float().t()).view(E, 2, 2 * half)
    yr, yi = y[:, :, :half], y[:, :, half:]
    ref = torch.stack((yr[:, 0] - yi[:, 1], yr[:, 1] + yi[:, 0]), dim=1).reshape(2 * E, half)
    assert float((C.float() - ref).abs().max()) < 1e-2 * float(ref.abs().max())


@pytest.mark.parametrize("Mred,N1,N2", [(5000, 1248, 224), (9999, 1536, 192), (4097, 2560, 160), (3000, 784, 896),
                                         (6000, 2688, 768), (64, 128, 64), (130, 144, 16), (131072, 2240, 640),
                                         (20000, 1120, 640), (7000, 1344, 768), (300, 48, 1120),
                                         # round 2 wgrad shapes (edge rows): N2 = [x_re | x_im] widths, bias column
                                         (5000, 1248, 240), (5000, 1536, 384), (5000, 1280, 320), (5000, 1344, 1536),
                                         (5000, 1120, 1280), (5000, 1504, 320), (5000, 2048, 512)])
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


def test_nt_and_tn_on_column_slices_of_wider_rows(pair_mode):
    """Operands and outputs as column blocks of wider per-edge rows (pitch != width, block offsets that are only
    16-byte aligned), exactly how ops._chunk_forward addresses the edge-row buffers."""
    from singa_b200 import tc_gemm
    torch.manual_seed(3)
    M = 3001
    rowA, rowC = _rand(M, 944), torch.zeros(M, 3712, dtype=torch.bfloat16, device=DEV)
    A = rowA[:, 240:624]                      # m = 1 block of an A row
    W = _rand(1536, 384, 0.1)
    tc_gemm.gemm_nt(0, A, W, None, rowC[:, 896:], 3712, None, 0, 0, M, 1536, 384)
    ref = A.float() @ W.float().t()
    assert float((rowC[:, 896:2432].float() - ref).abs().max()) < 1e-2 * float(ref.abs().max())
    assert float(rowC[:, :896].abs().max()) == 0.0 and float(rowC[:, 2432:].abs().max()) == 0.0
    G = _rand(M, 4064, 0.3)
    dW = torch.empty(1536, 384, dtype=torch.float32, device=DEV)
    tc_gemm.gemm_tn(G[:, 1248:2784], A, dW, False, M, 1536, 384)
    refw = G[:, 1248:2784].float().t() @ A.float()
    assert float((dW - refw).abs().max()) < 2e-3 * float(refw.abs().max())
