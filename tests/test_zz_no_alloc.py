"""efl_colsum_static: the column sum for callers that are not able to allocate device memory.

Who needs it: a process that sizes its device memory pool once at start-up and then closes it, a plugin that works
inside buffers its host program handed over, and code between cudaStreamBeginCapture / EndCapture (cudaMalloc fails
there).  efl_colsum asks such a caller for a workspace it has nowhere to get; efl_colsum_static keeps the partial sums
in a 4 MiB array that is part of libefl.so and takes inputs and outputs only.

This file is named test_zz_* so that pytest collects it after every other test file: with ``-x`` a failure here cannot
hide another test.  It must pass on its own in a fresh process and at the end of the whole suite.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_FLOATS = 1 << 20  # EFL_COLSUM_STATIC_MAX_FLOATS, include/efl.h


def _both(X, rows, cols, start):
    """(efl_colsum with a torch-allocated workspace, efl_colsum_static) on the same input, out preset to `start`."""
    from singa_b200 import _lib
    bf16 = int(X.dtype == torch.bfloat16)
    ws = torch.empty(((rows + 511) // 512 + 1) * cols, dtype=torch.float32, device=X.device)
    a = torch.full((cols,), start, dtype=torch.float32, device=X.device)
    b = torch.full((cols,), start, dtype=torch.float32, device=X.device)
    _lib.call("efl_colsum", bf16, _lib.ptr(X), X.stride(0), rows, cols, _lib.ptr(a), 1, _lib.ptr(ws))
    _lib.call("efl_colsum_static", bf16, _lib.ptr(X), X.stride(0), rows, cols, _lib.ptr(b), 1)
    torch.cuda.synchronize()
    return a, b


def test_static_scratch_column_sum_matches_the_workspace_entry_point_twice_in_a_row():
    """Two calls back to back through the C ABI, the second on a smaller, different matrix, so it runs over whatever
    the first left in the shared array.  Both entry points launch the same kernels in the same order, so the tolerance
    is the one the bias-gradient test uses for repeated calls: bit equality.  The fp64 sum is checked as well, so the
    two cannot agree on a wrong answer."""
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    g = torch.Generator(device="cpu").manual_seed(20261106)
    shapes = [(262144, 1248, 1248, torch.bfloat16),   # conv_1 bias gradient of a full edge chunk: 640 K scratch floats
              (5003, 784, 800, torch.bfloat16),        # ragged row count, pitch > cols
              (1000, 20, 24, torch.float32)]           # scalar fp32 kernel
    for rows, cols, pitch, dtype in shapes:
        X = torch.randn(rows, pitch, generator=g).to(dev).to(dtype)[:, :cols]
        a, b = _both(X, rows, cols, 3.0)
        assert torch.equal(a, b), (rows, cols, float((a - b).abs().max()))
        ref = X.double().sum(dim=0) + 3.0
        tol = 1e-5 * max(1.0, float(ref.abs().max())) + 2e-6 * rows ** 0.5
        assert float((b.double() - ref).abs().max()) <= tol


def test_static_scratch_column_sum_refuses_shapes_beyond_the_built_in_array():
    """One row block more than the array holds: EFL_ERR_ARG with both numbers in the message, output untouched."""
    from singa_b200 import _lib
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    cols = 2048
    rows = 512 * (MAX_FLOATS // cols)  # (rows/512 + 1) * cols = MAX_FLOATS + cols
    X = torch.zeros(8, cols, dtype=torch.bfloat16, device=dev)  # never read: the call is refused before any launch
    out = torch.full((cols,), 7.0, device=dev)
    with pytest.raises(_lib.EflError, match=str(MAX_FLOATS)):
        _lib.call("efl_colsum_static", 1, _lib.ptr(X), cols, rows, cols, _lib.ptr(out), 0)
    torch.cuda.synchronize()
    assert bool((out == 7.0).all())
    # the largest shape that fits is accepted
    rows_ok = rows - 512
    Xok = torch.ones(rows_ok, cols, dtype=torch.bfloat16, device=dev)
    _lib.call("efl_colsum_static", 1, _lib.ptr(Xok), cols, rows_ok, cols, _lib.ptr(out), 0)
    torch.cuda.synchronize()
    assert bool((out == float(rows_ok)).all())
