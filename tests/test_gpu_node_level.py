"""Node-level kernels through the C ABI against the CPU oracle, autograd of the oracle
giving the reference gradients.  fp32 mode: 1e-5 / 1e-4; bf16 tensor-core mode (operand layout kernels of
csrc/so3lin.cu + efl_tc_gemm_*): stated looser bound 2e-2 outputs / 6e-2 gradients relative to ||ref||_inf."""
import numpy as np
import pytest
import torch

from tests.parity_util import oracle_tables, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _rand(shape, seed, scale=1.0):
    return torch.from_numpy((np.random.RandomState(seed).normal(size=shape) * scale).astype(np.float32))


def _tol(precision):
    return (1e-5, 1e-4) if precision == "fp32" else (2e-2, 6e-2)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("cin,cout", [(16, 512), (512, 16), (112, 16), (16, 16)])
@pytest.mark.parametrize("n", [1, 77, 300])
def test_so3_linear_matches_oracle(precision, cin, cout, n):
    from oracle import ef_oracle as eo
    from singa_b200 import ops
    t = oracle_tables()
    ops.set_precision(precision)
    try:
        x = _rand((n, 49, cin), 1)
        w = _rand((7, cout, cin), 2, cin ** -0.5)
        b = _rand((cout,), 3)
        r = _rand((n, 49, cout), 4)
        xr, wr, br = (v.clone().requires_grad_(True) for v in (x, w, b))
        ref = eo.so3_linear({"weight": wr, "bias": br}, "", xr, t)
        (ref * r).sum().backward()
        xg, wg, bg = (v.to(DEV).requires_grad_(True) for v in (x, w, b))
        out = ops.SO3LinearFunction.apply(xg, wg, bg)
        (out * r.to(DEV)).sum().backward()
    finally:
        ops.set_precision("fp32")
    to, tg = _tol(precision)
    assert rel_err(out, ref) < to
    assert rel_err(xg.grad, xr.grad) < tg
    assert rel_err(wg.grad, wr.grad) < tg
    assert rel_err(bg.grad, br.grad) < tg


def test_so3_expand_select_exact():
    """The operand layout kernels are pure data movement: bit-exact against index arithmetic."""
    from singa_b200 import ops
    n, lmax, C = 5, 6, 16
    L = 49
    x = _rand((n, L, C), 5).to(DEV)
    KP = 128
    xe = ops._so3_expand(x, lmax, KP, True, 7 * C).float().cpu().view(n, L, KP)
    xb = x.to(torch.bfloat16).float().cpu()
    want = torch.zeros(n, L, KP)
    for i in range(L):
        l = int(np.floor(np.sqrt(i)))
        want[:, i, l * C:(l + 1) * C] = xb[:, i]
    want[:, 0, 7 * C] = 1.0
    assert torch.equal(xe, want)
    plain = ops._so3_expand(x, lmax, 24, False).float().cpu().view(n, L, 24)
    assert torch.equal(plain[:, :, :C], xb) and float(plain[:, :, C:].abs().max()) == 0.0
    ye = _rand((n * L, 7 * C + 4), 6).to(DEV)
    bias = _rand((C,), 7).to(DEV)
    y = ops._so3_select(ye, bias, n, lmax, C).cpu()
    yec = ye.cpu().view(n, L, -1)
    for i in range(L):
        l = int(np.floor(np.sqrt(i)))
        w = yec[:, i, l * C:(l + 1) * C] + (bias.cpu() if i == 0 else 0.0)
        assert torch.equal(y[:, i], w), i


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("n", [3, 200])
def test_feed_forward_network_matches_oracle(precision, n):
    from oracle import ef_oracle as eo
    from singa_b200 import EF_layers as ef
    from tests.parity_util import build_model
    t = oracle_tables()
    model = build_model(DEV, 5, precision)
    try:
        ffn = model.blocks[0].ffn
        p = {k: v.detach().cpu().clone().requires_grad_(True) for k, v in ffn.named_parameters()}
        x = _rand((n, 49, 16), 8)
        r = _rand((n, 49, 16), 9)
        xr = x.clone().requires_grad_(True)
        ref = eo.feed_forward(p, "", xr, t)
        (ref * r).sum().backward()
        xg = x.to(DEV).requires_grad_(True)
        e = ef.SO3_Embedding(0, [6], 16, dtype=torch.float32, device=DEV)
        e.set_embedding(xg * 1.0)
        for q in ffn.parameters():
            q.grad = None
        out = ffn(e).embedding
        (out * r.to(DEV)).sum().backward()
    finally:
        from singa_b200 import ops
        ops.set_precision("fp32")
    to, tg = _tol(precision)
    assert rel_err(out, ref) < to, rel_err(out, ref)
    assert rel_err(xg.grad, xr.grad) < tg
    for k, q in ffn.named_parameters():
        assert rel_err(q.grad, p[k].grad) < tg, (k, rel_err(q.grad, p[k].grad))


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("k0,no", [(48, 576), (48, 112), (16, 64)])
@pytest.mark.parametrize("e", [1, 31, 1000, 20011])
def test_radial_mlp_matches_oracle(k0, no, e, precision):
    """RadialFunction (row a6).  fp32 mode: one fused SIMT kernel each way, 1e-5 outputs / 1e-4 gradients.  bf16 mode:
    hidden layers SIMT + the 16 -> NO layer as K=16 tcgen05 GEMMs, stated looser bound.  Both deterministic."""
    from oracle import ef_oracle as eo
    from singa_b200 import ops
    ops.set_precision(precision)
    try:
        _radial_case(eo, ops, k0, no, e, *_tol(precision))
    finally:
        ops.set_precision("fp32")


def _radial_case(eo, ops, k0, no, e, tol_out, tol_grad):
    names = ["net.0.weight", "net.0.bias", "net.1.weight", "net.1.bias", "net.3.weight", "net.3.bias", "net.4.weight",
             "net.4.bias", "net.6.weight", "net.6.bias"]
    shapes = [(16, k0), (16,), (16,), (16,), (16, 16), (16,), (16,), (16,), (no, 16), (no,)]
    base = [_rand(s, 20 + i, 0.3 if len(s) == 2 else 0.2) + (1.0 if n in ("net.1.weight", "net.4.weight") else 0.0)
            for i, (n, s) in enumerate(zip(names, shapes))]
    x = _rand((e, k0), 40)
    r = _rand((e, no), 41)
    pr = {n: v.clone().requires_grad_(True) for n, v in zip(names, base)}
    xr = x.clone().requires_grad_(True)
    ref = eo.radial_function(pr, "", xr)
    (ref * r).sum().backward()
    pg = [v.to(DEV).requires_grad_(True) for v in base]
    xg = x.to(DEV).requires_grad_(True)
    out = ops.RadialMLPFunction.apply(xg, *pg)
    (out * r.to(DEV)).sum().backward()
    assert rel_err(out, ref) < tol_out, rel_err(out, ref)
    assert rel_err(xg.grad, xr.grad) < tol_grad
    for n, q in zip(names, pg):
        assert rel_err(q.grad, pr[n].grad) < tol_grad, (n, rel_err(q.grad, pr[n].grad))
    pg2 = [v.to(DEV).requires_grad_(True) for v in base]
    xg2 = x.to(DEV).requires_grad_(True)
    out2 = ops.RadialMLPFunction.apply(xg2, *pg2)
    (out2 * r.to(DEV)).sum().backward()
    assert torch.equal(out, out2) and torch.equal(xg.grad, xg2.grad)
    for a, b in zip(pg, pg2):
        assert torch.equal(a.grad, b.grad)


@pytest.mark.parametrize("rows,cols,pitch", [(262144, 1248, 1248), (5003, 784, 800), (37, 576, 576), (1000, 20, 24), (0, 64, 64)])
def test_bias_gradient_column_sum_bf16(rows, cols, pitch):
    """efl_colsum (bias gradients of the SO(2) linear maps): vectorised bf16 path (cols, pitch multiples of 8) and the
    scalar fallback; fp32 accumulation in a fixed order, so repeated calls agree bit for bit."""
    from singa_b200 import ops
    torch.manual_seed(rows + cols)
    X = torch.randn(max(rows, 1), pitch, device=DEV).to(torch.bfloat16)[:rows, :cols]
    out = torch.full((cols,), 3.0, device=DEV)
    ops._colsum(X, rows, cols, out, True)
    ref = X.double().sum(dim=0) + 3.0
    tol = 1e-5 * max(1.0, float(ref.abs().max())) + 2e-6 * rows ** 0.5
    assert float((out.double() - ref).abs().max()) <= tol
    out2 = torch.full((cols,), 3.0, device=DEV)
    ops._colsum(X, rows, cols, out2, True)
    assert torch.equal(out, out2)


@pytest.mark.parametrize("n,ch", [(1, 128), (37, 512)])
def test_node_s2_activation_lmax8_gemm_route_matches_fp32_kernel(n, ch):
    """FeedForwardNetwork's S2 activation on the (8,8) grid (81 coefficients, 342 points; BASELINE config 5): the bf16
    tensor-core GEMM route (ops._node_s2_gemm_*) against the fp32 SIMT kernel, outputs and gradients, at the stated
    bf16 tolerance (2e-2 of the tensor's max)."""
    from singa_b200 import ops
    from singa_b200 import so3_constants as sc
    to, fr = sc.s2_grid_matrices(8, 8)
    G = to.shape[0] * to.shape[1]
    to_g = torch.tensor(to.reshape(G, -1), dtype=torch.float32, device=DEV).contiguous()
    from_g = torch.tensor(fr.reshape(G, -1), dtype=torch.float32, device=DEV).contiguous()
    assert to_g.shape == (342, 81)
    g = torch.Generator(device="cpu").manual_seed(n)
    x0 = torch.randn(n, 81, ch, generator=g).to(DEV)
    gate0 = torch.randn(n, ch, generator=g).to(DEV)
    dy = torch.randn(n, 81, ch, generator=g).to(DEV)
    res = {}
    try:
        for mode in ("fp32", "bf16"):
            ops.set_precision(mode)
            x, gate = x0.clone().requires_grad_(True), gate0.clone().requires_grad_(True)
            y = ops.NodeS2ActFunction.apply(gate, x, to_g, from_g)
            y.backward(dy)
            res[mode] = (y.detach(), x.grad, gate.grad)
    finally:
        ops.set_precision("fp32")
    for a, b, name in zip(res["fp32"], res["bf16"], ("y", "dx", "dgate")):
        err = float((a - b).abs().max() / a.abs().max())
        assert err < 2e-2, (name, err)


@pytest.mark.parametrize("mode,tol", [("fp32", 2e-5), ("bf16", 2e-2)])
def test_node_s2_activation_lmax8_matches_dense_autograd(mode, tol):
    """The same activation against an INDEPENDENT fp32 reference: torch autograd of the dense einsum formulation
    (EF_layers.py:1736-1773) on the host - both the fp32 SIMT kernel and the bf16 GEMM route with its pack / unpack /
    SiLU-epilogue kernels (the kernel-vs-kernel test above compares the path with itself)."""
    from singa_b200 import ops
    from singa_b200 import so3_constants as sc
    to, fr = sc.s2_grid_matrices(8, 8)
    to_c = torch.tensor(to.reshape(342, 81), dtype=torch.float32)
    from_c = torch.tensor(fr.reshape(342, 81), dtype=torch.float32)
    N, CH = 5, 128
    torch.manual_seed(0)
    x = torch.randn(N, 81, CH, requires_grad=True)
    gate = torch.randn(N, CH, requires_grad=True)
    dy = torch.randn(N, 81, CH)
    grid = torch.einsum('gl,nlc->ngc', to_c, x)
    y = torch.einsum('gl,ngc->nlc', from_c, torch.nn.functional.silu(grid))
    y = torch.cat((torch.nn.functional.silu(gate).unsqueeze(1), y[:, 1:]), dim=1)
    y.backward(dy)
    try:
        ops.set_precision(mode)
        xg, gg = x.detach().to(DEV).requires_grad_(True), gate.detach().to(DEV).requires_grad_(True)
        y2 = ops.NodeS2ActFunction.apply(gg, xg, to_c.to(DEV), from_c.to(DEV))
        y2.backward(dy.to(DEV))
    finally:
        ops.set_precision("fp32")
    rel = lambda a, b: float((a - b.cpu()).abs().max() / a.abs().max())
    errs = dict(y=rel(y.detach(), y2.detach()), dx=rel(x.grad, xg.grad), dgate=rel(gate.grad, gg.grad))
    assert max(errs.values()) < tol, errs
