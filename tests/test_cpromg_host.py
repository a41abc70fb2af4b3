"""CProMG side on the host: the mirror of model/CProMG.py in singa_b200/cpromg.py against the
golden written by the UNMODIFIED reference ``Transformer`` (oracle/make_golden_cpromg.py -> cpromg_transformer.pt).

The message-passing ops have no CPU path (they are CUDA kernels); here they are stood in for by the oracle
(oracle/cpromg_oracle.py, test infrastructure) so that everything AROUND them -- module wiring, parameter names,
graph assembly, masks, decoder -- is checked where no GPU exists.  tests/test_gpu_cpromg.py runs the same golden
through the real kernels."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import cpromg_oracle as co
from singa_b200 import cpromg as cp
from singa_b200 import synthetic as syn
from tests.parity_util import digest_err, load_golden, rel_err


class _StubCSR:
    def __init__(self, edge_index, n):
        self.row, self.col, self.n = edge_index[0], edge_index[1], n


@pytest.fixture
def oracle_edge_ops(monkeypatch):
    monkeypatch.setattr(cp, "_build_centre_csr", lambda ei, n: _StubCSR(ei, n))
    monkeypatch.setattr(cp, "segment_softmax", lambda logits, csr: co.scatter_softmax(logits, csr.row, csr.n))
    monkeypatch.setattr(cp, "segment_sum", lambda v, csr: co.scatter_sum(v, csr.row, csr.n))
    monkeypatch.setattr(cp, "gather_rows", lambda h, csr, side: h[csr.col if side == "src" else csr.row])
    monkeypatch.setattr(cp, "edge_linear", lambda x, lin: F.linear(x, lin.weight, lin.bias))
    import singa_b200.radius as radius
    monkeypatch.setattr(radius, "knn_graph", lambda x, k, batch=None, flow="source_to_target", **kw:
                        co.knn_graph(x, k, batch=batch, flow=flow))
    cp._CSR_CACHE.clear()
    yield
    cp._CSR_CACHE.clear()


def _build(fx, device="cpu"):
    cfg = cp_cfg(fx)
    torch.manual_seed(0)
    model = cp.Transformer(cfg, 784, num_props=3, device=device).eval()
    syn.load_portable_parameters_fanin(model, fx["param_seed"])
    return model


def grad_errors(params, fx):
    """Digest errors of every parameter gradient.  weight_k_lin.bias shifts all logits of a centre by the same amount
    (q_i . b), which the per-centre softmax cancels: its gradient is zero up to rounding noise in the reference too,
    so it is checked to be noise (<= 1e-6), not compared digit for digit."""
    out = {}
    for n, d in fx["d_params"].items():
        if n.endswith("weight_k_lin.bias") or n.endswith("W_K.bias"):   # key biases: cancelled by the softmax
            ref_max = float(d["full"].abs().max()) if "full" in d else float(d["sample"].abs().max())
            assert ref_max < 1e-5, (n, ref_max)
            out[n] = 0.0 if float(params[n].grad.abs().max()) < 1e-4 else 1.0
        else:
            out[n] = digest_err(params[n].grad, d)
    return out


def cp_cfg(fx):
    from singa_b200.gan import _Cfg
    cfg = _Cfg(fx["cfg"])
    cfg.decoder["vocab_size"] = fx["vocab_size"]
    cfg.decoder["pad_index"] = fx["pad_index"]
    return cfg


def test_state_dict_surface_matches_reference_transformer():
    fx = load_golden("cpromg_transformer.pt")
    sd = _build(fx).state_dict()
    assert set(sd) == set(fx["surface"]), set(sd) ^ set(fx["surface"])
    for k, (shape, dtype) in fx["surface"].items():
        assert list(sd[k].shape) == shape and str(sd[k].dtype) == dtype, k


def test_transformer_wiring_matches_reference_golden(oracle_edge_ops):
    fx = load_golden("cpromg_transformer.pt")
    model = _build(fx)
    inp = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in fx["inputs"].items()}
    for k in ("node_attr", "aa_node_attr"):
        inp[k].requires_grad_(True)
    out = model(**inp)
    assert rel_err(out, fx["out"]) < 2e-5, rel_err(out, fx["out"])
    (out * fx["loss_weights"]).sum().backward()
    assert rel_err(inp["node_attr"].grad, fx["d_node_attr"]) < 2e-4
    assert rel_err(inp["aa_node_attr"].grad, fx["d_aa_node_attr"]) < 2e-4
    params = dict(model.named_parameters())
    bad = grad_errors(params, fx)
    bad = {n: e for n, e in bad.items() if e > 5e-4}
    assert not bad, bad
    assert sorted(n for n, p in params.items() if p.grad is None) == sorted(fx["params_without_grad"])


def test_graph_assembly_matches_the_loop_restatements():
    """to_undirected(mean) / to_dense_batch: sort-based device code vs the loop restatements of the pyg utilities."""
    rs = np.random.RandomState(3)
    n = 23
    pos = torch.from_numpy(rs.uniform(-4, 4, size=(n, 3)).astype(np.float32))
    batch = torch.from_numpy(np.repeat([0, 1, 2], [9, 1, 13]))
    di = co.knn_graph(pos, 4, batch=batch, flow="target_to_source")
    length = (pos[di[0]] - pos[di[1]]).norm(dim=1)
    ei_a, at_a = cp.to_undirected(di, length, n, reduce="mean")
    ei_b, at_b = co.to_undirected(di, length, reduce="mean")
    assert torch.equal(ei_a, ei_b) and torch.allclose(at_a, at_b, rtol=0, atol=1e-7)
    x = torch.from_numpy(rs.normal(size=(n, 5)).astype(np.float32))
    da, ma = cp.to_dense_batch(x, batch, 3, 13)
    db, mb = co.to_dense_batch(x, batch)
    assert torch.equal(da, db) and torch.equal(ma, mb)


def test_laplacian_pe_is_an_eigenbasis_and_is_cached():
    """dgl.lap_pe restated: columns are eigenvectors 1..k of I - D^-1/2 A D^-1/2 (up to sign), orthonormal; the cache
    computes a complex once."""
    g = syn.make_complex(3, n_protein=60, n_ligand=12)
    ei = g["ei_pp"].numpy()
    n = 60
    cache = cp.LapPECache(8)
    pe = cache.get(("c", 3), ei, n)
    assert pe.shape == (n, 8) and cache.misses == 1
    assert cache.get(("c", 3), ei, n) is pe and cache.hits == 1
    A = np.zeros((n, n)); A[ei[0], ei[1]] = 1          # dgl adj_external: rows = source
    assert np.array_equal(A, A.T)                      # this radius graph is symmetric: orientation-free check
    d = np.clip(A.sum(0), 1, None) ** -0.5
    L = np.eye(n) - d[:, None] * A * d[None, :]
    val = np.sort(np.linalg.eigvalsh(L))
    for j in range(8):
        v = pe[:, j].astype(np.float64)
        lam = v @ L @ v / (v @ v)
        assert abs(lam - val[j + 1]) < 1e-5, (j, lam, val[j + 1])
    assert np.allclose(pe.T @ pe, np.eye(8), atol=1e-4)
    small = cp.laplacian_pe(np.array([[0, 1], [1, 0]]), 2, 8)       # fewer nodes than k: zero padded
    assert small.shape == (2, 8) and np.all(small[:, 1:] == 0)
    # the cache holds the eigenvectors WITHOUT the random signs; a signed draw only flips whole columns
    signed = cp.laplacian_pe(ei, n, 8, np.random.RandomState(5))
    flips = (signed * pe).sum(0)
    assert np.allclose(np.abs(signed), np.abs(pe), atol=1e-6) and np.allclose(np.abs(flips), 1.0, atol=1e-4)
    assert (flips < 0).any() and (flips > 0).any()


def test_laplacian_pe_orientation_follows_dgl_on_an_asymmetric_graph():
    """dgl.lap_pe builds L from adj_external() (rows = source) and in-degrees.  On a directed 4-node graph the
    columns must be right eigenvectors of THAT matrix, not of its transpose (round-2 advisor finding)."""
    ei = np.array([[0, 0, 1, 2, 3, 3], [1, 2, 2, 3, 0, 1]])
    n = 4
    A = np.zeros((n, n)); A[ei[0], ei[1]] = 1
    d = np.clip(A.sum(0), 1, None) ** -0.5
    L = np.eye(n) - d[:, None] * A * d[None, :]
    val, vec = np.linalg.eig(L)
    order = np.argsort(val.real, kind="stable")
    pe = cp.laplacian_pe(ei, n, 2, random_sign=False)
    assert np.allclose(pe, np.real(vec[:, order[1:3]]).astype(np.float32), atol=1e-6)
    assert not np.allclose(np.abs(pe), np.abs(np.real(np.linalg.eig(L.T)[1][:, np.argsort(np.linalg.eig(L.T)[0].real, kind="stable")[1:3]])), atol=1e-3)


def test_collation_draws_fresh_eigenvector_signs_per_batch_when_asked():
    from singa_b200 import dataset as ds
    recs = ds.synthetic_records([1, 2], n_protein=30)
    plain = ds.collate_records(recs, pin=False)
    assert np.array_equal(plain.arena.views(plain.host)["lap_p"].numpy(), np.concatenate([r["lap_p"] for r in recs]))
    rng = np.random.RandomState(3)
    seen = set()
    for _ in range(6):
        b = ds.collate_records(recs, pin=False, lap_sign_rng=rng)
        lap = b.arena.views(b.host)["lap_p"].numpy()
        n0 = len(recs[0]["lap_p"])
        for part, rec in ((lap[:n0], recs[0]), (lap[n0:], recs[1])):
            assert np.array_equal(np.abs(part), np.abs(rec["lap_p"]))
            col = np.argmax(np.abs(rec["lap_p"]), axis=0)
            seen.add(tuple(np.sign(part[col, range(8)] * rec["lap_p"][col, range(8)]).astype(int)))
    assert len(seen) > 3      # signs differ between complexes and between batches
    import pytest
    with pytest.raises(ValueError):
        ds.pack_complex(syn.make_complex(1, n_protein=20, n_ligand=6), lap_cache=cp.LapPECache(8))


def test_lap_pe_reference_signature_uses_batch_encodings_or_caches_per_complex():
    g1, g2 = syn.make_complex(1, n_protein=30, n_ligand=8), syn.make_complex(2, n_protein=25, n_ligand=9)
    b = syn.collate([g1, g2])
    view = syn.GraphView(b)
    pe = cp.lap_pe(view, "protein_atoms")
    assert pe.shape == (55, 8)
    one = cp.laplacian_pe(g2["ei_pp"].numpy(), 25, 8)
    # per complex (not the batched graph), up to the cached random signs
    assert np.allclose(np.abs(pe[30:].numpy()), np.abs(one), atol=1e-5)
    misses = cp._LAP_CACHE.misses
    again = cp.lap_pe(view, "protein_atoms")     # same cached eigenvectors, fresh random signs (as dgl per call)
    assert torch.equal(again.abs(), pe.abs()) and cp._LAP_CACHE.misses == misses
    carried = dict(b, lap_p=torch.zeros(55, 8))
    assert cp.lap_pe(syn.GraphView(carried), "protein_atoms") is carried["lap_p"]


def test_centre_csr_cache_is_bounded_and_evicts_oldest_first():
    """Every training step registers the batch's two prebuilt encoder graphs; the cache must not grow with the steps
    (each entry pins an edge list and its CSR in device memory)."""
    from singa_b200 import cpromg as cp
    saved = dict(cp._CSR_CACHE)
    cp._CSR_CACHE.clear()
    try:
        eis = [torch.zeros(2, 3, dtype=torch.int64) for _ in range(3 * cp._CSR_CACHE_SLOTS)]
        for i, ei in enumerate(eis):
            cp._csr_cache_put((id(ei), ei._version, 5), ei, f"csr{i}")
            assert len(cp._CSR_CACHE) <= cp._CSR_CACHE_SLOTS
        kept = [v[1] for v in cp._CSR_CACHE.values()]
        assert kept == [f"csr{i}" for i in range(len(eis) - cp._CSR_CACHE_SLOTS, len(eis))]
        last = eis[-1]
        assert cp._CSR_CACHE[(id(last), last._version, 5)][0] is last
    finally:
        cp._CSR_CACHE.clear()
        cp._CSR_CACHE.update(saved)
