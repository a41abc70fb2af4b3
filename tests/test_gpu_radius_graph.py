"""radius_graph cell-list kernel (north star subsystem (a), BASELINE.json config 4) against the brute-force oracle:
the edge LIST is bit-exact (targets ascending, sources ascending), including ties on the cutoff, duplicates, the
neighbour cap, batches and self loops; at 1 M points: size-independent properties + sampled rows vs brute force."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _cloud(n, side, seed):
    return torch.from_numpy(np.random.RandomState(seed).uniform(0, side, size=(n, 3)).astype(np.float32))


@pytest.mark.parametrize("n,side,r", [(1, 1.0, 5.0), (2, 1.0, 5.0), (2, 30.0, 5.0), (257, 20.0, 5.0), (3000, 40.0, 8.0),
                                      (3000, 40.0, 10.0), (500, 3.0, 8.0)])
def test_edge_list_bit_exact(n, side, r):
    from oracle import ef_oracle as eo
    from singa_b200.radius import radius_graph_csr
    p = _cloud(n, side, n)
    ref = eo.radius_graph_bruteforce(p, r, max_num_neighbors=1 << 30)
    got = radius_graph_csr(p.to(DEV), r, max_num_neighbors=1 << 30)
    assert torch.equal(got.edge_index.cpu(), ref)
    assert not got.truncated
    deg = torch.bincount(ref[1], minlength=n)
    assert torch.equal(got.rowptr.cpu().long(), torch.cat((torch.zeros(1, dtype=torch.long), deg.cumsum(0))))
    assert torch.equal(got.col.cpu().long(), ref[0])
    assert got.max_degree == (int(deg.max()) if n else 0)


def test_ties_on_cutoff_duplicates_and_negative_coordinates():
    """Integer lattice: many pairs at exactly r (strict <, excluded) and just below; duplicated points at distance 0."""
    from oracle import ef_oracle as eo
    from singa_b200.radius import radius_graph
    g = torch.arange(-4, 5, dtype=torch.float32)
    p = torch.stack(torch.meshgrid(g, g, g, indexing="ij"), dim=-1).reshape(-1, 3)
    p = torch.cat((p, p[:50]))  # duplicates
    p = p[torch.from_numpy(np.random.RandomState(0).permutation(len(p)))]
    for r in (1.0, 2.0, 3.0, float(np.sqrt(np.float32(5.0)))):
        ref = eo.radius_graph_bruteforce(p, r, max_num_neighbors=1 << 30)
        assert torch.equal(radius_graph(p.to(DEV), r, max_num_neighbors=1 << 30).cpu(), ref), r


def test_batches_cap_loop_and_flow():
    from oracle import ef_oracle as eo
    from singa_b200.radius import radius_graph, radius_graph_csr
    rs = np.random.RandomState(3)
    sizes = [40, 1, 300, 77]
    p = torch.cat([_cloud(s, 18.0, 10 + i) + torch.from_numpy(rs.uniform(-500, 500, size=3).astype(np.float32))
                   for i, s in enumerate(sizes)])
    batch = torch.cat([torch.full((s,), i, dtype=torch.int64) for i, s in enumerate(sizes)])
    ref = eo.radius_graph_bruteforce(p, 6.0, batch, max_num_neighbors=1 << 30)
    assert torch.equal(radius_graph(p.to(DEV), 6.0, batch.to(DEV), max_num_neighbors=1 << 30).cpu(), ref)
    # default cap of torch_cluster (32) truncates: smallest source indices are kept
    refc = eo.radius_graph_bruteforce(p, 6.0, batch, max_num_neighbors=32)
    got = radius_graph_csr(p.to(DEV), 6.0, batch.to(DEV))
    assert got.truncated and torch.equal(got.edge_index.cpu(), refc)
    assert torch.equal(got.col.cpu().long(), refc[0])
    # upstream cap order (self hit counted, dropped after the cap): small caps, both loop modes
    for cap in (1, 4, 7):
        for loop in (False, True):
            refk = eo.radius_graph_bruteforce(p, 6.0, batch, loop=loop, max_num_neighbors=cap)
            gotk = radius_graph(p.to(DEV), 6.0, batch.to(DEV), loop=loop, max_num_neighbors=cap).cpu()
            assert torch.equal(gotk, refk), (cap, loop)
    refl = eo.radius_graph_bruteforce(p, 6.0, batch, loop=True, max_num_neighbors=1 << 30)
    assert torch.equal(radius_graph(p.to(DEV), 6.0, batch.to(DEV), loop=True, max_num_neighbors=1 << 30).cpu(), refl)
    t2s = radius_graph(p.to(DEV), 6.0, batch.to(DEV), max_num_neighbors=1 << 30, flow="target_to_source").cpu()
    assert torch.equal(t2s, ref.flip(0))


def test_long_rows_use_the_block_sort():
    """> 1024 neighbours per point (dense blob): rows sorted by the CTA-wide bitonic kernel."""
    from oracle import ef_oracle as eo
    from singa_b200.radius import radius_graph_csr
    p = _cloud(2500, 4.0, 5)
    ref = eo.radius_graph_bruteforce(p, 8.0, max_num_neighbors=1 << 30)
    got = radius_graph_csr(p.to(DEV), 8.0, max_num_neighbors=1 << 30)
    assert got.max_degree > 1024
    assert torch.equal(got.edge_index.cpu(), ref)


def test_synthetic_complex_edges_match_generator():
    from singa_b200 import synthetic as syn
    from singa_b200.radius import radius_graph
    g = syn.make_complex(7)
    ei = radius_graph(g["pos_p"].to(DEV), 8.0, max_num_neighbors=1 << 30).cpu()
    assert torch.equal(ei, g["ei_pp"])


def test_million_point_cloud_properties():
    """BASELINE.json config 4: 1 M points at 0.045 / A^3, r = 8 A.  Symmetry, sortedness, expected mean degree and
    400 sampled rows against brute force over all points."""
    from singa_b200.radius import radius_graph_csr
    n = 1_000_000
    side = (n / 0.045) ** (1.0 / 3.0)
    gen = torch.Generator(device=DEV).manual_seed(0)
    p = torch.rand(n, 3, device=DEV, generator=gen) * side
    got = radius_graph_csr(p, 8.0, max_num_neighbors=1 << 30)
    ei = got.edge_index
    E = ei.shape[1]
    mean_deg = E / n
    assert 85.0 < mean_deg < 97.0, mean_deg  # 0.045 * 4/3 pi 8^3 = 96.5 minus boundary losses
    key = ei[1] * n + ei[0]
    assert bool((key[1:] > key[:-1]).all()), "targets ascending, sources strictly ascending inside a target"
    rev = torch.sort(ei[0] * n + ei[1]).values
    assert torch.equal(rev, key), "edge set is symmetric"
    rows = torch.from_numpy(np.random.RandomState(1).choice(n, size=400, replace=False)).to(DEV)
    rp = got.rowptr.long()
    for i in rows.tolist():
        d = p - p[i]
        d2 = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
        want = torch.nonzero(d2 < 64.0).flatten()
        want = want[want != i]
        assert torch.equal(ei[0, rp[i]:rp[i + 1]], want), i
