"""efl_csr_from_edge_index (csrc/csr.cu) against the numpy stable-sort checker: every array bit-exact, on ragged,
empty, single-destination, bipartite (num_src != num_dst) and large edge lists; out-of-range ids raise."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
FIELDS = ("perm", "rowptr", "src_sorted", "dst_sorted", "dst_orig", "src_ptr", "src_idx")


def _check(ei, ns, nd):
    from oracle import ef_oracle as eo
    from singa_b200 import ops
    ref = eo.csr_from_edge_index(ei, ns, nd)
    got = ops.build_csr(ei.to(DEV), ns, nd)
    for f in FIELDS:
        g = getattr(got, f)
        assert g.dtype == torch.int32
        assert np.array_equal(g.cpu().numpy().astype(np.int64), ref[f]), f
    assert got.num_edges == ei.shape[1]
    # chunks tile the destination range with whole destinations
    if nd:
        assert got.chunks[0][0] == 0 and got.chunks[-1][1] == nd
        assert all(a[1] == b[0] and a[3] == b[2] for a, b in zip(got.chunks, got.chunks[1:]))
        assert got.chunks[-1][3] == ei.shape[1]
    return got


@pytest.mark.parametrize("ns,nd,E,seed", [(1, 1, 1, 0), (5, 5, 0, 1), (7, 3, 40, 2), (3, 900, 2000, 3),
                                          (465, 465, 930, 4), (14400, 14400, 900_000, 5), (70000, 1, 1000, 6),
                                          (1, 70000, 1000, 7), (1 << 20, 1 << 20, 3_000_000, 8)])
def test_csr_bit_exact(ns, nd, E, seed):
    rs = np.random.RandomState(seed)
    ei = torch.from_numpy(np.stack((rs.randint(0, ns, size=E), rs.randint(0, nd, size=E))).astype(np.int64))
    _check(ei, ns, nd)


def test_skewed_degrees_and_duplicate_edges():
    rs = np.random.RandomState(11)
    nd = 300
    dst = np.minimum((rs.pareto(1.0, size=50_000)).astype(np.int64), nd - 1)   # a few very heavy destinations
    src = rs.randint(0, 50, size=dst.size)                                       # many exact duplicates
    _check(torch.from_numpy(np.stack((src, dst))), 50, nd)


def test_non_contiguous_and_int32_edge_index():
    rs = np.random.RandomState(12)
    big = torch.from_numpy(rs.randint(0, 100, size=(2, 2000)).astype(np.int64))
    _check(big[:, ::2], 100, 100)
    from singa_b200 import ops
    a = ops.build_csr(big.to(DEV), 100, 100)
    b = ops.build_csr(big.to(DEV).to(torch.int32), 100, 100)
    assert all(torch.equal(getattr(a, f), getattr(b, f)) for f in FIELDS)


def test_out_of_range_ids_raise():
    from singa_b200 import ops
    ei = torch.tensor([[0, 1, 2], [0, 5, 1]], dtype=torch.int64, device=DEV)
    with pytest.raises(IndexError):
        ops.build_csr(ei, 3, 3)
    ei = torch.tensor([[0, -1, 2], [0, 1, 1]], dtype=torch.int64, device=DEV)
    with pytest.raises(IndexError):
        ops.build_csr(ei, 3, 3)


def test_batched_build_matches_single_builds_and_fills_the_cache():
    """ops.build_csr_many / EF_layers.prefetch_edge_csr: one device->host read for several edge lists, same arrays and
    chunk plans as one build per list; a bad list in the batch still raises."""
    from singa_b200 import EF_layers as ef
    from singa_b200 import ops
    rs = np.random.RandomState(21)
    shapes = [(40, 40, 500), (6, 40, 90), (40, 6, 90), (9, 9, 0)]
    eis = [torch.from_numpy(np.stack((rs.randint(0, ns, size=E), rs.randint(0, nd, size=E))).astype(np.int64)).to(DEV)
           for ns, nd, E in shapes]
    specs = [(ei, ns, nd) for ei, (ns, nd, _) in zip(eis, shapes)]
    many = ops.build_csr_many(specs, max_chunk_edges=64)
    for csr, (ei, ns, nd) in zip(many, specs):
        one = ops.build_csr(ei, ns, nd, max_chunk_edges=64)
        assert all(torch.equal(getattr(csr, f), getattr(one, f)) for f in FIELDS)
        assert csr.chunks == one.chunks
    ef._CSR_CACHE.clear()
    ef.prefetch_edge_csr(specs + specs[:1])
    assert len(ef._CSR_CACHE) == len(specs)
    for ei, ns, nd in specs:
        hit = ef._edge_csr(ei, ns, nd)
        assert hit is ef._CSR_CACHE[ef._csr_key(ei, ns, nd)][1]
    bad = torch.tensor([[0, 1], [0, 7]], dtype=torch.int64, device=DEV)
    with pytest.raises(IndexError):
        ops.build_csr_many([specs[0], (bad, 3, 3)])
