"""Pins oracle/ef_oracle.py (the travelling CPU restatement): against the reference-generated goldens everywhere,
and against the unmodified reference itself where /root/reference exists."""
import numpy as np
import pytest
import torch

from oracle import ef_oracle as eo
from oracle import ref_loader as rl
from singa_b200 import synthetic as syn
from singa_b200.embedding import DEFAULT_CONFIG
from tests.parity_util import LIGAND, PROTEIN, build_model, digest_err, load_golden, oracle_tables, rel_err

HP = dict(DEFAULT_CONFIG)


def _params(seed):
    model = build_model("cpu", seed)
    return {k: v.detach().clone() for k, v in model.state_dict().items()}, model


def test_wigner_rows_match_golden():
    fx = load_golden("wigner_small.pt")
    t = oracle_tables()
    rot = eo.edge_rot_mat(fx["edge_vec"], fx["rand"])
    assert rel_err(rot, fx["rot"]) < 1e-6
    W = eo.wigner_dense(fx["rot"], t)
    assert rel_err(W[:, t.keep, :], fx["rows_l6_m2"]) < 5e-6


@pytest.mark.parametrize("name", ["homo", "hetero"])
def test_block_and_edge_degree_match_golden(name):
    fx = load_golden("attention_small.pt")
    case = fx["cases"][name]
    p, model = _params(fx["param_seed"])
    p = {k: v.requires_grad_(v.dtype.is_floating_point) for k, v in p.items()}
    t = oracle_tables()
    W = eo.wigner_dense(case["rot"], t)
    x_s = case["x_src"].clone().requires_grad_(True)
    x_d = x_s if name == "homo" else case["x_dst"].clone().requires_grad_(True)
    x_edge = case["x_edge"].clone().requires_grad_(True)
    out, _, _ = eo.trans_block(p, fx["block"], x_s, x_d, x_edge, case["edge_index"], W, t, HP)
    ede = eo.edge_degree_embedding(p, "edge_degree_embedding.", x_edge, case["edge_index"], W, t, 16,
                                   case["n_dst"], eo._AVG_DEGREE)
    assert rel_err(out, case["block_out"]) < 2e-6
    assert rel_err(ede, case["edge_degree_out"]) < 2e-6
    seed = case["seed"]
    R = torch.from_numpy(np.random.RandomState(seed + 300).normal(size=tuple(out.shape)).astype(np.float32))
    R2 = torch.from_numpy(np.random.RandomState(seed + 301).normal(size=tuple(ede.shape)).astype(np.float32))
    ((out * R).sum() + (ede * R2).sum()).backward()
    assert rel_err(x_edge.grad, case["grad_x_edge"]) < 2e-5
    assert rel_err(x_s.grad, case["grad_x_src"]) < 2e-5
    if name != "homo":
        assert rel_err(x_d.grad, case["grad_x_dst"]) < 2e-5
    for n, dig in case["param_grads"].items():
        assert p[n].grad is not None, n
        assert digest_err(p[n].grad, dig) < 5e-5, n


def _frames_from_golden(fx):
    return fx["frames"]


def test_embedding_forward_matches_golden_example():
    fx = load_golden("example_3wi2.pt")
    p, _ = _params(fx["param_seed"])
    t = oracle_tables()
    with torch.no_grad():
        out = eo.embedding_forward(p, fx["graph"], fx["frames"], t, HP)
    for k, ref in fx["outputs"].items():
        assert rel_err(out[k], ref) < 5e-6, k


def test_dead_block_skip_is_exact():
    fx = load_golden("example_3wi2.pt")
    p, _ = _params(fx["param_seed"])
    t = oracle_tables()
    with torch.no_grad():
        a = eo.embedding_forward(p, fx["graph"], fx["frames"], t, HP, compute_dead_blocks=True)
        b = eo.embedding_forward(p, fx["graph"], fx["frames"], t, HP, compute_dead_blocks=False)
    for k in a:
        assert torch.equal(a[k], b[k]), k


@pytest.mark.skipif(not rl.reference_available(), reason="needs /root/reference (build container only)")
def test_oracle_matches_live_reference_on_synthetic_complex():
    """Independent of the goldens: unmodified reference vs oracle on a fresh synthetic complex."""
    ef, emb = rl.load_reference()
    torch.manual_seed(0)
    ref = emb.EquivariantEmbedding(rl.EasyDict(rl.DEFAULT_EMBEDDING_CONFIG), device="cpu")
    syn.load_portable_parameters(ref, 5)
    g = syn.make_complex(3, n_protein=60, n_ligand=12, cutoff=5.0)
    rnd = syn.frame_randoms(g, 9)
    draws = iter([rnd["pp"], rnd["ll"], rnd["lp"]])
    orig = torch.rand_like
    torch.rand_like = lambda x, **k: next(draws)
    try:
        with torch.no_grad():
            out = ref(syn.GraphView(g))
    finally:
        torch.rand_like = orig
    frames = {}
    for key, (s, d, ei) in dict(pp=("pos_p", "pos_p", "ei_pp"), ll=("pos_l", "pos_l", "ei_ll"),
                                lp=("pos_l", "pos_p", "ei_lp")).items():
        frames[key] = eo.edge_rot_mat(g[s][g[ei][0]] - g[d][g[ei][1]], rnd[key])
    p = {k: v.detach() for k, v in ref.state_dict().items()}
    G62, G66 = ref.SO3_grid[6][2], ref.SO3_grid[6][6]
    jd = torch.load(rl.REFERENCE_ROOT + "/model/Jd.pt", weights_only=False)
    t = eo.Tables(6, 2, jd, G62.to_grid_mat, G62.from_grid_mat, G66.to_grid_mat, G66.from_grid_mat)
    with torch.no_grad():
        o2 = eo.embedding_forward(p, g, frames, t, HP)
    for k in o2:
        assert rel_err(o2[k], out[k].embedding) < 5e-6, k


def test_radius_graph_oracle_neighbour_cap_follows_upstream_order():
    """torch-cluster 1.6.x radius_graph(loop=False) searches with max_num_neighbors + 1 INCLUDING the self hit, in
    ascending source index, then removes self loops: a target whose own index follows its first cap+1
    hits keeps cap+1 sources, the others min(deg, cap).  Hand-computed case: 6 collinear points, all within r."""
    from oracle import ef_oracle as eo
    p = torch.tensor([[float(i) * 0.1, 0.0, 0.0] for i in range(6)])
    ei = eo.radius_graph_bruteforce(p, 10.0, max_num_neighbors=2)
    rows = {t: ei[0][ei[1] == t].tolist() for t in range(6)}
    assert rows == {0: [1, 2], 1: [0, 2], 2: [0, 1], 3: [0, 1, 2], 4: [0, 1, 2], 5: [0, 1, 2]}
    eil = eo.radius_graph_bruteforce(p, 10.0, loop=True, max_num_neighbors=2)
    assert all(eil[0][eil[1] == t].tolist() == [0, 1] for t in range(6))
