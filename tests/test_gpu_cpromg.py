"""GPU parity of the CProMG side: knn_graph kernel, the message-passing MultiHeadAttention on the
deterministic segment kernels, the whole Transformer and a SINGA train step, against the reference-written goldens
(cpromg_attention.pt: the unmodified reference MultiHeadAttention; cpromg_transformer.pt: the unmodified reference
Transformer) and the brute-force oracle.  fp32 mode: outputs <= 1e-5 (2e-5 through the 6+6+6-layer transformer),
gradients <= 1e-4 (5e-4); bf16 tcgen05 mode: stated looser bounds 2e-2 / 6e-2."""
import numpy as np
import pytest
import torch

from tests.parity_util import digest_err, load_golden, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


# ------------------------------------------------------------------------------------------------ knn_graph
def _knn_cases():
    rs = np.random.RandomState(0)
    f = lambda a: torch.from_numpy(np.asarray(a, dtype=np.float32))
    yield "cloud k48", f(rs.uniform(-14, 14, size=(700, 3))), None, 48
    yield "cloud k1", f(rs.uniform(-5, 5, size=(200, 3))), None, 1
    yield "cloud k64", f(rs.uniform(-5, 5, size=(300, 3))), None, 64
    sizes = [40, 1, 300, 7, 2, 77]
    pos = torch.cat([f(rs.uniform(-9, 9, size=(s, 3))) + 300.0 * i for i, s in enumerate(sizes)])
    batch = torch.cat([torch.full((s,), i, dtype=torch.int64) for i, s in enumerate(sizes)])
    yield "batches smaller than k", pos, batch, 30
    g = np.stack(np.meshgrid(np.arange(7), np.arange(7), np.arange(5), indexing="ij"), -1).reshape(-1, 3)
    yield "lattice (ties)", f(g * 1.5), None, 12
    yield "duplicates", f(np.repeat(rs.uniform(-3, 3, size=(50, 3)), 3, axis=0)), None, 8
    yield "line", f(np.stack([np.linspace(0, 100, 400), np.zeros(400), np.zeros(400)], 1)), None, 6


@pytest.mark.parametrize("name,pos,batch,k", list(_knn_cases()), ids=[c[0] for c in _knn_cases()])
def test_knn_graph_kernel_matches_bruteforce_edge_list(name, pos, batch, k):
    from oracle import cpromg_oracle as co
    from singa_b200.radius import knn_graph
    ref = co.knn_graph_bruteforce(pos, k, batch)
    got = knn_graph(pos.to(DEV), k, batch=None if batch is None else batch.to(DEV), flow="target_to_source").cpu()
    assert got.shape == ref.shape, (got.shape, ref.shape)
    assert torch.equal(got, ref)
    s2t = knn_graph(pos.to(DEV), k, batch=None if batch is None else batch.to(DEV)).cpu()
    assert torch.equal(s2t, ref.flip(0))


def test_knn_graph_large_cloud_properties():
    """200 k points: every centre has exactly k neighbours, rows are sorted by distance, and sampled rows equal the
    brute force over ALL points."""
    from singa_b200.radius import knn_graph
    rs = np.random.RandomState(4)
    n, k = 200_000, 48
    pos = torch.from_numpy(rs.uniform(0, 164.0, size=(n, 3)).astype(np.float32))
    ei = knn_graph(pos.to(DEV), k, flow="target_to_source").cpu()
    assert ei.shape == (2, n * k)
    assert torch.equal(ei[0], torch.arange(n).repeat_interleave(k))
    d2 = ((pos[ei[0]] - pos[ei[1]]) ** 2).sum(1).view(n, k)
    assert bool((d2[:, 1:] >= d2[:, :-1] - 1e-4).all())
    for i in rs.choice(n, size=20, replace=False):
        d = pos - pos[i]
        full = ((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2])
        full[i] = float("inf")
        want = torch.argsort(full, stable=True)[:k]
        assert torch.equal(ei[1].view(n, k)[i], want), i


# ------------------------------------------------------------------------------------------------ segment ops
def test_segment_ops_match_scatter_restatements():
    from oracle import cpromg_oracle as co
    from singa_b200 import cpromg as cp
    rs = np.random.RandomState(2)
    n, e = 37, 400
    row = torch.from_numpy(rs.randint(0, n - 4, size=e))         # last 4 centres have no edge
    col = torch.from_numpy(rs.randint(0, n, size=e))
    ei = torch.stack((row, col)).to(DEV)
    csr = cp.centre_csr(ei, n)
    logits = torch.from_numpy(rs.normal(size=(e, 4)).astype(np.float32) * 3).to(DEV).requires_grad_(True)
    vals = torch.from_numpy(rs.normal(size=(e, 24)).astype(np.float32)).to(DEV).requires_grad_(True)
    h = torch.from_numpy(rs.normal(size=(n, 8)).astype(np.float32)).to(DEV).requires_grad_(True)
    a = cp.segment_softmax(logits, csr)
    s = cp.segment_sum(vals, csr)
    gs, gd = cp.gather_rows(h, csr, "src"), cp.gather_rows(h, csr, "dst")
    w = [torch.from_numpy(rs.normal(size=tuple(t.shape)).astype(np.float32)).to(DEV) for t in (a, s, gs, gd)]
    ((a * w[0]).sum() + (s * w[1]).sum() + (gs * w[2]).sum() + (gd * w[3]).sum()).backward()
    lc, vc, hc = (t.detach().cpu().requires_grad_(True) for t in (logits, vals, h))
    a2, s2 = co.scatter_softmax(lc, row, n), co.scatter_sum(vc, row, n)
    ((a2 * w[0].cpu()).sum() + (s2 * w[1].cpu()).sum() + (hc[col] * w[2].cpu()).sum() + (hc[row] * w[3].cpu()).sum()).backward()
    assert rel_err(a, a2) < 1e-6 and rel_err(s, s2) < 1e-6
    assert torch.equal(gs.cpu(), hc[col].detach()) and torch.equal(gd.cpu(), hc[row].detach())
    assert float(s[-4:].abs().max()) == 0.0
    assert rel_err(logits.grad, lc.grad) < 1e-5 and rel_err(vals.grad, vc.grad) < 1e-6 and rel_err(h.grad, hc.grad) < 1e-6


# ------------------------------------------------------------------------------------------------ attention
@pytest.mark.parametrize("precision,tol_out,tol_grad", [("fp32", 1e-5, 1e-4), ("bf16", 2e-2, 6e-2)])
def test_multi_head_attention_matches_reference_golden(precision, tol_out, tol_grad):
    from singa_b200 import cpromg as cp
    from singa_b200 import ops
    fx = load_golden("cpromg_attention.pt")
    d = fx["dims"]
    ops.set_precision(precision)
    try:
        mha = cp.MultiHeadAttention(d["hidden"], d["edge"], d["key"], d["heads"], device=DEV)
        mha.load_state_dict(fx["state_dict"], strict=True)
        node = fx["node_attr"].to(DEV).requires_grad_(True)
        eattr = fx["edge_attr"].to(DEV).requires_grad_(True)
        out = mha(node, fx["edge_index"].to(DEV), eattr)
        (out * fx["loss_weights"].to(DEV)).sum().backward()
    finally:
        ops.set_precision("fp32")
    errs = {"out": rel_err(out, fx["out"]), "d_node_attr": rel_err(node.grad, fx["d_node_attr"]),
            "d_edge_attr": rel_err(eattr.grad, fx["d_edge_attr"])}
    for n, p in mha.named_parameters():
        ref = fx["d_params"][n]
        if n == "weight_k_lin.bias":     # cancelled by the per-centre softmax: noise in the reference too
            assert float(p.grad.abs().max()) < 1e-4 * max(1.0, tol_grad / 1e-4)
            continue
        errs["grad:" + n] = rel_err(p.grad, ref)
    bad = {k: v for k, v in errs.items() if v > (tol_out if k == "out" else tol_grad)}
    assert not bad, bad


def test_attention_is_bitwise_reproducible():
    from singa_b200 import cpromg as cp
    fx = load_golden("cpromg_attention.pt")
    d = fx["dims"]
    mha = cp.MultiHeadAttention(d["hidden"], d["edge"], d["key"], d["heads"], device=DEV)
    mha.load_state_dict(fx["state_dict"])
    ei = fx["edge_index"].to(DEV)
    runs = []
    for _ in range(2):
        mha.zero_grad(set_to_none=True)
        node = fx["node_attr"].to(DEV).requires_grad_(True)
        out = mha(node, ei, fx["edge_attr"].to(DEV))
        out.square().sum().backward()
        runs.append((out.detach().clone(), node.grad.clone(), {n: p.grad.clone() for n, p in mha.named_parameters()}))
    assert torch.equal(runs[0][0], runs[1][0]) and torch.equal(runs[0][1], runs[1][1])
    for n in runs[0][2]:
        assert torch.equal(runs[0][2][n], runs[1][2][n]), n


# ------------------------------------------------------------------------------------------------ transformer
def test_transformer_on_cuda_matches_reference_golden():
    from singa_b200 import cpromg as cp
    from singa_b200 import synthetic as syn
    from tests.test_cpromg_host import cp_cfg, grad_errors
    fx = load_golden("cpromg_transformer.pt")
    torch.manual_seed(0)
    model = cp.Transformer(cp_cfg(fx), 784, num_props=3, device=DEV).eval()
    syn.load_portable_parameters_fanin(model, fx["param_seed"])
    inp = {k: (v.to(DEV) if torch.is_tensor(v) else v) for k, v in fx["inputs"].items()}
    for k in ("node_attr", "aa_node_attr"):
        inp[k].requires_grad_(True)
    out = model(**inp)
    assert rel_err(out, fx["out"]) < 2e-5, rel_err(out, fx["out"])
    (out * fx["loss_weights"].to(DEV)).sum().backward()
    assert rel_err(inp["node_attr"].grad, fx["d_node_attr"]) < 2e-4
    assert rel_err(inp["aa_node_attr"].grad, fx["d_aa_node_attr"]) < 2e-4
    bad = {n: e for n, e in grad_errors(dict(model.named_parameters()), fx).items() if e > 5e-4}
    assert not bad, bad


# ------------------------------------------------------------------------------------------------ dataset + SINGA
def test_prebuilt_csr_batch_equals_the_sorting_path():
    """Dataset format (8f-4): a batch collated from CSR-carrying records, with its CSRs registered (no sort, no
    device->host read), gives bit-identical embeddings to the same complexes through the sorting path."""
    from singa_b200 import EF_layers as ef
    from singa_b200 import dataset as ds
    from singa_b200 import synthetic as syn
    from tests.parity_util import build_model
    seeds = [3, 4, 5]
    graphs = [syn.make_complex(s, n_protein=70, n_ligand=12) for s in seeds]
    recs = [ds.pack_complex(g, key=s) for g, s in zip(graphs, seeds)]
    model = build_model(DEV, 1, "fp32")
    hb = ds.collate_records(recs)
    g1 = hb.to(DEV).graph()
    big = {k: v.to(DEV) for k, v in syn.collate(graphs).items()}
    frames = {}
    torch.manual_seed(3)
    for key, (sp, dp, ei) in dict(pp=("pos_p", "pos_p", "ei_pp"), ll=("pos_l", "pos_l", "ei_ll"),
                                  lp=("pos_l", "pos_p", "ei_lp")).items():
        e = big[ei]
        frames[key] = ef.init_edge_rot_mat(big[sp][e[0]] - big[dp][e[1]], device=DEV)
    with torch.no_grad():
        a = model(syn.GraphView(big), frames=frames)
        # the record stores the pp / ll edges destination-sorted: permute their frames the same way
        perm = {k: torch.cat([torch.from_numpy(r[f"{k}_orig"].astype(np.int64)) + off for r, off in
                              zip(recs, np.cumsum([0] + [len(r[f"{k}_src"]) for r in recs])[:-1])]).to(DEV)
                for k in ("pp", "ll")}
        b = model(syn.GraphView(g1), frames=dict(pp=frames["pp"][perm["pp"]], ll=frames["ll"][perm["ll"]],
                                                 lp=frames["lp"]))
    for k in ("protein_atoms", "ligand_atoms"):
        assert torch.equal(a[k].embedding, b[k].embedding), k


def test_prebuilt_encoder_graphs_give_the_same_transformer_output():
    """Encoder graphs carried by the dataset batch (host-built once) against the on-the-fly device construction
    (knn kernel + sort-based to_undirected + get_laplacian): identical edge lists, identical logits."""
    from singa_b200 import cpromg as cp
    from singa_b200 import dataset as ds
    from singa_b200 import synthetic as syn
    from singa_b200.embedding import batch_vector_from_ptr
    from singa_b200.gan import default_config
    recs = ds.synthetic_records(range(3), n_protein=70, tgt_len=12)
    hb = ds.collate_records(recs)
    g = hb.to(DEV).graph()
    cfg = default_config().model
    cfg.decoder["tgt_len"] = 12
    torch.manual_seed(0)
    model = cp.Transformer(cfg, 784, num_props=3, device=DEV).eval()
    syn.load_portable_parameters_fanin(model, 2)
    rs = np.random.RandomState(0)
    f = lambda n: torch.from_numpy(rs.normal(size=(n, 784)).astype(np.float32) * 0.3).to(DEV)
    n_p, n_l = g["pos_p"].shape[0], g["pos_l"].shape[0]
    args = dict(node_attr=f(n_p), pos=g["pos_p"], batch=batch_vector_from_ptr(g["ptr_p"], n_p), atom_laplacian=g["lap_p"],
                smiles_index=g["smi_in"], tgt_len=12, aa_node_attr=f(n_l), aa_pos=g["pos_l"],
                aa_batch=batch_vector_from_ptr(g["ptr_l"], n_l), aa_laplacian=g["lap_l"],
                prop=torch.ones(3, 3, device=DEV),
                sizes=dict(batch_size=3, max_atoms=hb.meta["max_p"], max_aa=hb.meta["max_l"]))
    with torch.no_grad():
        a = model(**args)
        b = model(**args, graphs=g["enc_graphs"])
        for role, (pos, bt, k) in dict(atom=(g["pos_p"], args["batch"], cfg.encoder.knn),
                                       aa=(g["pos_l"], args["aa_batch"], 30)).items():
            ei, _ = model.encoder._graph(pos, bt, k)
            assert torch.equal(ei, g["enc_graphs"][role].edge_index), role
    assert torch.equal(a, b)


def test_singa_train_step_runs_and_is_reproducible():
    """BASELINE.json config 3 in miniature: SINGA (EquivariantEmbedding + CProMG transformer) through Trainer.step on
    dataset-format batches; finite loss, every trainable parameter that can receive a gradient gets one, and two
    trainers from the same seed produce bit-identical losses (no float atomics on the step)."""
    from singa_b200 import dataset as ds
    from singa_b200 import ops
    from singa_b200 import synthetic as syn
    from singa_b200.gan import SINGA, default_config
    from singa_b200.train import Trainer
    recs = ds.synthetic_records(range(4), n_protein=60)
    losses = []
    try:
        ops.set_precision("fp32")
        for _ in range(2):
            torch.manual_seed(0)
            cfg = default_config()
            model = SINGA(cfg, device=DEV)
            syn.load_portable_parameters(model.embedding, 0)
            syn.load_portable_parameters_fanin(model.model, 1)
            model.model.decoder.pos_emb.dropout.p = 0.0
            tr = Trainer(model, ds.ComplexDataset(recs), batch_size_per_rank=2, seed=5)
            out = [tr.step() for _ in range(2)]
            losses.append(torch.stack([l for l, _ in out]).cpu())
            assert bool(torch.isfinite(losses[-1]).all())
    finally:
        ops.set_precision("fp32")
    assert torch.equal(losses[0], losses[1]), losses
    assert float(losses[0][0]) > 1.0      # ~ln(116) at initialisation
