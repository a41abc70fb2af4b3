"""Shared helpers for the parity tests (and __graft_entry__.smoke): run the CUDA path and the CPU oracle on the
same inputs.  The oracle (oracle/ef_oracle.py) is used here only as the checker."""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")

PROTEIN, LIGAND = "protein_atoms", "ligand_atoms"


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    """Per-tensor ||a-b||_inf / ||b||_inf (SURVEY.md 8c tolerance definition)."""
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    denom = float(b.abs().max())
    return float((a - b).abs().max()) / (denom if denom > 0 else 1.0)


def digest_err(grad: torch.Tensor, dig: dict) -> float:
    """Compare a gradient with a digest written by oracle/make_golden.py."""
    flat = grad.detach().reshape(-1).float().cpu()
    if "full" in dig:
        return rel_err(flat, dig["full"].reshape(-1))
    sample = flat[::dig["stride"]][:dig["sample"].numel()]
    scale = max(float(dig["sample"].abs().max()), dig["l2"] / max(flat.numel(), 1) ** 0.5, 1e-30)
    e1 = float((sample - dig["sample"]).abs().max()) / scale
    e2 = abs(float(flat.double().norm()) - dig["l2"]) / max(dig["l2"], 1e-30)
    return max(e1, e2)


def load_golden(name):
    return torch.load(os.path.join(GOLDEN, name), map_location="cpu", weights_only=False)


def build_model(device, param_seed, precision="fp32", compute_dead_blocks=False, lmax=6):
    from singa_b200 import ops, synthetic
    from singa_b200.embedding import DEFAULT_CONFIG, EquivariantEmbedding
    ops.set_precision(precision)
    torch.manual_seed(0)
    cfg = dict(DEFAULT_CONFIG)
    cfg["lmax_list"] = [lmax]
    model = EquivariantEmbedding(cfg, device=device, compute_dead_blocks=compute_dead_blocks)
    synthetic.load_portable_parameters(model, param_seed)
    return model


def oracle_tables(model=None):
    """Oracle tables; grid matrices come from the oracle-side shim formulas (ref_loader), J from the product's
    generator only when the reference's Jd.pt is not on this machine."""
    from oracle import ef_oracle as eo
    from oracle import ref_loader as rl
    if rl.reference_available():
        jd = torch.load(os.path.join(rl.REFERENCE_ROOT, "model", "Jd.pt"), weights_only=False)
    else:
        from singa_b200 import so3_constants as sc
        jd = [torch.tensor(sc.j_matrix(l)) for l in range(9)]
    ex = load_golden("example_3wi2.pt")["grid"]  # grid buffers exactly as the reference built them
    return eo.Tables(6, 2, jd, ex["to62"], ex["from62"], ex["to66"], ex["from66"])


def run_attention_case(name: str, device="cuda:0", precision="fp32", return_all=False, fixture="attention_small.pt",
                       model=None):
    """TransBlockV2 + EdgeDegreeEmbedding on tests/golden/attention_small.pt[name]: CUDA path vs the
    reference-generated golden.  Returns the max relative error over outputs and gradients."""
    from singa_b200 import EF_layers as ef
    fx = load_golden(fixture)
    case = fx["cases"][name]
    lmax = fx.get("lmax", 6)
    if model is None:
        model = build_model(device, fx["param_seed"], precision, lmax=lmax)
    else:   # a model built by the caller (smoke(): one construction, two precisions)
        from singa_b200 import ops
        ops.set_precision(precision)
        model.zero_grad(set_to_none=True)
    block = model.blocks[1]
    dev = torch.device(device)
    ei = case["edge_index"].to(dev)
    model.SO3_rotation[0].set_wigner(case["rot"].to(dev))
    x_edge = case["x_edge"].to(dev).requires_grad_(True)
    x_s = case["x_src"].to(dev).requires_grad_(True)
    homo = name == "homo"
    x_d = x_s if homo else case["x_dst"].to(dev).requires_grad_(True)
    seed = case["seed"]
    rs = np.random.RandomState(seed)
    z_s = torch.from_numpy(rs.choice([6, 7, 8, 16], size=x_s.shape[0])).to(dev)
    z_d = z_s if homo else torch.from_numpy(np.random.RandomState(seed + 1).choice([6, 7, 8, 16], size=x_d.shape[0])).to(dev)

    def emb(t):
        e = ef.SO3_Embedding(0, [lmax], 16, dtype=torch.float32, device=dev)
        e.set_embedding(t * 1.0)
        return e

    if homo:
        out = block(x=emb(x_s), atomic_numbers=z_s, edge_distance=x_edge, edge_index=ei, batch=len(z_s), hetero=False)
        ede = model.edge_degree_embedding(z_s, x_edge, ei, hetero=False)
    else:
        xd = {LIGAND: emb(x_s), PROTEIN: emb(x_d)}
        zd = {LIGAND: z_s, PROTEIN: z_d}
        out = block(x=xd, atomic_numbers=zd, edge_distance=x_edge, edge_index=ei, batch=len(z_d), hetero=True,
                    source_target=(LIGAND, PROTEIN))
        ede = model.edge_degree_embedding(zd, x_edge, ei, hetero=True, source_target=(LIGAND, PROTEIN))

    def ptensor(shape, s):
        return torch.from_numpy(np.random.RandomState(s).normal(size=tuple(shape)).astype(np.float32)).to(dev)

    loss = (out.embedding * ptensor(out.embedding.shape, seed + 300)).sum() + \
           (ede.embedding * ptensor(ede.embedding.shape, seed + 301)).sum()
    loss.backward()
    errs = {
        "block_out": rel_err(out.embedding, case["block_out"]),
        "edge_degree_out": rel_err(ede.embedding, case["edge_degree_out"]),
        "grad_x_edge": rel_err(x_edge.grad, case["grad_x_edge"]),
        "grad_x_src": rel_err(x_s.grad, case["grad_x_src"]),
        "grad_x_dst": rel_err(x_d.grad, case["grad_x_dst"]),
    }
    params = dict(model.named_parameters())
    for n, dig in case["param_grads"].items():
        assert params[n].grad is not None, f"missing gradient for {n}"
        errs["grad:" + n] = digest_err(params[n].grad, dig)
    if return_all:
        return errs
    return max(errs.values())
