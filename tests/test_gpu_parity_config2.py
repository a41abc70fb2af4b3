"""Parity at BASELINE.json config-2 SIZE: one full ~460-atom / 30 k-edge complex through all four
passes, forward AND backward, CUDA path vs the CPU oracle, in both precisions, with the MEASURED errors written to
gpurun_out/r2_parity_config2.json (copied to profiles/); and a batch that crosses a 262 144-edge chunk boundary,
checked against the oracle on the complex that straddles the boundary (ops.MAX_CHUNK_EDGES = 524 288).

Stated bounds (||d||_inf / ||ref||_inf per tensor): fp32 mode outputs <= 2e-5, parameter gradients <= 2e-4 (sums over
~3e4 edges x 3 layers); bf16x3 accurate tensor-core mode outputs <= 1e-3, gradients <= 5e-3; bf16 tcgen05 mode
outputs <= 3e-2, gradients <= 8e-2.
"""
import json
import os

import numpy as np
import pytest
import torch

from tests.parity_util import LIGAND, PROTEIN, build_model, oracle_tables, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BOUNDS = {"fp32": (2e-5, 2e-4), "bf16x3": (1e-3, 5e-3), "bf16": (3e-2, 8e-2)}
_REF = {}


def _reference_run():
    """CPU oracle fwd+bwd on the 460-atom complex, once per session (about 10 s on the box's host cores)."""
    if not _REF:
        import bench
        graph, _ = bench.cpu_sample_graph(bench.PARITY_SAMPLE_ATOMS)
        model = build_model("cpu", 0)
        sec, outs, grads = bench.run_cpu_port(graph, 1, 0, state=model.state_dict(), keep=True)
        _REF.update(graph=graph, outs=outs, grads=grads, sec=sec)
    return _REF


def _record(key, value):
    path = os.path.join(ROOT, "gpurun_out", "r2_parity_config2.json")
    if not os.path.isdir(os.path.dirname(path)):
        return
    data = json.load(open(path)) if os.path.isfile(path) else {}
    data[key] = value
    json.dump(data, open(path, "w"), indent=1)


@pytest.mark.parametrize("precision", ["fp32", "bf16x3", "bf16"])
def test_full_size_complex_fwd_bwd_matches_oracle(precision):
    import bench
    from singa_b200 import ops
    ref = _reference_run()
    try:
        model = build_model(DEV, 0, precision)
        par = bench.parity_against_port(model, ref["graph"], ref["outs"], ref["grads"], torch.device(DEV))
    finally:
        ops.set_precision("fp32")
    par["edges"] = bench.count_edges(ref["graph"])
    _record(precision, par)
    tol_out, tol_grad = BOUNDS[precision]
    assert par["n_grads"] > 100
    assert par["out_rel_err"] < tol_out, par
    assert par["grad_rel_err"] < tol_grad, par


def test_batch_across_a_chunk_boundary_matches_oracle_on_the_straddling_complex():
    """20 complexes = ~575 k protein-protein edges: the second chunk starts inside one complex's destination range
    (chunks hold whole destinations, ops.plan_chunks).  That complex's outputs, computed inside the batch, against the
    oracle run on the complex alone (complexes share no edges)."""
    import bench
    from oracle import ef_oracle as eo
    from singa_b200 import ops
    from singa_b200 import synthetic as syn
    from singa_b200.embedding import DEFAULT_CONFIG
    graphs = [syn.make_complex(i) for i in range(20)]
    batch = syn.collate(graphs)
    cum = np.cumsum([0] + [g["ei_pp"].shape[1] for g in graphs])
    assert cum[-1] > ops.MAX_CHUNK_EDGES, "batch must span two chunks"
    k = int(np.searchsorted(cum, ops.MAX_CHUNK_EDGES, side="right")) - 1   # complex holding the boundary edge
    assert cum[k] < ops.MAX_CHUNK_EDGES < cum[k + 1]
    per = [bench.oracle_frames(g, seed=20 + i) for i, g in enumerate(graphs)]
    frames = {key: torch.cat([p[key] for p in per]).to(DEV) for key in ("pp", "ll", "lp")}
    model = build_model(DEV, 0, "fp32")
    with torch.no_grad():
        out = model(syn.GraphView({a: b.to(DEV) for a, b in batch.items()}), frames=frames)
        p = {a: b.detach().cpu() for a, b in model.state_dict().items()}
        ref = eo.embedding_forward(p, graphs[k], per[k], oracle_tables(), dict(DEFAULT_CONFIG), compute_dead_blocks=False)
    pp, pl = batch["ptr_p"].tolist(), batch["ptr_l"].tolist()
    ep = rel_err(out[PROTEIN].embedding[pp[k]:pp[k + 1]], ref[PROTEIN])
    el = rel_err(out[LIGAND].embedding[pl[k]:pl[k + 1]], ref[LIGAND])
    _record("chunk_boundary_fp32", {"complex": k, "protein_rel_err": ep, "ligand_rel_err": el,
                                    "pp_edges_in_batch": int(cum[-1])})
    assert ep < 2e-5 and el < 2e-5, (ep, el)
