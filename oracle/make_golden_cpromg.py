"""TEST INFRASTRUCTURE ONLY -- writes tests/golden/cpromg_attention.pt from the UNMODIFIED reference class
``MultiHeadAttention`` (model/CProMG.py:19-78), run on CPU in the build container.

  python -m oracle.make_golden_cpromg

model/CProMG.py imports five absent, un-vendored packages at module level (dgl, torch_geometric, torch_scatter,
torch_cluster: environment.yml:171-175).  ``MultiHeadAttention`` only calls two of their functions,
``torch_scatter.scatter_softmax`` and ``scatter_sum`` (CProMG.py:69, 76); they are stood in for here by per-segment
Python loops over ``torch.softmax`` / ``sum`` -- a route independent of the vectorised restatement in
oracle/cpromg_oracle.py that the fixture pins.  Every other imported name is an inert placeholder.

Fixture: config/train.yml:53-63 sizes (hidden 256, edge 64, key 128, 4 heads), a 40-node kNN-6 graph made undirected,
default-initialised parameters under torch.manual_seed(0), inputs from numpy RandomState streams.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

import numpy as np
import torch

from oracle import cpromg_oracle as co
from oracle import ref_loader as rl

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _loop_scatter_softmax(src, index, dim=0):
    out = torch.empty_like(src)
    for seg in torch.unique(index).tolist():
        m = index == seg
        out[m] = torch.softmax(src[m], dim=0)
    return out


def _loop_scatter_sum(src, index, dim=0, dim_size=None):
    n = int(index.max()) + 1 if dim_size is None else dim_size
    out = torch.zeros((n,) + src.shape[1:], dtype=src.dtype)
    for seg in torch.unique(index).tolist():
        out[seg] = src[index == seg].sum(dim=0)
    return out


def load_cpromg():
    """Import the reference's model/CProMG.py with placeholder modules for its absent dependencies."""
    def mod(name, **attrs):
        m = types.ModuleType(name)
        for k, v in attrs.items():
            setattr(m, k, v)
        sys.modules[name] = m
        return m
    unused = lambda *a, **k: (_ for _ in ()).throw(NotImplementedError("placeholder"))
    saved = {k: sys.modules.get(k) for k in ("dgl", "torch_geometric", "torch_geometric.data", "torch_geometric.nn",
                                             "torch_geometric.utils", "torch_scatter", "torch_cluster")}
    try:
        mod("dgl")
        pyg = mod("torch_geometric")
        pyg.data = mod("torch_geometric.data", Data=object, HeteroData=object)
        pyg.nn = mod("torch_geometric.nn", knn_graph=unused)
        pyg.utils = mod("torch_geometric.utils", get_laplacian=unused, to_dense_batch=unused, to_undirected=unused)
        mod("torch_scatter", scatter_sum=_loop_scatter_sum, scatter_softmax=_loop_scatter_softmax)
        mod("torch_cluster")
        spec = importlib.util.spec_from_file_location("_ref_cpromg", os.path.join(rl.REFERENCE_ROOT, "model", "CProMG.py"))
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
        return m
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def main():
    ref = load_cpromg()
    hidden, edge_ch, key_ch, heads, n, k = 256, 64, 128, 4, 40, 6
    torch.manual_seed(0)
    mha = ref.MultiHeadAttention(hidden, edge_ch, key_ch, heads, device="cpu")
    rs = np.random.RandomState(11)
    pos = torch.from_numpy(rs.uniform(-6, 6, size=(n, 3)).astype(np.float32))
    batch = torch.from_numpy(np.repeat([0, 1], [25, 15]))
    node_attr = torch.from_numpy(rs.normal(size=(n, hidden)).astype(np.float32)).requires_grad_(True)
    di = co.knn_graph_bruteforce(pos, k, batch)
    pairs = sorted(set(map(tuple, di.t().tolist())) | set(map(tuple, di.flip(0).t().tolist())))   # to_undirected
    edge_index = torch.tensor(pairs, dtype=torch.long).t().contiguous()
    edge_index = edge_index[:, torch.from_numpy(rs.permutation(edge_index.shape[1]))]            # unsorted by centre
    length = (pos[edge_index[0]] - pos[edge_index[1]]).norm(dim=1)
    edge_attr = ref.GaussianSmearing(stop=15, num_gaussians=edge_ch, device="cpu")(length).detach()
    edge_attr.requires_grad_(True)
    out = mha(node_attr, edge_index, edge_attr)
    w = torch.from_numpy(rs.normal(size=tuple(out.shape)).astype(np.float32))
    (out * w).sum().backward()
    fixture = dict(
        dims=dict(hidden=hidden, edge=edge_ch, key=key_ch, heads=heads, knn=k),
        state_dict={k_: v.detach().clone() for k_, v in mha.state_dict().items()},
        pos=pos, batch=batch, node_attr=node_attr.detach().clone(), edge_index=edge_index,
        edge_attr=edge_attr.detach().clone(), out=out.detach().clone(), loss_weights=w,
        d_node_attr=node_attr.grad.clone(), d_edge_attr=edge_attr.grad.clone(),
        d_params={k_: p.grad.clone() for k_, p in mha.named_parameters()}, knn_directed=di)
    path = os.path.join(GOLDEN_DIR, "cpromg_attention.pt")
    torch.save(fixture, path)
    print(path, "edges", edge_index.shape[1], "out", tuple(out.shape), os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
