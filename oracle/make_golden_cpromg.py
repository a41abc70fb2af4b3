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
        pyg.nn = mod("torch_geometric.nn", knn_graph=co.knn_graph)
        pyg.utils = mod("torch_geometric.utils", get_laplacian=co.get_laplacian, to_dense_batch=co.to_dense_batch,
                        to_undirected=co.to_undirected)
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


class _Cfg(dict):
    def __init__(self, d=None, **kw):
        super().__init__()
        for k, v in dict(d or {}, **kw).items():
            self[k] = _Cfg(v) if isinstance(v, dict) else v
    __getattr__ = dict.__getitem__


def transformer_inputs(seed=5, tgt_len=12, vocab=116, pad=110):
    """Two small complexes (27 + 33 'atoms', 7 + 9 'residue' nodes), portable numpy streams."""
    rs = np.random.RandomState(seed)
    n_p, n_l = [27, 33], [7, 9]
    f32 = lambda a: torch.from_numpy(np.asarray(a, dtype=np.float32))
    pos = f32(np.concatenate([rs.uniform(-5, 5, size=(n, 3)) + 40.0 * i for i, n in enumerate(n_p)]))
    aa_pos = f32(np.concatenate([rs.uniform(-3, 3, size=(n, 3)) + 40.0 * i for i, n in enumerate(n_l)]))
    batch = torch.from_numpy(np.repeat([0, 1], n_p))
    aa_batch = torch.from_numpy(np.repeat([0, 1], n_l))
    toks = rs.randint(1, vocab, size=(2, tgt_len))
    toks[toks == pad] = 1
    toks[0, 8:] = pad
    return dict(node_attr=f32(rs.normal(size=(sum(n_p), 784)) * 0.3), pos=pos, batch=batch,
                atom_laplacian=f32(rs.normal(size=(sum(n_p), 8)) * 0.3), smiles_index=torch.from_numpy(toks),
                tgt_len=tgt_len, aa_node_attr=f32(rs.normal(size=(sum(n_l), 784)) * 0.3), aa_pos=aa_pos,
                aa_batch=aa_batch, aa_laplacian=f32(rs.normal(size=(sum(n_l), 8)) * 0.3),
                prop=f32(rs.randint(0, 2, size=(2, 3))))


TRANSFORMER_CFG = dict(hidden_channels=256, featurizer_feat_dim=784,
                       encoder=dict(hidden_channels=256, edge_channels=64, key_channels=128, num_heads=4,
                                    num_interactions=6, cutoff=10.0, knn=6, lap_dim=8),
                       decoder=dict(tgt_len=12, hidden_channels=256, key_channels=128, num_heads=4, num_interactions=6,
                                    padding="right"))


def main_transformer():
    """The whole reference ``Transformer`` (CProMG.py:423-464: Encoder, Encoder2, Decoder) in eval mode on CPU."""
    import yaml
    ref = load_cpromg()
    # keep the graph utilities bound after load_cpromg restored sys.modules: the reference module holds references
    voc = yaml.safe_load(open(os.path.join(rl.REFERENCE_ROOT, "config", "train.yml")))["model"]["decoder"]["smiVoc"]
    cfg = _Cfg(TRANSFORMER_CFG)
    cfg.decoder["smiVoc"] = voc
    from singa_b200 import synthetic as syn
    torch.manual_seed(0)
    model = ref.Transformer(cfg, 784, num_props=3, device="cpu").eval()
    syn.load_portable_parameters_fanin(model, 5)      # parameters are regenerated from the seed, not stored
    inp = transformer_inputs()
    for k in ("node_attr", "aa_node_attr"):
        inp[k].requires_grad_(True)
    out = model(**inp)
    rs = np.random.RandomState(9)
    w = torch.from_numpy(rs.normal(size=tuple(out.shape)).astype(np.float32))
    (out * w).sum().backward()
    from oracle.make_golden import digest
    fixture = dict(cfg=TRANSFORMER_CFG, vocab_size=len(voc), pad_index=voc.index("^"),
                   param_seed=5, surface={k: [list(v.shape), str(v.dtype)] for k, v in model.state_dict().items()},
                   inputs={k: (v.detach().clone() if torch.is_tensor(v) else v) for k, v in inp.items()},
                   out=out.detach().clone(), loss_weights=w, d_node_attr=inp["node_attr"].grad.clone(),
                   d_aa_node_attr=inp["aa_node_attr"].grad.clone(),
                   d_params={k: digest(p.grad) for k, p in model.named_parameters() if p.grad is not None},
                   params_without_grad=[k for k, p in model.named_parameters() if p.grad is None])
    path = os.path.join(GOLDEN_DIR, "cpromg_transformer.pt")
    torch.save(fixture, path)
    print(path, "out", tuple(out.shape), os.path.getsize(path), "bytes; no grad:", fixture["params_without_grad"])


if __name__ == "__main__":
    if "--attention" in sys.argv:
        main()
    main_transformer()
