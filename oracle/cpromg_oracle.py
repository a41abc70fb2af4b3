"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the next row of the scope table: the CProMG
encoder's graph message passing, ``MultiHeadAttention.forward`` (reference model/CProMG.py:50-78), and the k-nearest-
neighbour graph it runs on (``knn_graph``, CProMG.py:293, 330).  No CUDA path exists for this row yet: this file and
tests/golden/cpromg_attention.pt are the checker that path will be held to.  Nothing under singa_b200/ may import it.

Third-party pieces restated here (absent from /root/reference, pinned in environment.yml:171-175):
  torch_scatter 2.1.1  scatter_softmax(src, index, dim=0): exp(src - max_seg) / sum_seg   (no epsilon, unlike pyg softmax)
                       scatter_sum(src, index, dim=0, dim_size=N)
  torch_cluster 1.6.3  knn_graph(x, k, batch, flow='target_to_source'): for every centre i the k nearest OTHER points of
                       the same batch item by squared Euclidean distance; row 0 = centre, row 1 = neighbour for this
                       flow.  "parity unpinned": no reference vector; ties are broken here by the smaller index.
Parity status of the attention restatement: pinned against the reference class itself (imported with the two
torch_scatter functions above as its only stand-ins) by oracle/make_golden_cpromg.py -> tests/golden/cpromg_attention.pt.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


def scatter_softmax(src: torch.Tensor, index: torch.Tensor, num_segments: int) -> torch.Tensor:
    """torch_scatter.scatter_softmax(src, index, dim=0) for src [E, H] (CProMG.py:69)."""
    mx = torch.full((num_segments,) + src.shape[1:], -float("inf"), dtype=src.dtype)
    mx = mx.scatter_reduce(0, index.view(-1, *([1] * (src.dim() - 1))).expand_as(src), src, reduce="amax")
    e = torch.exp(src - mx[index])
    den = torch.zeros_like(mx).index_add_(0, index, e)
    return e / den[index]


def scatter_sum(src: torch.Tensor, index: torch.Tensor, num_segments: int) -> torch.Tensor:
    """torch_scatter.scatter_sum(src, index, dim=0, dim_size=num_segments) (CProMG.py:76)."""
    return torch.zeros((num_segments,) + src.shape[1:], dtype=src.dtype).index_add_(0, index, src)


def shifted_softplus(x: torch.Tensor) -> torch.Tensor:
    """ShiftedSoftplus (CProMG.py:489-496)."""
    return F.softplus(x) - math.log(2.0)


def grouped_pointwise(x: torch.Tensor, w: torch.Tensor, groups: int) -> torch.Tensor:
    """Conv1d(C_in, C_out, 1, groups, bias=False) on x.unsqueeze(-1) (CProMG.py:27-29, 55-57):
    x [N, C_in], w [C_out, C_in/groups, 1] -> [N, groups, C_out/groups]."""
    N = x.shape[0]
    cin_g, cout_g = x.shape[1] // groups, w.shape[0] // groups
    xg = x.view(N, groups, cin_g)
    wg = w.view(groups, cout_g, cin_g)
    return torch.einsum("ngi,goi->ngo", xg, wg)


def multi_head_attention(p: dict, node_attr: torch.Tensor, edge_index: torch.Tensor, edge_attr: torch.Tensor,
                         num_heads: int) -> torch.Tensor:
    """MultiHeadAttention.forward (CProMG.py:50-78).  ``p`` is the module's state_dict (same key names).
    row = edge_index[0] is the aggregating centre, col = edge_index[1] the neighbour."""
    N = node_attr.shape[0]
    row, col = edge_index
    lin = lambda name, x: F.linear(x, p[name + ".weight"], p.get(name + ".bias"))
    h_keys = grouped_pointwise(node_attr, p["k_lin.weight"], num_heads)
    h_queries = grouped_pointwise(node_attr, p["q_lin.weight"], num_heads)
    h_values = grouped_pointwise(node_attr, p["v_lin.weight"], num_heads)
    W_k = lin("weight_k_net.2", shifted_softplus(lin("weight_k_net.0", edge_attr)))
    keys_j = lin("weight_k_lin", W_k.unsqueeze(1) * h_keys[col])
    qk = (h_queries[row] * keys_j).sum(-1) / math.sqrt(keys_j.shape[-1])
    alpha = scatter_softmax(qk, row, N)
    W_v = lin("weight_v_net.2", shifted_softplus(lin("weight_v_net.0", edge_attr)))
    msg = alpha.unsqueeze(-1) * lin("weight_v_lin", W_v.unsqueeze(1) * h_values[col])
    out = lin("centroid_lin", node_attr) + scatter_sum(msg, row, N).view(N, -1)
    out = lin("out_transform", shifted_softplus(out))
    return F.layer_norm(out, (out.shape[-1],), p["layer_norm.weight"], p["layer_norm.bias"])


def knn_graph_bruteforce(pos: torch.Tensor, k: int, batch: torch.Tensor = None) -> torch.Tensor:
    """knn_graph(pos, k, batch, flow='target_to_source') (CProMG.py:293): [2, <= N*k] int64, row 0 = centre (ascending),
    row 1 = its nearest neighbours in order of increasing squared distance (fp32 ((dx^2+dy^2)+dz^2), ties by index)."""
    p = pos.to(torch.float32)
    N = p.shape[0]
    d2 = torch.zeros(N, N, dtype=torch.float32)
    for a in range(3):
        diff = p[:, None, a] - p[None, :, a]
        d2 = d2 + diff * diff
    bad = torch.eye(N, dtype=torch.bool)
    if batch is not None:
        bad |= batch[:, None] != batch[None, :]
    d2 = d2.masked_fill(bad, float("inf"))
    order = torch.argsort(d2, dim=1, stable=True)[:, :k]
    keep = torch.gather(d2, 1, order) < float("inf")
    centre = torch.arange(N).unsqueeze(1).expand_as(order)
    return torch.stack((centre[keep], order[keep]))


# ---------------------------------------------------------------------------------------------------------------
# torch_geometric 2.3.1 utilities the encoders call around the attention (CProMG.py:293-300, 304, 307, 330-334),
# restated from their published behaviour with plain loops (an independent route from the sort-based device code in
# singa_b200/cpromg.py).  Un-vendored, no reference vector: "parity unpinned".
# ---------------------------------------------------------------------------------------------------------------
def to_undirected(edge_index, edge_attr=None, num_nodes=None, reduce="add"):
    """Both directions of every edge; duplicate pairs merged, attribute = mean (reduce='mean') of the merged entries;
    output sorted by (row, col)."""
    assert reduce == "mean" and edge_attr is not None
    acc = {}
    for (r, c), a in zip(edge_index.t().tolist(), edge_attr.tolist()):
        for key in ((r, c), (c, r)):
            acc.setdefault(key, []).append(a)
    keys = sorted(acc)
    ei = torch.tensor(keys, dtype=torch.long).t().contiguous()
    attr = torch.tensor([sum(acc[k]) / len(acc[k]) for k in keys], dtype=edge_attr.dtype)
    return ei, attr


def get_laplacian(edge_index, edge_weight=None, normalization=None, dtype=None, num_nodes=None):
    """normalization=None: L = D - A.  Self loops removed, then one loop per node appended; weights cat(-w, deg) with
    deg[i] = sum of w over the edges whose row is i.  Works for per-edge weight vectors [E, C] (CProMG.py:300)."""
    assert normalization is None
    keep = edge_index[0] != edge_index[1]
    edge_index, edge_weight = edge_index[:, keep], edge_weight[keep]
    n = int(edge_index.max()) + 1 if num_nodes is None else num_nodes
    deg = torch.zeros((n,) + edge_weight.shape[1:], dtype=edge_weight.dtype).index_add_(0, edge_index[0], edge_weight)
    loops = torch.arange(n, dtype=torch.long)
    return torch.cat((edge_index, torch.stack((loops, loops))), dim=1), torch.cat((-edge_weight, deg), dim=0)


def to_dense_batch(x, batch=None, fill_value=0.0, max_num_nodes=None, batch_size=None):
    """[N, D] rows of a sorted batch vector -> ([B, Nmax, D], mask [B, Nmax])."""
    B = int(batch.max()) + 1
    counts = torch.bincount(batch, minlength=B)
    nmax = int(counts.max())
    out = x.new_full((B, nmax, x.shape[1]), fill_value)
    mask = torch.zeros(B, nmax, dtype=torch.bool)
    start = 0
    for b in range(B):
        c = int(counts[b])
        out[b, :c] = x[start:start + c]
        mask[b, :c] = True
        start += c
    return out, mask


def knn_graph(x, k, batch=None, loop=False, flow="source_to_target", **kw):
    """torch_geometric.nn.knn_graph signature over knn_graph_bruteforce."""
    assert not loop
    ei = knn_graph_bruteforce(x, k, batch)       # row 0 = centre, row 1 = neighbour
    return ei if flow == "target_to_source" else ei.flip(0)
