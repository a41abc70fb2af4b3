"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.pt from the UNMODIFIED reference.

Run in the build container (needs /root/reference):  python -m oracle.make_golden

Fixtures (all float32 / int64 tensors, torch.save):
  example_3wi2.pt      reference EquivariantEmbedding.forward on example/3wi2_4tpp.pt
                       (BASELINE.json config 1): inputs, the three edge frames the
                       reference's init_edge_rot_mat produced from a portable random
                       stream, the four output tensors, and parameter-gradient digests
                       for loss = <out_protein, R_p> + <out_ligand, R_l>.
  attention_small.pt   reference TransBlockV2 (homogeneous and hetero) and
                       EdgeDegreeEmbedding on a small dense random graph, with
                       outputs and full gradients w.r.t. inputs and digests w.r.t. parameters.
  wigner_small.pt      reference SO3_Rotation.set_wigner on random frames (|m|<=mmax rows).
Parameters are NOT stored: both sides regenerate them with
singa_b200.synthetic.portable_parameters(seed).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle import ref_loader as rl  # noqa: E402
from singa_b200 import synthetic as syn  # noqa: E402

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
DIGEST_N = 2048


def digest(t: torch.Tensor) -> dict:
    """Compact summary of a (gradient) tensor: full if small, else strided sample + norms."""
    flat = t.detach().reshape(-1).to(torch.float32)
    if flat.numel() <= DIGEST_N:
        return dict(full=flat.clone().reshape(t.shape))
    stride = flat.numel() // DIGEST_N
    return dict(sample=flat[::stride][:DIGEST_N].clone(), stride=stride,
                sum=float(flat.double().sum()), l2=float(flat.double().norm()))


def portable_tensor(shape, seed, scale=1.0):
    rs = np.random.RandomState(seed)
    return torch.from_numpy((rs.normal(size=shape) * scale).astype(np.float32))


class _RandFeed:
    """Feeds torch.rand_like calls inside init_edge_rot_mat from a fixed list."""

    def __init__(self, draws):
        self.draws = list(draws)
        self.orig = torch.rand_like

    def __enter__(self):
        it = iter(self.draws)
        torch.rand_like = lambda x, **k: next(it).to(x.dtype)
        return self

    def __exit__(self, *a):
        torch.rand_like = self.orig


def example_graph(name="3wi2_4tpp"):
    g = rl.load_example(os.path.join(rl.REFERENCE_ROOT, "example", f"{name}.pt"))
    graph = dict(
        z_p=g["atomicnum"][syn.PROTEIN], z_l=g["atomicnum"][syn.LIGAND],
        x_p=g[syn.PROTEIN]["x"][:, -15:].clone(), x_l=g[syn.LIGAND]["x"][:, -15:].clone(),
        pos_p=g[syn.PROTEIN]["pos"], pos_l=g[syn.LIGAND]["pos"],
        ei_pp=g[syn.PP]["edge_index"], ei_ll=g[syn.LL]["edge_index"],
        ei_lp=g[syn.LP]["edge_index"], ei_pl=g[syn.PL]["edge_index"])
    return g, graph


def make_example(ef, emb):
    torch.manual_seed(0)
    cfg = rl.EasyDict(rl.DEFAULT_EMBEDDING_CONFIG)
    model = emb.EquivariantEmbedding(cfg, device="cpu")
    syn.load_portable_parameters(model, 0)
    g, graph = example_graph()
    rnd = syn.frame_randoms(graph, 7)
    frames = {}
    with _RandFeed([rnd["pp"], rnd["ll"], rnd["lp"]]):
        for key, (s, d, ei) in dict(pp=("pos_p", "pos_p", "ei_pp"), ll=("pos_l", "pos_l", "ei_ll"),
                                    lp=("pos_l", "pos_p", "ei_lp")).items():
            vec = graph[s][graph[ei][0]] - graph[d][graph[ei][1]]
            frames[key] = ef.init_edge_rot_mat(vec)
    with _RandFeed([rnd["pp"], rnd["ll"], rnd["lp"]]):
        out = model(g)
    R_p = portable_tensor(out[syn.PROTEIN].embedding.shape, 101)
    R_l = portable_tensor(out[syn.LIGAND].embedding.shape, 102)
    loss = (out[syn.PROTEIN].embedding * R_p).sum() + (out[syn.LIGAND].embedding * R_l).sum()
    loss.backward()
    grads = {n: digest(p.grad) for n, p in model.named_parameters() if p.grad is not None}
    no_grad = [n for n, p in model.named_parameters() if p.grad is None]
    fixture = dict(
        graph=graph, frames=frames, frame_seed=7, param_seed=0, config=dict(cfg),
        outputs={k: v.embedding.detach().clone() for k, v in out.items()},
        loss=float(loss), loss_seeds=dict(protein=101, ligand=102), grads=grads, params_without_grad=no_grad,
        grid=dict(to62=model.SO3_grid[6][2].to_grid_mat.clone(), from62=model.SO3_grid[6][2].from_grid_mat.clone(),
                  to66=model.SO3_grid[6][6].to_grid_mat.clone(), from66=model.SO3_grid[6][6].from_grid_mat.clone()),
    )
    torch.save(fixture, os.path.join(GOLDEN_DIR, "example_3wi2.pt"))
    print("example_3wi2.pt: loss", float(loss), "params w/o grad:", no_grad)


def random_graph(seed, n_src, n_dst, n_edges, homo):
    rs = np.random.RandomState(seed)
    pos_s = rs.uniform(-4, 4, size=(n_src, 3)).astype(np.float32)
    pos_d = pos_s if homo else rs.uniform(-4, 4, size=(n_dst, 3)).astype(np.float32)
    pairs = set()
    while len(pairs) < n_edges:
        s, d = int(rs.randint(n_src)), int(rs.randint(max(1, n_dst - 3)))  # last 3 targets keep in-degree 0
        if homo and s == d:
            continue
        pairs.add((s, d))
    ei = np.array(sorted(pairs), dtype=np.int64)
    ei = ei[rs.permutation(len(ei))].T.copy()  # unsorted by destination, like the shipped graphs
    vec = pos_s[ei[0]] - pos_d[ei[1]]
    return torch.from_numpy(ei), torch.from_numpy(vec), torch.from_numpy(rs.uniform(size=(n_edges, 3)).astype(np.float32))


def make_attention(ef, emb, lmax=6, fname="attention_small.pt"):
    """TransBlockV2 + EdgeDegreeEmbedding of the unmodified reference on small random graphs; lmax=8 gives the
    BASELINE.json config-5 coefficient layout (L=81, M=39)."""
    torch.manual_seed(0)
    cfg = rl.EasyDict(rl.DEFAULT_EMBEDDING_CONFIG)
    cfg.lmax_list = [lmax]
    L = (lmax + 1) ** 2
    model = emb.EquivariantEmbedding(cfg, device="cpu")
    syn.load_portable_parameters(model, 3)
    block = model.blocks[1]
    cases = {}
    for name, (homo, n_s, n_d, n_e, seed) in dict(homo=(True, 24, 24, 200, 11), hetero=(False, 9, 30, 70, 12)).items():
        ei, vec, rnd = random_graph(seed, n_s, n_d, n_e, homo)
        with _RandFeed([rnd]):
            rot = ef.init_edge_rot_mat(vec)
        model.SO3_rotation[0].set_wigner(rot)
        x_edge = portable_tensor((n_e, 48), seed + 100, 0.7).requires_grad_(True)
        x_s = portable_tensor((n_s, L, 16), seed + 200).requires_grad_(True)
        x_d = x_s if homo else portable_tensor((n_d, L, 16), seed + 201).requires_grad_(True)
        z_s = torch.from_numpy(np.random.RandomState(seed).choice([6, 7, 8, 16], size=n_s))
        z_d = z_s if homo else torch.from_numpy(np.random.RandomState(seed + 1).choice([6, 7, 8, 16], size=n_d))
        model.zero_grad()
        if homo:
            x = ef.SO3_Embedding(0, [lmax], 16, dtype=torch.float32, device="cpu")
            x.set_embedding(x_s * 1.0)
            out = block(x=x, atomic_numbers=z_s, edge_distance=x_edge, edge_index=ei, batch=n_s, hetero=False)
            ede = model.edge_degree_embedding(z_s, x_edge, ei, hetero=False)
        else:
            xs = ef.SO3_Embedding(0, [lmax], 16, dtype=torch.float32, device="cpu")
            xs.set_embedding(x_s * 1.0)
            xd = ef.SO3_Embedding(0, [lmax], 16, dtype=torch.float32, device="cpu")
            xd.set_embedding(x_d * 1.0)
            xdict = {syn.LIGAND: xs, syn.PROTEIN: xd}
            zdict = {syn.LIGAND: z_s, syn.PROTEIN: z_d}
            out = block(x=xdict, atomic_numbers=zdict, edge_distance=x_edge, edge_index=ei, batch=n_d,
                        hetero=True, source_target=(syn.LIGAND, syn.PROTEIN))
            ede = model.edge_degree_embedding(zdict, x_edge, ei, hetero=True, source_target=(syn.LIGAND, syn.PROTEIN))
        R = portable_tensor(out.embedding.shape, seed + 300)
        R2 = portable_tensor(ede.embedding.shape, seed + 301)
        loss = (out.embedding * R).sum() + (ede.embedding * R2).sum()
        loss.backward()
        pg = {n: digest(p.grad) for n, p in model.named_parameters() if p.grad is not None}
        cases[name] = dict(
            edge_index=ei, edge_vec=vec, rot=rot, x_edge=x_edge.detach().clone(), x_src=x_s.detach().clone(),
            x_dst=x_d.detach().clone(), n_dst=n_d, seed=seed, block_out=out.embedding.detach().clone(),
            edge_degree_out=ede.embedding.detach().clone(), grad_x_edge=x_edge.grad.clone(),
            grad_x_src=x_s.grad.clone(), grad_x_dst=x_d.grad.clone(), param_grads=pg, loss=float(loss))
        print(fname, name, "loss", float(loss))
    torch.save(dict(param_seed=3, block="blocks.1.", lmax=lmax, cases=cases), os.path.join(GOLDEN_DIR, fname))


def make_wigner(ef, emb):
    rs = np.random.RandomState(5)
    vec = torch.from_numpy(rs.normal(size=(64, 3)).astype(np.float32))
    vec[0] = torch.tensor([0.0, 1.0, 0.0])   # edge along the polar axis
    vec[1] = torch.tensor([0.0, -2.0, 0.0])  # and against it
    vec[2] = torch.tensor([1.0, 0.0, 0.0])
    rnd = torch.from_numpy(rs.uniform(size=(64, 3)).astype(np.float32))
    with _RandFeed([rnd]):
        rot = ef.init_edge_rot_mat(vec)
    out = {}
    for lmax, mmax in ((6, 2), (8, 2)):
        r = ef.SO3_Rotation(lmax, device="cpu")
        r.set_wigner(rot)
        keep = r.mapping.coefficient_idx(lmax, mmax)
        out[f"rows_l{lmax}_m{mmax}"] = r.wigner[:, keep, :].clone()
    torch.save(dict(edge_vec=vec, rand=rnd, rot=rot, **out), os.path.join(GOLDEN_DIR, "wigner_small.pt"))
    print("wigner_small.pt")


def main():
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    ef, emb = rl.load_reference()
    make_wigner(ef, emb)
    make_attention(ef, emb)
    make_attention(ef, emb, lmax=8, fname="attention_l8.pt")
    make_example(ef, emb)


if __name__ == "__main__":
    main()
