"""Synthetic CrossDocked-shaped complexes and portable parameter values.

Both are pure numpy (legacy ``RandomState`` streams, frozen across numpy
versions) so that the build container (reference on CPU) and the GPU box (CUDA
path) can regenerate bit-identical inputs from a seed without shipping data.

Input format mirrors what ``EquivariantEmbedding.forward`` reads from the pyg
``HeteroData`` (SURVEY.md 2.3 / Embedding.py:227-231, 306-310, 384, 438).
"""
from __future__ import annotations

import zlib

import numpy as np
import torch

PROTEIN, LIGAND = "protein_atoms", "ligand_atoms"
PP = (PROTEIN, "linked_to", PROTEIN)
LL = (LIGAND, "linked_to", LIGAND)
LP = (LIGAND, "interact_with", PROTEIN)
PL = (PROTEIN, "interact_with", LIGAND)

_ELEMENTS = np.array([6, 7, 8, 16])
_ELEMENT_P = np.array([0.6, 0.17, 0.2, 0.03])


def portable_parameters(named_shapes, seed: int = 0) -> dict:
    """Deterministic, platform-independent values for every parameter.

    named_shapes: iterable of (name, shape).  Scale depends on the role so that
    activations stay O(1): weights ~ U(-1/sqrt(fan_in), 1/sqrt(fan_in)), norm
    gains ~ 1 + U(-.2,.2), biases ~ U(-.1,.1), embeddings ~ N(0,1),
    edge element embeddings ~ U(-.05, .05).
    """
    out = {}
    for name, shape in named_shapes:
        shape = tuple(int(s) for s in shape)
        rs = np.random.RandomState((seed * 1000003 + zlib.crc32(name.encode())) % (2 ** 32))
        leaf = name.rsplit(".", 1)[-1]
        if "sphere_embedding" in name:
            v = rs.normal(size=shape)
        elif "source_embedding" in name or "target_embedding" in name:
            v = rs.uniform(-0.05, 0.05, size=shape)
        elif leaf == "affine_weight" or (leaf == "weight" and len(shape) == 1):
            v = 1.0 + rs.uniform(-0.2, 0.2, size=shape)
        elif leaf in ("bias", "affine_bias"):
            v = rs.uniform(-0.1, 0.1, size=shape)
        elif leaf == "alpha_dot":
            b = 1.0 / np.sqrt(shape[-1])
            v = rs.uniform(-b, b, size=shape)
        else:
            fan_in = shape[-1]
            b = 1.0 / np.sqrt(fan_in)
            v = rs.uniform(-b, b, size=shape)
            if ".so2_m_conv." in name:
                v = v / np.sqrt(2.0)
        out[name] = torch.from_numpy(np.ascontiguousarray(v, dtype=np.float32))
    return out


def load_portable_parameters_fanin(module: torch.nn.Module, seed: int = 0) -> None:
    """Portable values for modules outside the EF_layers naming scheme (the CProMG transformer): >= 2-D tensors
    ~ U(+-1/sqrt(fan_in)) with fan_in = prod(shape[1:]) (Linear, grouped Conv1d, Embedding), 1-D ``weight`` (norm
    gains) ~ 1 + U(+-0.2), other 1-D tensors ~ U(+-0.1).  Buffers keep their constructor values."""
    with torch.no_grad():
        for name, p in module.named_parameters():
            rs = np.random.RandomState((seed * 1000003 + zlib.crc32(name.encode())) % (2 ** 32))
            shape = tuple(p.shape)
            if len(shape) >= 2:
                b = 1.0 / np.sqrt(float(np.prod(shape[1:])))
                v = rs.uniform(-b, b, size=shape)
            elif name.rsplit(".", 1)[-1] == "weight":
                v = 1.0 + rs.uniform(-0.2, 0.2, size=shape)
            else:
                v = rs.uniform(-0.1, 0.1, size=shape)
            p.copy_(torch.from_numpy(np.ascontiguousarray(v, dtype=np.float32)).to(p.device))


def load_portable_parameters(module: torch.nn.Module, seed: int = 0) -> None:
    vals = portable_parameters(((n, p.shape) for n, p in module.named_parameters()), seed)
    with torch.no_grad():
        for n, p in module.named_parameters():
            p.copy_(vals[n].to(p.device))


def radius_graph_bruteforce(pos: np.ndarray, r: float) -> np.ndarray:
    """All ordered pairs (j -> i), i != j, with fp32 squared distance < r^2.
    Returns int64 [2, E] sorted by target then source (row 0 = source j, row 1 = target i)."""
    p = pos.astype(np.float32)
    d2 = np.zeros((len(p), len(p)), dtype=np.float32)
    for k in range(3):
        diff = (p[:, None, k] - p[None, :, k]).astype(np.float32)
        d2 = (d2 + diff * diff).astype(np.float32)
    mask = d2 < np.float32(r) * np.float32(r)
    np.fill_diagonal(mask, False)
    tgt, src = np.nonzero(mask)
    return np.stack([src, tgt]).astype(np.int64)


def _ball_points(rs, n, radius, min_sep):
    pts = np.zeros((0, 3))
    while len(pts) < n:
        cand = rs.uniform(-radius, radius, size=(4 * n, 3))
        cand = cand[np.linalg.norm(cand, axis=1) <= radius]
        for c in cand:
            if len(pts) == 0 or np.min(np.linalg.norm(pts - c, axis=1)) >= min_sep:
                pts = np.vstack([pts, c])
                if len(pts) == n:
                    break
    return pts


def make_complex(seed: int, n_protein=None, n_ligand=None, cutoff: float = 8.0, contact: float = 4.0,
                 density: float = 0.045) -> dict:
    """One synthetic complex: protein atoms uniform in a ball
    at ``density`` atoms/A^3 with 1.2 A minimum separation, ligand in a 5 A ball at the
    centre; linked_to edges = radius graph at ``cutoff``; interact_with = P-L pairs
    within ``contact`` (PL is LP with rows swapped, same order)."""
    rs = np.random.RandomState(seed)
    n_p = int(rs.randint(400, 501)) if n_protein is None else n_protein
    n_l = int(rs.randint(20, 46)) if n_ligand is None else n_ligand
    radius = (3.0 * n_p / (4.0 * np.pi * density)) ** (1.0 / 3.0)
    pos_p = _ball_points(rs, n_p, radius, 1.2).astype(np.float32)
    pos_l = _ball_points(rs, n_l, min(5.0, radius), 1.2).astype(np.float32)

    def feats(n):
        x = np.zeros((n, 59), dtype=np.float32)
        x[:, -15:] = rs.randint(0, 2, size=(n, 15))
        x[:, -8] = rs.uniform(-0.5, 0.5, size=n)
        return x

    g = {
        "z_p": rs.choice(_ELEMENTS, size=n_p, p=_ELEMENT_P).astype(np.int64),
        "z_l": rs.choice(_ELEMENTS, size=n_l, p=_ELEMENT_P).astype(np.int64),
        "x_p": feats(n_p), "x_l": feats(n_l), "pos_p": pos_p, "pos_l": pos_l,
        "ei_pp": radius_graph_bruteforce(pos_p, cutoff),
        "ei_ll": radius_graph_bruteforce(pos_l, cutoff),
    }
    d = np.linalg.norm(pos_l[:, None, :] - pos_p[None, :, :], axis=-1)
    li, pi = np.nonzero(d <= contact)
    g["ei_lp"] = np.stack([li, pi]).astype(np.int64)
    g["ei_pl"] = np.stack([pi, li]).astype(np.int64)
    return {k: torch.from_numpy(v) for k, v in g.items()}


def collate(graphs) -> dict:
    """Disjoint union of complexes (what pyg batching does, GAN.py:48-55)."""
    out = {}
    off_p = off_l = 0
    parts = {k: [] for k in graphs[0]}
    ptr_p, ptr_l = [0], [0]
    for g in graphs:
        for k, v in g.items():
            if k == "ei_pp":
                v = v + off_p
            elif k == "ei_ll":
                v = v + off_l
            elif k == "ei_lp":
                v = v + torch.tensor([[off_l], [off_p]])
            elif k == "ei_pl":
                v = v + torch.tensor([[off_p], [off_l]])
            parts[k].append(v)
        off_p += len(g["z_p"])
        off_l += len(g["z_l"])
        ptr_p.append(off_p)
        ptr_l.append(off_l)
    for k, v in parts.items():
        out[k] = torch.cat(v, dim=1 if k.startswith("ei_") else 0)
    out["ptr_p"] = torch.tensor(ptr_p)
    out["ptr_l"] = torch.tensor(ptr_l)
    return out


def frame_randoms(graph: dict, seed: int) -> dict:
    """The uniform [0,1) draws init_edge_rot_mat takes (EF_layers.py:2301), one per pass
    that calls it (Embedding.py:235, 314, 389), from a portable stream."""
    rs = np.random.RandomState(seed)
    return {name: torch.from_numpy(rs.uniform(size=(graph[key].shape[1], 3)).astype(np.float32))
            for name, key in (("pp", "ei_pp"), ("ll", "ei_ll"), ("lp", "ei_lp"))}


class GraphView:
    """Presents the flat dict in the indexing form EquivariantEmbedding.forward uses."""

    def __init__(self, g: dict):
        self._g = g

    def __getitem__(self, key):
        g = self._g
        if key == "atomicnum":
            return {PROTEIN: g["z_p"], LIGAND: g["z_l"]}
        if key == PROTEIN:   # "x": feature rows, or their packed integer codes (dataset format)
            return {"x": g["x_p"] if "x_p" in g else g["bits_p"], "pos": g["pos_p"], "ptr": g.get("ptr_p"),
                    "lap": g.get("lap_p")}
        if key == LIGAND:
            return {"x": g["x_l"] if "x_l" in g else g["bits_l"], "pos": g["pos_l"], "ptr": g.get("ptr_l"),
                    "lap": g.get("lap_l")}
        if key == "ligand_data":
            return {"sas": g["sas"], "logP": g["logP"], "qed": g["qed"], "weight": g["weight"], "tpsa": g["tpsa"],
                    "vina_score": g["vina_score"], "smiIndices_input": g["smi_in"], "smiIndices_tgt": g["smi_tgt"]}
        table = {PP: "ei_pp", LL: "ei_ll", LP: "ei_lp", PL: "ei_pl"}
        if key in table:
            return {"edge_index": g[table[key]]}
        raise KeyError(key)
