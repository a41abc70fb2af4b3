"""CSR-carrying on-disk format and loader.

The reference stores one pickled pyg ``HeteroData`` per complex and, per item and inside the DataLoader worker, does
``torch.load(path).to(device)`` (utils/Data.py:179-182): a CUDA context per worker, unsorted edge lists that every
step has to sort again, the 59-column one-hot feature rows of which the model reads 15 columns as a bit string
(Embedding.py:250-253), and a Laplacian eigendecomposition on every forward (CProMG.py:562-571).

A *record* here is a flat dict of numpy arrays per complex, written once by ``pack_complex``:

  z_p, z_l            int16   atomic numbers
  bits_p, bits_l      int32   the 15 feature columns already packed into the integer the model embeds
                              (``embedding.pack_feature_bits``: same truncation of the charge column)
  pos_p, pos_l        float32
  for each of the four edge lists (pp, ll, lp, pl):
    {k}_src, {k}_dst  int32   the stored edge list: DESTINATION-SORTED (stable) for pp / ll, with {k}_orig = stored
                              position -> input edge id; input order for lp / pl (their pairing is semantic: pass 4
                              reuses pass 3's Wigner matrices edge for edge), with {k}_perm = sorted position ->
                              stored id and the sorted copies {k}_ssrc / {k}_sdst
    {k}_rowptr        int32   CSR over destinations
    {k}_srcptr/_srcidx int32  the source-side grouping the backward of the gather needs
  lap_p, lap_l        float32 [n, 8] Laplacian positional encodings (cpromg.laplacian_pe), computed once
  ligand scalars      float32 sas, logP, qed, weight, tpsa, vina_score; smi_in, smi_tgt int16 [tgt_len]

``collate_records`` concatenates records into one batch with offset indices -- a concatenation of destination-sorted
lists over disjoint node ranges is destination-sorted, so the batch CSR needs NO sort and no device->host read (the
chunk plan comes from the host row pointers) -- inside ONE pinned host buffer, moved with ONE host->device copy;
``BatchLoader`` keeps its iterator alive across steps (train.py:116 re-creates the DataLoader iterator, and with it
the worker processes, on every step).
"""
from __future__ import annotations

import os
from typing import Dict, Iterable, List, Optional

import numpy as np
import torch

from . import synthetic as syn

EDGE_KEYS = ("pp", "ll", "lp", "pl")
_NODE_OF = {"pp": ("p", "p"), "ll": ("l", "l"), "lp": ("l", "p"), "pl": ("p", "l")}   # (source side, destination side)
LIGAND_SCALARS = ("sas", "logP", "qed", "weight", "tpsa", "vina_score")


SORTED_KEYS = ("pp", "ll")   # stored destination-sorted (the bulk of the edges); lp / pl keep their input order


def _csr_fields(edge_index: np.ndarray, n_src: int, n_dst: int, store_sorted: bool) -> Dict[str, np.ndarray]:
    """CSR of one edge list.  ``store_sorted``: the list itself is stored destination-sorted (``perm`` = identity,
    ``orig`` maps a stored position to the input edge id).  Otherwise the list is stored as given and ``perm`` maps a
    sorted position to the stored edge id: the two interaction lists must keep their pairing, because pass 4 of the
    caller reuses, edge for edge, the Wigner matrices of pass 3 (Embedding.py:437-471)."""
    src, dst = edge_index[0].astype(np.int64), edge_index[1].astype(np.int64)
    perm = np.argsort(dst, kind="stable")
    src_s, dst_s = src[perm], dst[perm]
    rowptr = np.concatenate(([0], np.cumsum(np.bincount(dst, minlength=n_dst))))
    srcidx = np.argsort(src_s, kind="stable")
    srcptr = np.concatenate(([0], np.cumsum(np.bincount(src, minlength=n_src))))
    i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
    out = dict(rowptr=i32(rowptr), srcptr=i32(srcptr), srcidx=i32(srcidx))
    if store_sorted:
        out.update(src=i32(src_s), dst=i32(dst_s), orig=i32(perm))
    else:
        out.update(src=i32(src), dst=i32(dst), perm=i32(perm), ssrc=i32(src_s), sdst=i32(dst_s))
    return out


def pack_bits(x: np.ndarray) -> np.ndarray:
    """Last 15 feature columns cast to integers (truncation toward zero) and read as a big-endian bit string:
    numpy twin of embedding.pack_feature_bits (Embedding.py:250-253)."""
    w = 2 ** np.arange(14, -1, -1, dtype=np.int64)
    return (x[:, -15:].astype(np.int64) * w).sum(axis=1).astype(np.int32)


def _encoder_graph_fields(pos: np.ndarray, k: int) -> Dict[str, np.ndarray]:
    """Host-built message-passing graph of one encoder for one complex (cpromg.encoder_graph_host) with the CSR of its
    Laplacian edge list [undirected pairs..., one loop per node] grouped by centre (row 0)."""
    from .cpromg import encoder_graph_host
    n = pos.shape[0]
    g = encoder_graph_host(pos, k)
    und = g["und"]
    loops = np.arange(n, dtype=np.int32)
    lap = np.concatenate((und, np.stack((loops, loops))), axis=1)
    f = _csr_fields(lap[::-1], n, n, store_sorted=False)      # destination = centre = row 0, source = neighbour = row 1
    return dict(und=und, und_rowptr=g["und_rowptr"], perm=f["perm"], rowptr=f["rowptr"], ssrc=f["ssrc"], sdst=f["sdst"],
                srcptr=f["srcptr"], srcidx=f["srcidx"])


ENCODER_GRAPHS = (("ga", "p", 48), ("gl", "l", 30))   # (record prefix, node side, k): CProMG.py:293 (config knn), :330


def pack_complex(g: Dict[str, torch.Tensor], ligand_data: Optional[Dict] = None, lap_k: int = 8, key=None,
                 lap_cache=None, encoder_graphs: bool = True) -> Dict[str, np.ndarray]:
    """One complex in the flat-dict form of ``synthetic.make_complex`` (z_*, x_*, pos_*, ei_*) -> record."""
    from .cpromg import LapPECache
    npy = {k: (v.numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in g.items()}
    n = {"p": len(npy["z_p"]), "l": len(npy["z_l"])}
    rec = dict(z_p=npy["z_p"].astype(np.int16), z_l=npy["z_l"].astype(np.int16), bits_p=pack_bits(npy["x_p"]),
               bits_l=pack_bits(npy["x_l"]), pos_p=npy["pos_p"].astype(np.float32), pos_l=npy["pos_l"].astype(np.float32))
    for k in EDGE_KEYS:
        s, d = _NODE_OF[k]
        for name, arr in _csr_fields(npy["ei_" + k], n[s], n[d], k in SORTED_KEYS).items():
            rec[f"{k}_{name}"] = arr
    if lap_cache is not None and key is None:
        raise ValueError("pack_complex: a shared lap_cache needs a per-complex key (key=None would return the first "
                         "complex's encodings for every complex)")
    cache = lap_cache or LapPECache(lap_k)
    rec["lap_p"] = cache.get((key, "p"), npy["ei_pp"], n["p"])
    rec["lap_l"] = cache.get((key, "l"), npy["ei_ll"], n["l"])
    if encoder_graphs:
        for prefix, side, k in ENCODER_GRAPHS:
            for name, arr in _encoder_graph_fields(rec["pos_" + side], k).items():
                rec[f"{prefix}_{name}"] = arr
            rec[f"{prefix}_k"] = np.int32(k)   # checked against the encoder's knn when the graph is used
    if ligand_data is not None:
        for name in LIGAND_SCALARS:
            rec[name] = np.float32(ligand_data[name])
        rec["smi_in"] = np.asarray(ligand_data["smiIndices_input"], dtype=np.int16)
        rec["smi_tgt"] = np.asarray(ligand_data["smiIndices_tgt"], dtype=np.int16)
    return rec


def save_record(path: str, rec: Dict[str, np.ndarray]) -> None:
    np.savez(path, **rec)


def load_record(path: str) -> Dict[str, np.ndarray]:
    with np.load(path) as f:
        return {k: f[k] for k in f.files}


class ComplexDataset:
    """Directory of ``*.npz`` records (or an in-memory list of records)."""

    def __init__(self, source):
        if isinstance(source, str):
            self.paths = sorted(os.path.join(source, f) for f in os.listdir(source) if f.endswith(".npz"))
            self.records = None
        else:
            self.paths, self.records = None, list(source)

    def __len__(self):
        return len(self.paths) if self.records is None else len(self.records)

    def __getitem__(self, i):
        return load_record(self.paths[i]) if self.records is None else self.records[i]

    def edge_counts(self) -> List[int]:
        return [int(sum(len(self[i][f"{k}_src"]) for k in EDGE_KEYS)) for i in range(len(self))]


# ---------------------------------------------------------------------------------------------------------------
# collation: one pinned buffer, one H2D copy, batch CSR without a sort
# ---------------------------------------------------------------------------------------------------------------
_ALIGN = 64


class _Arena:
    """Lays arrays out back to back (64-byte aligned) in one pinned uint8 buffer; after ``to`` the same offsets
    address views of one device buffer."""

    def __init__(self):
        self.items, self.size = [], 0

    def add(self, name: str, arr: np.ndarray):
        arr = np.ascontiguousarray(arr)
        off = (self.size + _ALIGN - 1) // _ALIGN * _ALIGN
        self.items.append((name, off, arr))
        self.size = off + arr.nbytes

    def build(self, pin: bool = True):
        host = torch.empty(max(self.size, 1), dtype=torch.uint8, pin_memory=pin and torch.cuda.is_available())
        hv = host.numpy()
        for _, off, arr in self.items:
            hv[off:off + arr.nbytes] = arr.view(np.uint8).reshape(-1)
        return host

    def views(self, buf: torch.Tensor) -> Dict[str, torch.Tensor]:
        out = {}
        for name, off, arr in self.items:
            t = buf[off:off + arr.nbytes].view(_TORCH[arr.dtype.type])
            out[name] = t.view(*arr.shape)
        return out


_TORCH = {np.int16: torch.int16, np.int32: torch.int32, np.int64: torch.int64, np.float32: torch.float32,
          np.uint8: torch.uint8}


class HostBatch:
    """A collated batch in pinned host memory; ``to(device)`` is one copy."""

    def __init__(self, arena: _Arena, host: torch.Tensor, meta: dict):
        self.arena, self.host, self.meta = arena, host, meta

    @property
    def nbytes(self) -> int:
        return int(self.host.numel())

    def to(self, device) -> "DeviceBatch":
        dev = self.host.to(device, non_blocking=True)
        return DeviceBatch(self.arena.views(dev), self.meta)


class DeviceBatch:
    """Device-side views + host-side metadata (sizes, host row pointers for the chunk plans)."""

    def __init__(self, t: Dict[str, torch.Tensor], meta: dict):
        self.t, self.meta = t, meta
        self._graph = None

    def graph(self) -> dict:
        """Flat dict in the form ``synthetic.GraphView`` / ``EquivariantEmbedding`` read, with the prebuilt CSRs
        registered so that no edge list is sorted on the step (EF_layers.register_edge_csr)."""
        if self._graph is None:
            from . import EF_layers as ef
            from . import ops
            t, m = self.t, self.meta
            g = dict(z_p=t["z_p"].long(), z_l=t["z_l"].long(), pos_p=t["pos_p"], pos_l=t["pos_l"],
                     bits_p=t["bits_p"].long(), bits_l=t["bits_l"].long(), ptr_p=t["ptr_p"].long(),
                     ptr_l=t["ptr_l"].long(), lap_p=t["lap_p"], lap_l=t["lap_l"])
            n = {"p": m["n_p"], "l": m["n_l"]}
            for k in EDGE_KEYS:
                s, d = _NODE_OF[k]
                ei = torch.stack((t[f"{k}_src"], t[f"{k}_dst"])).long()
                E = ei.shape[1]
                if k in SORTED_KEYS:   # the stored order IS the sorted order
                    perm = torch.arange(E, dtype=torch.int32, device=ei.device)
                    ssrc, sdst = t[f"{k}_src"], t[f"{k}_dst"]
                else:
                    perm, ssrc, sdst = t[f"{k}_perm"], t[f"{k}_ssrc"], t[f"{k}_sdst"]
                csr = ops.EdgeCSR(n[s], n[d], E, perm, t[f"{k}_rowptr"], ssrc, sdst, t[f"{k}_dst"],
                                  t[f"{k}_srcptr"], t[f"{k}_srcidx"])
                csr.chunks = ops.plan_chunks(m[f"{k}_rowptr_host"], ops.MAX_CHUNK_EDGES)
                ef.register_edge_csr(ei, n[s], n[d], csr)
                g["ei_" + k] = ei
            for name in LIGAND_SCALARS + ("smi_in", "smi_tgt"):
                if name in t:
                    g[name] = t[name]
            enc = {}
            for (prefix, side, _k), role in zip(ENCODER_GRAPHS, ("atom", "aa")):
                if f"{prefix}_ei" not in t:
                    continue
                from .cpromg import PrebuiltGraph
                N, E = n[side], m[f"{prefix}_und_edges"]
                ei32 = t[f"{prefix}_ei"]
                ei = ei32.long()
                ident = torch.arange(E, dtype=torch.int32, device=ei.device)
                und = ops.EdgeCSR(N, N, E, ident, t[f"{prefix}_und_rowptr"], ei32[1, :E].contiguous(),
                                  ei32[0, :E].contiguous(), ei32[0, :E].contiguous(), t[f"{prefix}_und_rowptr"], ident)
                und.dst_only = True    # src_ptr / src_idx above are placeholders: segment_sum never reads them
                lap = ops.EdgeCSR(N, N, E + N, t[f"{prefix}_perm"], t[f"{prefix}_rowptr"], t[f"{prefix}_ssrc"],
                                  t[f"{prefix}_sdst"], ei32[0].contiguous(), t[f"{prefix}_srcptr"], t[f"{prefix}_srcidx"])
                enc[role] = PrebuiltGraph(ei, E, und, lap, k=m.get(f"{prefix}_k"))
            g["enc_graphs"] = enc
            self._graph = g
        return self._graph


def collate_records(recs: List[Dict[str, np.ndarray]], pin: bool = True, lap_sign_rng=None) -> HostBatch:
    """``lap_sign_rng`` (a numpy RandomState): multiply every complex's stored Laplacian eigenvectors by a fresh random
    sign per column, the draw dgl.lap_pe makes on each forward of the reference (GAN.py:70,77; cpromg.random_signs).
    None leaves the stored orientation (deterministic tests and benches)."""
    a = _Arena()
    cat = lambda name: np.concatenate([r[name] for r in recs])
    for name in ("z_p", "z_l", "bits_p", "bits_l", "pos_p", "pos_l"):
        a.add(name, cat(name))
    for name in ("lap_p", "lap_l"):
        if lap_sign_rng is None:
            a.add(name, cat(name))
        else:
            from .cpromg import random_signs
            a.add(name, np.concatenate([r[name] * random_signs(r[name].shape[1], lap_sign_rng) for r in recs]))
    n = {"p": np.array([len(r["z_p"]) for r in recs]), "l": np.array([len(r["z_l"]) for r in recs])}
    off = {s: np.concatenate(([0], np.cumsum(v))) for s, v in n.items()}
    a.add("ptr_p", off["p"].astype(np.int32))
    a.add("ptr_l", off["l"].astype(np.int32))
    meta = dict(n_p=int(off["p"][-1]), n_l=int(off["l"][-1]), batch_size=len(recs), max_p=int(n["p"].max()),
                max_l=int(n["l"].max()), edges={})
    for k in EDGE_KEYS:
        s, d = _NODE_OF[k]
        ne = np.array([len(r[f"{k}_src"]) for r in recs])
        eoff = np.concatenate(([0], np.cumsum(ne)))
        a.add(f"{k}_src", np.concatenate([r[f"{k}_src"] + off[s][i] for i, r in enumerate(recs)]).astype(np.int32))
        a.add(f"{k}_dst", np.concatenate([r[f"{k}_dst"] + off[d][i] for i, r in enumerate(recs)]).astype(np.int32))
        rowptr = np.concatenate([r[f"{k}_rowptr"][:-1] + eoff[i] for i, r in enumerate(recs)] + [eoff[-1:]])
        srcptr = np.concatenate([r[f"{k}_srcptr"][:-1] + eoff[i] for i, r in enumerate(recs)] + [eoff[-1:]])
        a.add(f"{k}_rowptr", rowptr.astype(np.int32))
        a.add(f"{k}_srcptr", srcptr.astype(np.int32))
        a.add(f"{k}_srcidx", np.concatenate([r[f"{k}_srcidx"] + eoff[i] for i, r in enumerate(recs)]).astype(np.int32))
        if k not in SORTED_KEYS:
            a.add(f"{k}_perm", np.concatenate([r[f"{k}_perm"] + eoff[i] for i, r in enumerate(recs)]).astype(np.int32))
            a.add(f"{k}_ssrc", np.concatenate([r[f"{k}_ssrc"] + off[s][i] for i, r in enumerate(recs)]).astype(np.int32))
            a.add(f"{k}_sdst", np.concatenate([r[f"{k}_sdst"] + off[d][i] for i, r in enumerate(recs)]).astype(np.int32))
        meta[f"{k}_rowptr_host"] = rowptr.astype(np.int64)
        meta["edges"][k] = int(eoff[-1])
    for prefix, side, _k in ENCODER_GRAPHS:
        if f"{prefix}_und" not in recs[0]:
            continue
        ne = np.array([r[f"{prefix}_und"].shape[1] for r in recs])
        eoff = np.concatenate(([0], np.cumsum(ne)))
        E, N = int(eoff[-1]), int(off[side][-1])
        soff = eoff[:-1] + off[side][:-1]                       # offset of a complex in sorted-position space
        und = np.concatenate([r[f"{prefix}_und"] + off[side][i] for i, r in enumerate(recs)], axis=1)
        loops = np.arange(N, dtype=np.int64)
        a.add(f"{prefix}_ei", np.concatenate((und, np.stack((loops, loops))), axis=1).astype(np.int32))
        a.add(f"{prefix}_und_rowptr", np.concatenate([r[f"{prefix}_und_rowptr"][:-1] + eoff[i] for i, r in enumerate(recs)]
                                                     + [eoff[-1:]]).astype(np.int32))
        perm = []
        for i, r in enumerate(recs):   # stored ids: undirected pairs first (all complexes), then the loops
            pc = r[f"{prefix}_perm"].astype(np.int64)
            perm.append(np.where(pc < ne[i], pc + eoff[i], pc - ne[i] + E + off[side][i]))
        a.add(f"{prefix}_perm", np.concatenate(perm).astype(np.int32))
        for name in ("rowptr", "srcptr"):
            a.add(f"{prefix}_{name}", np.concatenate([r[f"{prefix}_{name}"][:-1] + soff[i] for i, r in enumerate(recs)]
                                                     + [np.array([E + N])]).astype(np.int32))
        a.add(f"{prefix}_srcidx", np.concatenate([r[f"{prefix}_srcidx"] + soff[i] for i, r in enumerate(recs)]).astype(np.int32))
        for name in ("ssrc", "sdst"):
            a.add(f"{prefix}_{name}", np.concatenate([r[f"{prefix}_{name}"] + off[side][i] for i, r in enumerate(recs)]).astype(np.int32))
        meta[f"{prefix}_und_edges"] = E
        ks = {int(r.get(f"{prefix}_k", _k)) for r in recs}   # records written before the field existed used the default
        assert len(ks) == 1, f"records of one batch carry encoder graphs built with different k: {sorted(ks)}"
        meta[f"{prefix}_k"] = ks.pop()
    if "smi_in" in recs[0]:
        for name in LIGAND_SCALARS:
            a.add(name, np.array([r[name] for r in recs], dtype=np.float32))
        a.add("smi_in", np.stack([r["smi_in"] for r in recs]).astype(np.int64))
        a.add("smi_tgt", np.stack([r["smi_tgt"] for r in recs]).astype(np.int64))
    return HostBatch(a, a.build(pin), meta)


class BatchLoader:
    """Iterates a dataset shard in fixed-size batches, endlessly, shuffling per epoch with a seeded generator; the
    iterator (and its state) lives as long as the loader.  ``indices`` = this rank's shard (parallel.shard_by_edges)."""

    def __init__(self, dataset: ComplexDataset, batch_size: int, indices: Optional[Iterable[int]] = None,
                 shuffle: bool = True, seed: int = 0, pin: bool = True, prefetch: int = 0, lap_sign: bool = True):
        self.ds, self.bs, self.shuffle, self.pin = dataset, batch_size, shuffle, pin
        # fresh Laplacian-eigenvector signs per batch (the reference's per-forward dgl draw), from a seeded stream
        self.sign_rng = np.random.RandomState((seed + 0x5151) % (2 ** 32)) if lap_sign else None
        self.indices = list(range(len(dataset))) if indices is None else list(indices)
        self.rng = np.random.RandomState(seed)
        self._order, self._pos = [], 0
        self._queue = None
        if prefetch > 0:   # one background thread collates the next batches while the step runs (numpy drops the GIL)
            import queue
            import threading
            self._queue = queue.Queue(maxsize=prefetch)

            def work():
                while True:
                    self._queue.put(self._make())
            threading.Thread(target=work, daemon=True).start()

    def __iter__(self):
        return self

    def __next__(self) -> HostBatch:
        return self._queue.get() if self._queue is not None else self._make()

    def _make(self) -> HostBatch:
        take = []
        while len(take) < self.bs:
            if self._pos >= len(self._order):
                self._order = list(self.rng.permutation(self.indices)) if self.shuffle else list(self.indices)
                self._pos = 0
            take.append(self._order[self._pos])
            self._pos += 1
        return collate_records([self.ds[i] for i in take], self.pin, self.sign_rng)


def synthetic_records(seeds: Iterable[int], tgt_len: int = 200, vocab: int = 116, pad: int = 110, n_protein=None):
    """Records of synthetic CrossDocked-shaped complexes (synthetic.make_complex) with synthetic ligand data."""
    from .cpromg import LapPECache
    cache = LapPECache(8)
    recs = []
    for s in seeds:
        g = syn.make_complex(s, n_protein=n_protein)
        rs = np.random.RandomState(77000 + s)
        length = int(rs.randint(min(20, tgt_len // 2), min(80, tgt_len)))
        toks = rs.randint(1, vocab, size=length + 1)
        toks[toks == pad] = 1
        smi = np.full(tgt_len + 1, pad, dtype=np.int64)
        smi[:length + 1] = toks
        lig = dict(sas=rs.uniform(1, 7), logP=rs.uniform(-2, 6), qed=rs.uniform(0, 1), weight=rs.uniform(150, 600),
                   tpsa=rs.uniform(20, 150), vina_score=rs.uniform(-11, -4), smiIndices_input=smi[:-1],
                   smiIndices_tgt=smi[1:])
        recs.append(pack_complex(g, lig, key=("syn", s), lap_cache=cache))
    return recs
