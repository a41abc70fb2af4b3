"""On-disk format and loader on the host: records carry destination-sorted CSRs, packed feature
codes and cached Laplacian encodings; collation needs no sort."""
import os

import numpy as np
import torch

from oracle import ef_oracle as eo
from singa_b200 import dataset as ds
from singa_b200 import synthetic as syn
from singa_b200.embedding import pack_feature_bits


def _graphs():
    return [syn.make_complex(s, n_protein=50 + 7 * s, n_ligand=9 + s) for s in range(3)]


def test_record_fields_match_the_oracle_csr_and_bit_packing(tmp_path):
    g = _graphs()[1]
    rec = ds.pack_complex(g, key=1)
    for k, (s, d) in ds._NODE_OF.items():
        n = {"p": len(g["z_p"]), "l": len(g["z_l"])}
        ref = eo.csr_from_edge_index(g["ei_" + k], n[s], n[d])
        assert np.array_equal(rec[f"{k}_rowptr"], ref["rowptr"])
        if k in ds.SORTED_KEYS:
            assert np.array_equal(rec[f"{k}_orig"], ref["perm"])
            assert np.array_equal(rec[f"{k}_src"], ref["src_sorted"]) and np.array_equal(rec[f"{k}_dst"], ref["dst_sorted"])
        else:   # interaction lists keep their input order (Wigner reuse pairs lp and pl edge for edge)
            assert np.array_equal(rec[f"{k}_src"], g["ei_" + k][0].numpy()) and np.array_equal(rec[f"{k}_perm"], ref["perm"])
            assert np.array_equal(rec[f"{k}_ssrc"], ref["src_sorted"]) and np.array_equal(rec[f"{k}_sdst"], ref["dst_sorted"])
        assert np.array_equal(rec[f"{k}_srcptr"], ref["src_ptr"]) and np.array_equal(rec[f"{k}_srcidx"], ref["src_idx"])
    assert np.array_equal(rec["bits_p"], pack_feature_bits(g["x_p"]).numpy())
    assert rec["lap_p"].shape == (len(g["z_p"]), 8) and rec["lap_l"].shape == (len(g["z_l"]), 8)
    path = os.path.join(tmp_path, "c1.npz")
    ds.save_record(path, rec)
    back = ds.load_record(path)
    assert set(back) == set(rec) and all(np.array_equal(back[k], rec[k]) for k in rec)
    # 59 float columns per atom shrink to one int32 code; edges to two int32 (+ the CSR)
    assert rec["bits_p"].nbytes * 59 == g["x_p"].numel() * 4


def test_collation_is_a_sorted_disjoint_union_without_a_sort():
    graphs = _graphs()
    recs = [ds.pack_complex(g, key=i, encoder_graphs=False) for i, g in enumerate(graphs)]
    hb = ds.collate_records(recs, pin=False)
    t = hb.arena.views(hb.host)
    big = syn.collate(graphs)
    n = {"p": hb.meta["n_p"], "l": hb.meta["n_l"]}
    assert np.array_equal(t["ptr_p"].numpy(), big["ptr_p"].numpy())
    for k, (s, d) in ds._NODE_OF.items():
        ref = eo.csr_from_edge_index(big["ei_" + k], n[s], n[d])      # the sort the loader avoids
        sk = ("src", "dst") if k in ds.SORTED_KEYS else ("ssrc", "sdst")
        assert np.array_equal(t[f"{k}_{sk[0]}"].numpy(), ref["src_sorted"]), k
        assert np.array_equal(t[f"{k}_{sk[1]}"].numpy(), ref["dst_sorted"]), k
        if k not in ds.SORTED_KEYS:
            assert np.array_equal(t[f"{k}_perm"].numpy(), ref["perm"]), k
            assert np.array_equal(t[f"{k}_src"].numpy(), big["ei_" + k][0].numpy()), k
        assert np.array_equal(t[f"{k}_rowptr"].numpy(), ref["rowptr"]), k
        assert np.array_equal(t[f"{k}_srcptr"].numpy(), ref["src_ptr"]), k
        assert np.array_equal(t[f"{k}_srcidx"].numpy(), ref["src_idx"]), k
        assert np.array_equal(hb.meta[f"{k}_rowptr_host"], ref["rowptr"])
    assert hb.nbytes < sum(v.numel() * v.element_size() for v in big.values())   # smaller than the raw batch


def test_batch_loader_keeps_its_iterator_and_shards():
    recs = ds.synthetic_records(range(5), n_protein=40, tgt_len=10)
    data = ds.ComplexDataset(recs)
    assert len(data.edge_counts()) == 5
    loader = ds.BatchLoader(data, 2, indices=[0, 2, 4], seed=1, pin=False)
    seen = [next(loader).meta["batch_size"] for _ in range(4)]       # 8 draws from a 3-complex shard: wraps epochs
    assert seen == [2, 2, 2, 2]
    hb = next(loader)
    t = hb.arena.views(hb.host)
    assert t["smi_in"].shape == (2, 10) and t["smi_tgt"].shape == (2, 10) and t["vina_score"].shape == (2,)
    assert int(t["smi_in"].max()) < 116


def test_encoder_graphs_in_the_record_equal_the_on_the_fly_construction():
    """The CProMG encoders' graphs (knn_graph -> to_undirected -> get_laplacian, CProMG.py:293-300, 330-334) depend only
    on positions; the record carries them, built on the host.  The collated batch graph and its centre-sorted CSR must
    equal what the loop restatements of the pyg / torch-cluster functions give on the batched positions."""
    from oracle import cpromg_oracle as co
    graphs = [syn.make_complex(s, n_protein=40 + 5 * s, n_ligand=7 + s) for s in range(3)]
    recs = [ds.pack_complex(g, key=i) for i, g in enumerate(graphs)]
    hb = ds.collate_records(recs, pin=False)
    t = hb.arena.views(hb.host)
    big = syn.collate(graphs)
    for prefix, side, k in ds.ENCODER_GRAPHS:
        pos = big["pos_" + side]
        n = pos.shape[0]
        batch = torch.repeat_interleave(torch.arange(3), torch.tensor([len(g["z_" + side]) for g in graphs]))
        di = co.knn_graph(pos, k, batch=batch, flow="target_to_source")
        ei, attr = co.to_undirected(di, (pos[di[0]] - pos[di[1]]).norm(dim=1), reduce="mean")
        lap_ei, _ = co.get_laplacian(ei, attr.unsqueeze(1), num_nodes=n)
        assert torch.equal(t[prefix + "_ei"].long(), lap_ei), prefix
        assert hb.meta[prefix + "_und_edges"] == ei.shape[1]
        assert hb.meta[prefix + "_k"] == k and int(recs[0][prefix + "_k"]) == k   # checked by the encoder on use
        ref = eo.csr_from_edge_index(lap_ei.flip(0), n, n)
        for a, b in (("perm", "perm"), ("rowptr", "rowptr"), ("ssrc", "src_sorted"), ("sdst", "dst_sorted"),
                     ("srcptr", "src_ptr"), ("srcidx", "src_idx")):
            assert np.array_equal(t[f"{prefix}_{a}"].numpy(), ref[b]), (prefix, a)
        und_rowptr = np.concatenate(([0], np.cumsum(np.bincount(ei[0].numpy(), minlength=n))))
        assert np.array_equal(t[prefix + "_und_rowptr"].numpy(), und_rowptr)


def test_prebuilt_encoder_graph_declares_its_k_and_its_destination_only_csr(tmp_path):
    """the record's encoder graphs are built with a fixed k; the batch carries that k so an encoder
    configured with another knn falls back to the device search instead of silently training on the record's graph,
    and a record round-trips the field through np.savez.  The undirected-part CSR is marked destination-only."""
    import pytest
    from singa_b200 import ops
    recs = [ds.pack_complex(syn.make_complex(s, n_protein=30, n_ligand=8), key=s) for s in range(2)]
    ds.save_record(str(tmp_path / "r.npz"), recs[0])
    back = ds.load_record(str(tmp_path / "r.npz"))
    assert int(back["ga_k"]) == 48 and int(back["gl_k"]) == 30
    old = [{k: v for k, v in r.items() if not k.endswith("_k")} for r in recs]      # records from before the field
    assert ds.collate_records(old, pin=False).meta["ga_k"] == 48
    mixed = [dict(recs[0]), dict(recs[1], ga_k=np.int32(16))]
    with pytest.raises(AssertionError):
        ds.collate_records(mixed, pin=False)
    i32 = lambda *v: torch.tensor(v, dtype=torch.int32)
    csr = ops.EdgeCSR(2, 2, 2, i32(0, 1), i32(0, 1, 2), i32(1, 0), i32(0, 1), i32(0, 1), i32(0, 1, 2), i32(0, 1),
                      dst_only=True)
    assert csr.index("dst_orig").dtype == torch.int64
    with pytest.raises(ValueError):
        csr.index("src_group")
