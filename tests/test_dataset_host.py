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
    recs = [ds.pack_complex(g, key=i) for i, g in enumerate(graphs)]
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
