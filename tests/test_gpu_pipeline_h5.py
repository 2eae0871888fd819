"""The per-slide pipeline of the drop-in (process_single_file / rebuild / CSV loop) on the GPU,
with an in-memory h5py stand-in (h5py is not installed in this image)."""
import json
import os
import sys

import numpy as np
import pytest
import torch

import fake_h5py
import oracle_np as oracle
from hg_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture()
def h5(monkeypatch):
    fake_h5py.reset()
    monkeypatch.setitem(sys.modules, "h5py", fake_h5py)
    return fake_h5py


def _make_slide_file(h5, path, n_wsi=600, n_tma=90, d=64, seed=5, with_tma=True):
    feats, pos = synth.make_slide(n_wsi, d, seed, "gaussian")
    with h5.File(path, "w") as f:
        g = f.create_group("wsi")
        g["features"] = feats
        g["positions"] = pos
        if with_tma:
            f.create_group("tma")["features"] = synth.make_features(n_tma, d, seed + 1, "gaussian")
    return feats, pos


def test_process_single_file_end_to_end(h5, tmp_path):
    import hg_b200 as hg
    path = str(tmp_path / "case0.h5")
    feats, pos = _make_slide_file(h5, path)
    stats = hg.process_single_file(path, num_wsi_super_patches=20, num_groups=4, hypergraph_k=5, num_hyperedges=3,
                                   lambda_h=1.0, lambda_g=1.0 / 128)
    assert set(stats) == {"wsi_aggregation", "similarity", "grouping", "hypergraph"}
    with h5.File(path, "r") as f:
        hgp = f["hypergraph"]
        sup = hgp["wsi_super"]["features"][:]
        tma = hgp["tma"]["features"][:]
        ei = hgp["edge_index"][:]
        w = hgp["edge_weights"][:]
        assert sup.shape == (20, 64) and hgp["wsi_super"]["positions"][:].shape == (20, 2)
        assert ei.dtype == np.int64 and ei.shape[0] == 2 and w.dtype == np.float32 and w.shape == (ei.shape[1],)
        assert hgp["group_labels"][:].shape == (20,)
        K = hgp["similarity"]["wsi_internal"][:]
        assert K.shape == (600, 600)
        np.testing.assert_allclose(K, oracle.combined_similarity(feats, pos, 1.0, 1.0 / 128), rtol=1e-5)
        sim = hgp["similarity"]["wsi_tma"][:]
        np.testing.assert_allclose(sim, oracle.wsi_tma_similarity(sup, tma, 1.0), rtol=1e-5)
        saved = json.loads(hgp.attrs["stats"])
        assert saved["hypergraph"]["num_edges"] == ei.shape[1] == stats["hypergraph"]["num_edges"]
    # K statistics (A9) against the oracle on the stored matrix
    ref = oracle.similarity_stats(K)
    for key, val in stats["wsi_aggregation"]["wsi_similarity_matrix_stats"].items():
        assert np.isclose(val, ref[key], rtol=1e-5), key
    # the edge stage given the stored nodes: kNN part must be contained, weights must match
    allf = np.concatenate([sup, tma], 0)
    idx, _ = oracle.knn_feature(allf, 5, "euclidean")
    knn_edges = set(map(tuple, oracle.edges_from_knn(idx).T.tolist()))
    got = set(map(tuple, ei.T.tolist()))
    assert knn_edges <= got and (ei[0] < ei[1]).all()
    np.testing.assert_allclose(w, oracle.edge_weights_cosine(allf, ei), rtol=1e-5, atol=1e-6)
    # second run on the same file: the reference would fail on the existing datasets, we overwrite
    hg.process_single_file(path, num_wsi_super_patches=10, num_groups=3, hypergraph_k=4, num_hyperedges=2)
    with h5.File(path, "r") as f:
        assert f["hypergraph"]["wsi_super"]["features"][:].shape == (10, 64)


def test_skip_without_tma_and_rebuild(h5, tmp_path):
    import hg_b200 as hg
    p0 = str(tmp_path / "no_tma.h5")
    _make_slide_file(h5, p0, with_tma=False)
    assert hg.process_single_file(p0) == {"status": "skipped", "reason": "no_tma"}       # reference :560-562
    p1 = str(tmp_path / "case1.h5")
    _make_slide_file(h5, p1, n_wsi=300, seed=9)
    hg.process_single_file(p1, num_wsi_super_patches=12, num_groups=3, hypergraph_k=4, num_hyperedges=2)
    stats = hg.rebuild_hypergraph_from_similarity(p1, num_wsi_super_patches=8, num_groups=2, hypergraph_k=3,
                                                  num_hyperedges=2, threshold_median_ratio=0.5)
    assert stats["hypergraph"]["num_edges_after_threshold"] <= stats["hypergraph"]["num_edges"]
    stats2 = hg.rebuild_hypergraph_from_similarity(p1, hypergraph_k=3, num_hyperedges=2)   # reuse stored stages
    assert stats2["grouping"]["method"] == "existing"
    with pytest.raises(ValueError):
        hg.rebuild_hypergraph_from_similarity(p0)                                          # TMA features not found


def test_process_dataset_csv_loop(h5, tmp_path, monkeypatch):
    import hg_b200 as hg
    root = str(tmp_path)
    names = ["a.h5", "b.h5", "missing.h5"]
    for i, nme in enumerate(names[:2]):
        _make_slide_file(h5, os.path.join(root, nme), n_wsi=200 + 50 * i, seed=20 + i)
    # a.h5 / b.h5 exist only in the in-memory store: make os.path.exists agree with it
    real_exists = os.path.exists
    monkeypatch.setattr(os.path, "exists", lambda p: p in fake_h5py._FILES or real_exists(p))
    csv = tmp_path / "cases.csv"
    csv.write_text("case_id,h5_file_path\n" + "\n".join(f"c{i},{n}" for i, n in enumerate(names)) + "\n")
    out = tmp_path / "stats.json"
    res = hg.process_dataset(str(csv), root, num_wsi_super_patches=10, num_groups=3, hypergraph_k=3,
                             num_hyperedges=2, output_stats_path=str(out))
    assert [r["case_id"] for r in res] == ["c0", "c1"]           # the missing file is reported and skipped
    assert len(json.loads(out.read_text())) == 2
    bad = tmp_path / "bad.csv"
    bad.write_text("case_id,path\nx,y\n")
    with pytest.raises(ValueError):
        hg.process_dataset(str(bad), root)


def test_pipeline_on_npz_container_matches_h5(h5, tmp_path, monkeypatch):
    """(f)2: the same slide through the .npz container on disk and through h5py gives the same outputs;
    beyond the dense limit the top-k lists of K_wsi are stored instead of the N x N matrix."""
    import hg_b200 as hg
    from hg_b200 import preprocess_hypergraph as pp, store
    p_h5, p_npz = str(tmp_path / "case.h5"), str(tmp_path / "case.npz")
    feats, pos = _make_slide_file(h5, p_h5, n_wsi=500, seed=31)
    tma = synth.make_features(90, 64, 32, "gaussian")
    with store.open_store(p_npz, "w") as f:
        g = f.create_group("wsi")
        g["features"] = feats
        g["positions"] = pos
        f.create_group("tma")["features"] = tma
    kw = dict(num_wsi_super_patches=16, num_groups=4, hypergraph_k=5, num_hyperedges=3, lambda_h=1.0, lambda_g=1.0 / 128)
    s_h5 = hg.process_single_file(p_h5, **kw)
    s_npz = hg.process_single_file(p_npz, **kw)
    assert s_h5["hypergraph"] == s_npz["hypergraph"]
    with h5.File(p_h5, "r") as a, store.open_store(p_npz, "r") as b:
        for key in ("edge_index", "edge_weights", "group_labels"):
            assert np.array_equal(a["hypergraph"][key][:], b["hypergraph"][key][:]), key
        assert np.array_equal(a["hypergraph"]["similarity"]["wsi_internal"][:], b["hypergraph/similarity/wsi_internal"][:])
        assert b["hypergraph"]["similarity"].attrs["wsi_shape"] == [500, 500]
        assert json.loads(b["hypergraph"].attrs["stats"])["hypergraph"]["num_edges"] == s_npz["hypergraph"]["num_edges"]
        assert "wsi_topk_indices" not in b["hypergraph"]["similarity"]
    assert hg.load_similarity_topk(p_npz) is None
    # re-processing and rebuilding an already processed container replaces its datasets
    hg.rebuild_hypergraph_from_similarity(p_npz, num_wsi_super_patches=8, num_groups=2, hypergraph_k=3, num_hyperedges=2)
    with store.open_store(p_npz, "r") as b:
        assert b["hypergraph/wsi_super/features"].shape == (8, 64)
    # large-slide layout: no dense kernel on disk, top-k lists of K_wsi instead
    monkeypatch.setattr(pp, "MAX_DENSE_N", 400)
    monkeypatch.setattr(pp, "STORED_TOPK", 12)
    p_big = str(tmp_path / "big.npz")
    with store.open_store(p_big, "w") as f:
        g = f.create_group("wsi")
        g["features"] = feats
        g["positions"] = pos
        f.create_group("tma")["features"] = tma
    s_big = hg.process_single_file(p_big, **kw)
    with store.open_store(p_big, "r") as b:
        assert "wsi_internal" not in b["hypergraph"]["similarity"] and "wsi_tma" in b["hypergraph"]["similarity"]
    idx, val = hg.load_similarity_topk(p_big)
    assert idx.shape == (500, 12) and idx.dtype == torch.int64 and val.dtype == torch.float32
    K = oracle.combined_similarity(feats, pos, 1.0, 1.0 / 128)
    np.fill_diagonal(K, -np.inf)
    exp_val = -np.sort(-K, axis=1)[:, :12]
    np.testing.assert_allclose(val.numpy(), exp_val, rtol=1e-5)
    np.testing.assert_allclose(K[np.arange(500)[:, None], idx.numpy()], exp_val, rtol=1e-5)
    # streamed statistics of the kernel that was never stored, and the same edges as the dense run
    ref = oracle.similarity_stats(oracle.combined_similarity(feats, pos, 1.0, 1.0 / 128))
    got = s_big["wsi_aggregation"]["wsi_similarity_matrix_stats"]
    for key in ("mean", "std", "min", "max"):
        assert np.isclose(got[key], ref[key], rtol=1e-5, atol=1e-7), key
    assert np.isclose(s_big["wsi_aggregation"]["avg_intra_cluster_similarity"],
                      s_npz["wsi_aggregation"]["avg_intra_cluster_similarity"], rtol=1e-5)
    assert s_big["hypergraph"] == s_npz["hypergraph"]
