"""Row (f)4: device-side super-patch aggregation (segmented means + intra-cluster similarity) against
the reference's own output (tests/golden/aggregate_*.npz) and the oracle restatement of
preprocess_hypergraph.py:157-190 on seeded inputs."""
import numpy as np
import pytest
import torch

import hg_b200
import oracle_np as oracle
from hg_b200 import ops, synth
from hg_b200 import preprocess_hypergraph as pp
from test_gpu_parity import _cuda

pytestmark = pytest.mark.gpu


def _check(agg, ref, s):
    sf, sp, counts, intra, avg = ref
    np.testing.assert_allclose(agg["features"].cpu().numpy(), sf, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(agg["positions"].cpu().numpy(), sp, rtol=1e-5, atol=1e-5)
    assert np.array_equal(agg["counts"].cpu().numpy(), counts)                      # integer work: exact
    got = agg["intra"].cpu().numpy().astype(np.float64)
    assert np.array_equal(np.isnan(got), np.isnan(intra))
    ok = ~np.isnan(intra)
    np.testing.assert_allclose(got[ok], intra[ok], rtol=1e-5)
    summary = agg["summary"].tolist()
    assert abs(summary[0] - avg) <= 1e-5 * max(abs(avg), 1e-30)
    assert summary[1] == ok.sum() and summary[2] == -1 and summary[3] == counts.sum()


def test_aggregate_matches_reference_golden(golden):
    g = golden("aggregate_n600_d96_s12")
    feats, pos = synth.make_slide(int(g["n"]), int(g["d"]), int(g["seed"]), "gaussian")
    s, lh, lg = int(g["num_super_patches"]), float(g["lambda_h"]), float(g["lambda_g"])
    for dense in (False, True):
        K = hg_b200.compute_combined_similarity(_cuda(feats), _cuda(pos), lh, lg) if dense else None
        agg = ops.aggregate_super_patches(_cuda(feats), _cuda(pos), g["labels"], s, lh, lg, dense_k=K)
        np.testing.assert_allclose(agg["features"].cpu().numpy(), g["super_features"], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(agg["positions"].cpu().numpy(), g["super_positions"], rtol=1e-5, atol=1e-5)
        assert abs(agg["summary"][0].item() - float(g["avg_intra"])) <= 1e-5 * float(g["avg_intra"])


@pytest.mark.parametrize("n,d,s,pd,seed", [(777, 50, 9, 2, 1), (4096, 128, 100, 2, 2), (3000, 67, 300, 3, 3),
                                           (2500, 1024, 10, 2, 4), (300, 32, 1, 2, 5)])
def test_aggregate_matches_oracle(n, d, s, pd, seed):
    feats, pos2 = synth.make_slide(n, d, seed, "gaussian")
    rng = np.random.default_rng(seed)
    pos = pos2 if pd == 2 else np.concatenate([pos2, rng.integers(0, 4, size=(n, 1)).astype(np.float32) * 64], 1)
    labels = rng.permutation(np.arange(n) % s)                 # every cluster has members
    if s >= 9:
        labels[labels == s - 1] = s - 2                        # ... except a merged one, then re-seeded with a single row
        labels[5] = s - 1
    lh, lg = 0.5, 1.0 / 512.0
    ref = oracle.aggregate_super_patches(feats, pos, labels, s, lambda_h=lh, lambda_g=lg)
    agg = ops.aggregate_super_patches(_cuda(feats), _cuda(pos), labels, s, lh, lg, return_order=True)
    _check(agg, ref, s)
    order = agg["order"].cpu().numpy()
    assert np.array_equal(order, np.lexsort((np.arange(n), labels)))               # stable sort by label
    K = hg_b200.compute_combined_similarity(_cuda(feats), _cuda(pos), lh, lg)
    _check(ops.aggregate_super_patches(_cuda(feats), _cuda(pos), labels, s, lh, lg, dense_k=K), ref, s)
    # deterministic: bit-identical on a second run
    again = ops.aggregate_super_patches(_cuda(feats), _cuda(pos), labels, s, lh, lg)
    assert torch.equal(again["features"], agg["features"]) and torch.equal(again["summary"], agg["summary"])


def test_aggregate_empty_cluster_and_stray_labels():
    n, d, s = 500, 40, 6
    feats, pos = synth.make_slide(n, d, 9, "gaussian")
    labels = np.arange(n) % s
    labels[labels == 2] = 4                                     # cluster 2 empty
    agg = ops.aggregate_super_patches(_cuda(feats), _cuda(pos), labels, s, 1.0, 0.0)
    assert agg["summary"][2].item() == 2 and agg["counts"][2].item() == 0
    assert torch.isnan(agg["features"][2]).all()
    labels = np.arange(n) % s
    labels[:7] = [-1, s, s + 5, 10 ** 9, -7, s, s]              # rows outside [0, S) are left out
    keep = (labels >= 0) & (labels < s)
    ref = oracle.aggregate_super_patches(feats[keep], pos[keep], labels[keep], s, lambda_h=1.0, lambda_g=0.0)
    agg = ops.aggregate_super_patches(_cuda(feats), _cuda(pos), labels, s, 1.0, 0.0)
    _check(agg, ref, s)
    with pytest.raises(ValueError):
        ops.aggregate_super_patches(_cuda(feats), _cuda(pos), labels[:-1], s)


def test_aggregate_wsi_super_patches_api(monkeypatch):
    """The reference-facing function: same outputs with and without the dense kernel (large-N path)."""
    n, d, s = 1500, 64, 20
    feats, pos = synth.make_slide(n, d, 12, "gaussian")
    labels = np.random.default_rng(3).permutation(np.arange(n) % s)
    monkeypatch.setattr(pp, "_cluster", lambda x, k: labels)
    lh, lg = 0.25, 1.0 / 1024.0
    sf, sp, st, K = hg_b200.aggregate_wsi_super_patches(_cuda(feats), _cuda(pos), s, lh, lg)
    ref = oracle.aggregate_super_patches(feats, pos, labels, s, lambda_h=lh, lambda_g=lg)
    np.testing.assert_allclose(sf.cpu().numpy(), ref[0], rtol=1e-5, atol=1e-6)
    assert abs(st["avg_intra_cluster_similarity"] - ref[4]) <= 1e-5 * ref[4]
    assert K is not None and K.shape == (n, n)
    monkeypatch.setattr(pp, "MAX_DENSE_N", 1000)
    sf2, sp2, st2, K2 = hg_b200.aggregate_wsi_super_patches(_cuda(feats), _cuda(pos), s, lh, lg)
    assert K2 is None and torch.equal(sf2, sf) and torch.equal(sp2, sp)
    assert abs(st2["avg_intra_cluster_similarity"] - ref[4]) <= 1e-5 * ref[4]
    bad = labels.copy()
    bad[bad == 7] = 8
    monkeypatch.setattr(pp, "_cluster", lambda x, k: bad)
    with pytest.raises(ValueError, match="Cluster 7 is empty"):
        hg_b200.aggregate_wsi_super_patches(_cuda(feats), _cuda(pos), s, lh, lg)
