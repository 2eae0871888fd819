"""The CUDA path against the UNMODIFIED reference executed on the same machine's CPU (baseline/_ref,
installed by baseline/install_ref.py and shipped to the GPU box like the built .so).

BASELINE config C1 at full size: N = 4,096 x 1024, cosine-normalised rows, k = 8.  The reference's
build_hypergraph_knn_kmeans (preprocess_hypergraph.py:335-433) runs untouched; only the KMeans class
it instantiates is swapped for a stand-in, once so that no cliques are added (kNN + de-dup + weights)
and once with fixed labels (a deterministic clique set)."""
import os
import sys

import numpy as np
import pytest
import torch

import hg_b200 as hg
from hg_b200 import synth

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline"))
import ref_loader  # noqa: E402

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_loader.available(), reason="baseline/_ref not installed (python baseline/install_ref.py)")]


def _sorted(ei, w):
    ei = ei.cpu().numpy().astype(np.int64)
    w = w.cpu().numpy()
    order = np.lexsort((ei[1], ei[0]))
    return ei[:, order], w[order]


class _FixedLabels:
    labels = None

    def __init__(self, n_clusters=8, **kw):
        pass

    def fit_predict(self, x):
        return np.asarray(type(self).labels[: x.shape[0]])


@pytest.mark.parametrize("cliques", [False, True])
def test_c1_against_the_reference_run_here(cliques):
    pkg, pp, _ = ref_loader.load()
    n_wsi, n_tma, d, k = 3840, 256, 1024, 8
    n = n_wsi + n_tma
    feats = synth.make_features(n, d, 1235, "cosine")
    if cliques:
        num_h = 40                                             # ~100 rows per cluster: 4,096 * 99 / 2 = 2e5 clique pairs
        labels = np.random.default_rng(7).integers(0, num_h + 2, size=n)        # two label values out of range
        _FixedLabels.labels = labels
        pp.KMeans = _FixedLabels
    else:
        num_h, labels = 10, None
        pp.KMeans = ref_loader.NoCliqueKMeans
    r_ei, r_w, r_stats = pkg.build_hypergraph_knn_kmeans(torch.from_numpy(feats[:n_wsi]), torch.from_numpy(feats[n_wsi:]),
                                                         np.zeros(n_wsi, dtype=np.int64), k=k, num_hyperedges=num_h)
    r_ei, r_w = _sorted(r_ei, r_w)
    kw = dict(hyperedge_labels=labels) if cliques else dict(hyperedge_labels=np.full(n, 10 ** 6))
    ei, w, stats = hg.build_hypergraph_knn_kmeans(torch.from_numpy(feats[:n_wsi]).cuda(), torch.from_numpy(feats[n_wsi:]).cuda(),
                                                  np.zeros(n_wsi, dtype=np.int64), k=k, num_hyperedges=num_h, **kw)
    ei, w = ei.cpu().numpy(), w.cpu().numpy()
    assert stats["num_edges"] == r_stats["num_edges"] and stats["num_nodes"] == r_stats["num_nodes"] == n
    assert ei.shape == r_ei.shape and (ei == r_ei).all()              # bit-exact incidence against the reference itself
    np.testing.assert_allclose(w, r_w, rtol=1e-5, atol=1e-7)          # max(0, cos): 1e-5 relative
    print(f"C1 vs reference (cliques={cliques}): E = {ei.shape[1]}, max |dw| = {np.abs(w - r_w).max():.2e}")
