"""Generate tests/golden/*.npz by EXECUTING THE UNMODIFIED REFERENCE in the build
container -- test infrastructure only (see oracle/oracle_np.py header).

    python oracle/make_golden.py            # needs /root/reference; writes tests/golden/

The reference (pure Python, torch + scikit-learn) imports ``h5py`` at module
scope (build_hypergraph/preprocess_hypergraph.py:15) but only its I/O functions
use it; an empty stub module lets the package import.  ``KMeans`` inside
``preprocess_hypergraph`` is monkey-patched to return FIXED labels so that the
clique stage (:391-400) is deterministic: labels >= num_hyperedges produce no
cliques (pure kNN oracle), labels in range produce a known clique set.
Nothing else of the reference is touched.  /root/reference does not exist on
the GPU box, hence the committed fixtures.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("HG_REFERENCE_ROOT", "/root/reference")
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))


def import_reference():
    if "h5py" not in sys.modules:
        try:
            import h5py  # noqa: F401
        except ModuleNotFoundError:
            sys.modules["h5py"] = types.ModuleType("h5py")
    sys.path.insert(0, REF)
    import build_hypergraph as ref  # the reference package
    from build_hypergraph import preprocess_hypergraph as ref_pp
    assert ref.__file__.startswith(REF), ref.__file__
    return ref, ref_pp


class _FixedKMeans:
    """Stand-in for sklearn.cluster.KMeans with injected labels."""
    labels = None

    def __init__(self, n_clusters=8, **kw):
        self.n_clusters = n_clusters

    def fit_predict(self, x):
        return np.asarray(type(self).labels[: x.shape[0]])


def sorted_edges(edge_index, edge_weights):
    ei = edge_index.cpu().numpy().astype(np.int64)
    w = edge_weights.cpu().numpy()
    order = np.lexsort((ei[1], ei[0]))
    return ei[:, order], w[order]


def main():
    import torch
    from hg_b200 import synth

    torch.manual_seed(0)
    ref, ref_pp = import_reference()
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)

    # ---- dense kernels (A1, A2, A7, A8, A9) --------------------------------
    n, d, seed = 192, 64, 11
    feats, pos = synth.make_slide(n, d, seed, "gaussian")
    lam_h, lam_g = 1.0, 1.0 / 128.0
    tf, tp = torch.from_numpy(feats), torch.from_numpy(pos)
    K_h = ref.compute_morphological_similarity(tf, lam_h).numpy()
    K_g = ref.compute_spatial_similarity(tp, lam_g).numpy()
    K = ref.compute_combined_similarity(tf, tp, lam_h, lam_g).numpy()
    tma = synth.make_features(80, d, seed + 1, "gaussian")
    sim, sim_stats = ref.compute_wsi_tma_similarity(tf[:100], tp[:100], torch.from_numpy(tma), lam_h, lam_g)
    nt = 48   # N^2 .item() calls in the reference: keep it small
    thr_ei, thr_w = ref.build_weighted_hypergraph(tf[:nt], tp[:nt], lam_h, lam_g, threshold_median_ratio=1.0)
    thr2_ei, thr2_w = ref.build_weighted_hypergraph(tf[:nt], tp[:nt], lam_h, lam_g, threshold_median_ratio=0.5)
    pooled = ref.mean_pool_with_similarity(tf).numpy()
    np.savez_compressed(
        os.path.join(out_dir, "dense_n192_d64.npz"),
        n=n, d=d, seed=seed, lambda_h=lam_h, lambda_g=lam_g, n_tma=80, n_thr=nt,
        feat_sum=np.float64(feats.astype(np.float64).sum()),
        K_h=K_h, K_g=K_g, K=K, wsi_tma=sim.numpy(),
        wsi_tma_stats=np.array([sim_stats[s] for s in ("mean", "std", "min", "max", "median")]),
        K_stats=np.array([float(torch.from_numpy(K).mean()), float(torch.from_numpy(K).std()),
                          float(K.min()), float(K.max()), float(torch.from_numpy(K).median())]),
        thr_edge_index=thr_ei.numpy(), thr_edge_weights=thr_w.numpy(),
        thr2_edge_index=thr2_ei.numpy(), thr2_edge_weights=thr2_w.numpy(),
        pooled=pooled)

    # ---- kNN + dedup + cosine weights (A4, A5, A6), kNN-only and with cliques
    ref_pp.KMeans = _FixedKMeans
    for name, (n_wsi, n_tma, d, k, mode, seed, hyper) in {
        "knn_n512_d128_k8": (400, 112, 128, 8, "cosine", 21, 0),
        "knn_n768_d256_k5_cliques": (600, 168, 256, 5, "gaussian", 22, 6),
        "knn_n1024_d1024_k8": (896, 128, 1024, 8, "cosine", 23, 0),
        # BASELINE config C1 at full size (N = 4,096 x 1024, cosine kNN k = 8, the bench generator's seed)
        "knn_c1_n4096_d1024_k8": (3840, 256, 1024, 8, "cosine", 1235, 0),
    }.items():
        n = n_wsi + n_tma
        feats = synth.make_features(n, d, seed, mode)
        rng = np.random.default_rng(seed)
        if hyper:
            labels = rng.integers(0, hyper + 3, size=n)   # some labels out of range
            num_h = hyper
        else:
            labels = np.full(n, 10 ** 6)
            num_h = 10
        _FixedKMeans.labels = labels
        ei, w, stats = ref.build_hypergraph_knn_kmeans(
            torch.from_numpy(feats[:n_wsi]), torch.from_numpy(feats[n_wsi:]),
            np.zeros(n_wsi, dtype=np.int64), k=k, num_hyperedges=num_h)
        ei, w = sorted_edges(ei, w)
        # the raw sklearn neighbour lists, exactly as the reference asks for them (:379-382)
        from sklearn.neighbors import NearestNeighbors
        dist, idx = NearestNeighbors(n_neighbors=k + 1, metric="euclidean").fit(feats).kneighbors(feats)
        np.savez_compressed(
            os.path.join(out_dir, name + ".npz"),
            n_wsi=n_wsi, n_tma=n_tma, d=d, k=k, mode=mode, seed=seed,
            num_hyperedges=num_h, labels=labels.astype(np.int64),
            feat_sum=np.float64(feats.astype(np.float64).sum()),
            edge_index=ei, edge_weights=w.astype(np.float32),
            num_edges=stats["num_edges"],
            sk_indices=idx.astype(np.int32), sk_distances=dist.astype(np.float64))
        print(name, "E =", ei.shape[1])
    # ---- super-patch aggregation ((f)4): reference :87-199 with injected cluster labels ------
    n, d, seed, s_super = 600, 96, 31, 12
    feats, pos = synth.make_slide(n, d, seed, "gaussian")
    rng = np.random.default_rng(seed)
    labels = rng.integers(0, s_super - 1, size=n)
    labels[17] = s_super - 1                               # one single-member cluster (no intra term)
    _FixedKMeans.labels = labels
    lam_h, lam_g = 0.5, 1.0 / 256.0
    sf, sp, st, K = ref_pp.aggregate_wsi_super_patches(torch.from_numpy(feats), torch.from_numpy(pos), s_super,
                                                       lam_h, lam_g)
    np.savez_compressed(
        os.path.join(out_dir, "aggregate_n600_d96_s12.npz"),
        n=n, d=d, seed=seed, num_super_patches=s_super, lambda_h=lam_h, lambda_g=lam_g,
        labels=labels.astype(np.int64), feat_sum=np.float64(feats.astype(np.float64).sum()),
        super_features=sf.numpy(), super_positions=sp.numpy(),
        avg_intra=np.float64(st["avg_intra_cluster_similarity"]),
        K_stats=np.array([st["wsi_similarity_matrix_stats"][q] for q in ("mean", "std", "min", "max", "median")]))
    print("aggregate_n600_d96_s12 avg_intra =", st["avg_intra_cluster_similarity"])
    print("golden fixtures written to", out_dir)


if __name__ == "__main__":
    main()
