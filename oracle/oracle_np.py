"""CPU oracle for the hypergraph-construction hot path -- TEST INFRASTRUCTURE ONLY.

This file is a numpy restatement of the reference algorithm
(zz9tf/multimodal-fusion, ``build_hypergraph/``).  It is the checker for the
CUDA path: only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import it.  Nothing under
``multimodal-fusion_b200/`` imports it and the product path has no CPU fallback.

Pinning status: the reference ships NO tests / golden vectors for this path
(SURVEY.md section 4), so the oracle is pinned against outputs of the reference
itself, executed in the build container by ``oracle/make_golden.py`` and
committed under ``tests/golden/`` (``tests/test_oracle_golden.py`` checks every
function below against them).  The kNN arithmetic of the reference lives in
scikit-learn (pinned 1.3.2 by the reference, 1.9.0 here) -- that boundary has
no reference-side test, so kNN parity is "pinned to the reference as executed
here", not to a published vector.

Every function cites the reference file:line it follows (paths relative to the
reference root).
"""
from __future__ import annotations

import numpy as np

__all__ = [
    "morphological_similarity", "spatial_similarity", "combined_similarity",
    "wsi_tma_similarity", "similarity_stats", "sq_dists_f64", "knn_feature",
    "knn_spatial", "edges_from_knn", "edge_weights_cosine", "threshold_edges",
    "csr_from_coo", "build_hypergraph_knn", "lower_median", "hypergraph_conv_propagate",
    "aggregate_super_patches", "kmeans_fit",
]


# --------------------------------------------------------------------------
# A1 / A2: dense Gaussian kernels
# --------------------------------------------------------------------------
def _gauss_expansion(x: np.ndarray, lam: float, dtype) -> np.ndarray:
    """exp(-lam * (|x_i|^2 + |x_j|^2 - 2 x_i.x_j)), un-clamped.

    Follows build_hypergraph/similarity_kernel.py:43-52 (features) and :79-84
    (positions): row norms, Gram matrix, expansion form, exp.  ``dtype``
    float32 reproduces the reference arithmetic; float64 is the "truth" used for
    the 1e-5 tolerance checks.
    """
    x = np.asarray(x, dtype=dtype)
    nrm = np.sum(x * x, axis=1, keepdims=True)
    dot = x @ x.T
    d2 = nrm + nrm.T - dtype(2) * dot
    return np.exp(-dtype(lam) * d2)


def morphological_similarity(features, lambda_h=1.0, dtype=np.float64):
    """similarity_kernel.py:17-54 ``compute_morphological_similarity``."""
    return _gauss_expansion(features, lambda_h, dtype)


def spatial_similarity(positions, lambda_g=1.0, dtype=np.float64):
    """similarity_kernel.py:57-86 ``compute_spatial_similarity``."""
    return _gauss_expansion(positions, lambda_g, dtype)


def combined_similarity(features, positions, lambda_h=1.0, lambda_g=1.0, dtype=np.float64):
    """similarity_kernel.py:88-124: K = K_h * K_g (product of two exps, :122)."""
    return (morphological_similarity(features, lambda_h, dtype)
            * spatial_similarity(positions, lambda_g, dtype))


# --------------------------------------------------------------------------
# A7: rectangular WSI x TMA kernel (direct-difference form)
# --------------------------------------------------------------------------
def wsi_tma_similarity(wsi_features, tma_features, lambda_h=1.0, dtype=np.float64):
    """preprocess_hypergraph.py:248-257: for every WSI row, sum((a-b)^2) over D
    then exp(-lambda_h * .).  Positions / lambda_g are accepted by the reference
    signature but unused (:245-246)."""
    a = np.asarray(wsi_features, dtype=dtype)
    b = np.asarray(tma_features, dtype=dtype)
    out = np.empty((a.shape[0], b.shape[0]), dtype=dtype)
    for i in range(a.shape[0]):
        diff = a[i:i + 1] - b
        out[i] = np.exp(-dtype(lambda_h) * np.sum(diff * diff, axis=1))
    return out


def lower_median(values: np.ndarray) -> float:
    """torch.median semantics: the LOWER of the two middle values for an even
    count (used at similarity_kernel.py:186, preprocess_hypergraph.py:195,264)."""
    v = np.sort(np.asarray(values).ravel())
    return float(v[(v.size - 1) // 2])


def similarity_stats(mat: np.ndarray) -> dict:
    """mean / std (unbiased, torch default) / min / max / lower-median of all
    entries: preprocess_hypergraph.py:190-196 and :259-265."""
    m = np.asarray(mat, dtype=np.float64).ravel()
    return {
        "mean": float(m.mean()),
        "std": float(m.std(ddof=1)) if m.size > 1 else float("nan"),
        "min": float(m.min()),
        "max": float(m.max()),
        "median": lower_median(m),
    }


# --------------------------------------------------------------------------
# A4: feature-space kNN (the sklearn brute-force path, restated)
# --------------------------------------------------------------------------
def sq_dists_f64(xq: np.ndarray, xall: np.ndarray) -> np.ndarray:
    """Squared Euclidean distances in float64 with sklearn's formula and clamp:
    d2 = |x|^2 - 2 x.y + |y|^2, max(0, .)  (scikit-learn
    metrics/_pairwise_distances_reduction/_middle_term_computer.pyx.tp and
    _argkmin.pyx.tp: float32 chunks are up-cast to float64 before the GEMM).
    This is what NearestNeighbors(metric='euclidean', algorithm->brute) ranks by
    at preprocess_hypergraph.py:379-382."""
    q = np.asarray(xq, dtype=np.float64)
    a = np.asarray(xall, dtype=np.float64)
    d2 = (q * q).sum(1)[:, None] - 2.0 * (q @ a.T) + (a * a).sum(1)[None, :]
    np.maximum(d2, 0.0, out=d2)
    return d2


def _topk_rows(score: np.ndarray, k: int):
    """k smallest per row, ordered by (score, index) ascending -- the
    deterministic tie rule the CUDA path uses."""
    n_cols = score.shape[1]
    k = min(k, n_cols)
    part = np.argpartition(score, k - 1, axis=1)[:, :k] if k < n_cols else \
        np.tile(np.arange(n_cols), (score.shape[0], 1))
    # argpartition is not tie-stable: widen to everything <= the k-th value
    kth = np.take_along_axis(score, part, 1).max(1)
    idx_out = np.empty((score.shape[0], k), dtype=np.int64)
    val_out = np.empty((score.shape[0], k), dtype=score.dtype)
    for r in range(score.shape[0]):
        cand = np.nonzero(score[r] <= kth[r])[0]
        order = np.lexsort((cand, score[r, cand]))[:k]
        idx_out[r] = cand[order]
        val_out[r] = score[r, cand[order]]
    return idx_out, val_out


def knn_feature(features, k, metric="euclidean", positions=None, lambda_h=1.0,
                lambda_g=0.0, rows=None, exclude_self=True, chunk=2048):
    """k nearest neighbours per row, self excluded BY INDEX.

    metric 'euclidean': ranks by sklearn's float64 d2 (preprocess_hypergraph.py
    :379-388; the reference asks for k+1 and drops column 0, which equals
    index-exclusion whenever no other row is at distance 0 -- see DESIGN.md).
    metric 'gaussian': ranks by lambda_h*d2_h (+ lambda_g*d2_g when positions
    are given) == descending K of similarity_kernel.py:52/:84/:122; value
    returned is K = exp(-score) in float64.
    metric 'cosine': ranks by -cos(f_i, f_j) (F.cosine_similarity semantics,
    preprocess_hypergraph.py:419; sklearn metric='cosine' at :311).
    Returns (idx [R,k] int64, val [R,k] float64): val = d2 / K / cos.
    """
    x = np.asarray(features, dtype=np.float64)
    n = x.shape[0]
    rows = np.arange(n) if rows is None else np.asarray(rows)
    idx_out = np.empty((rows.size, k), dtype=np.int64)
    val_out = np.empty((rows.size, k), dtype=np.float64)
    if metric == "cosine":
        xn = x / np.maximum(np.linalg.norm(x, axis=1, keepdims=True), 1e-8)
    for s in range(0, rows.size, chunk):
        r = rows[s:s + chunk]
        if metric == "cosine":
            score = -(xn[r] @ xn.T)
        else:
            score = sq_dists_f64(x[r], x)
            if metric == "gaussian":
                score *= lambda_h
                if positions is not None and lambda_g != 0.0:
                    p = np.asarray(positions, dtype=np.float64)
                    score += lambda_g * sq_dists_f64(p[r], p)
        if exclude_self:
            score[np.arange(r.size), r] = np.inf
        idx, val = _topk_rows(score, k)
        idx_out[s:s + r.size] = idx
        if metric == "gaussian":
            val = np.exp(-val)
        elif metric == "cosine":
            val = -val
        val_out[s:s + r.size] = val
    return idx_out, val_out


def knn_spatial(positions, k, rows=None, chunk=4096):
    """A3: k smallest |g_i-g_j|^2 excluding self, ties by (d2, index).
    Semantics pinned by compute_spatial_similarity (similarity_kernel.py:57-86):
    top-k of a K_g row == k smallest squared distances.  Direct-difference form
    in float64 (exact for integer grid coordinates)."""
    p = np.asarray(positions, dtype=np.float64)
    n = p.shape[0]
    rows = np.arange(n) if rows is None else np.asarray(rows)
    idx_out = np.empty((rows.size, k), dtype=np.int64)
    val_out = np.empty((rows.size, k), dtype=np.float64)
    for s in range(0, rows.size, chunk):
        r = rows[s:s + chunk]
        diff = p[r][:, None, :] - p[None, :, :]
        d2 = (diff * diff).sum(-1)
        d2[np.arange(r.size), r] = np.inf
        idx, val = _topk_rows(d2, k)
        idx_out[s:s + r.size] = idx
        val_out[s:s + r.size] = val
    return idx_out, val_out


# --------------------------------------------------------------------------
# A5: edge assembly (kNN pairs  U  intra-cluster cliques, undirected dedup)
# --------------------------------------------------------------------------
def edges_from_knn(knn_idx, labels=None, num_hyperedges=0):
    """preprocess_hypergraph.py:385-410.

    Directed kNN pairs (i, nbr) (:386-388) plus, for every cluster id in
    [0, num_hyperedges), all ordered pairs i != j of its members (:395-400);
    each pair is canonicalised to (min, max) and set-deduplicated (:403).  The
    reference order is CPython set-iteration order; the canonical order here --
    and in the CUDA path -- is lexicographic by (min, max).
    Returns edge_index int64 [2, E]."""
    knn_idx = np.asarray(knn_idx, dtype=np.int64)
    n, k = knn_idx.shape
    src = np.repeat(np.arange(n, dtype=np.int64), k)
    dst = knn_idx.ravel()
    keep = dst >= 0
    src, dst = src[keep], dst[keep]
    lo, hi = [np.minimum(src, dst)], [np.maximum(src, dst)]
    if labels is not None and num_hyperedges > 0:
        labels = np.asarray(labels)
        for c in range(num_hyperedges):
            members = np.nonzero(labels == c)[0].astype(np.int64)
            if members.size > 1:
                a, b = np.triu_indices(members.size, 1)
                lo.append(members[a])
                hi.append(members[b])
    lo = np.concatenate(lo)
    hi = np.concatenate(hi)
    if lo.size == 0:
        return np.empty((2, 0), dtype=np.int64)
    width = np.int64(max(int(hi.max()) + 1, n))
    key = np.unique(lo * width + hi)
    return np.stack([key // width, key % width]).astype(np.int64)


# --------------------------------------------------------------------------
# A6: cosine edge weights
# --------------------------------------------------------------------------
def edge_weights_cosine(features, edge_index, dtype=np.float64, eps=1e-8):
    """preprocess_hypergraph.py:413-422: w = max(0, cos(f_i, f_j)) with
    F.cosine_similarity's eps=1e-8 (each norm clamped from below)."""
    x = np.asarray(features, dtype=dtype)
    i, j = np.asarray(edge_index)
    a, b = x[i], x[j]
    na = np.maximum(np.sqrt((a * a).sum(1)), eps)
    nb = np.maximum(np.sqrt((b * b).sum(1)), eps)
    w = (a * b).sum(1) / (na * nb)
    return np.maximum(w, 0.0)


# --------------------------------------------------------------------------
# A8: dense-threshold builder
# --------------------------------------------------------------------------
def threshold_edges(K: np.ndarray, threshold_median_ratio):
    """similarity_kernel.py:183-210: lower median of the off-diagonal of K times
    the ratio; keep every ordered (i, j) -- diagonal included, both directions --
    with K[i,j] >= threshold, row-major order.  ratio=None keeps everything
    (the documented intent at :147-152; the reference line :188 raises TypeError
    for None)."""
    K = np.asarray(K)
    n = K.shape[0]
    if n <= 1:
        raise ValueError("Number of nodes must be greater than 1")  # :176-178
    if threshold_median_ratio is None:
        keep = np.ones_like(K, dtype=bool)
        thr = None
    else:
        off = K[~np.eye(n, dtype=bool)]
        thr = K.dtype.type(lower_median(off)) * K.dtype.type(threshold_median_ratio)
        keep = ~(K < thr)
    i, j = np.nonzero(keep)
    return np.stack([i, j]).astype(np.int64), K[i, j], thr


# --------------------------------------------------------------------------
# CSR incidence (new output; validated by round-trip to COO)
# --------------------------------------------------------------------------
def csr_from_coo(edge_index, num_nodes):
    """Node -> (neighbour, edge id) adjacency of the undirected pair list.
    Each edge e=(u,v) appears in row u (col v) and row v (col u); rows are
    sorted by neighbour id.  rowptr int64 [N+1], col int64 [2E], eid int64 [2E].
    Consumer semantics: downstream_survival/models/cust_omics.py:101 reads
    edge_index as (node, hyperedge) incidence; see DESIGN.md."""
    u, v = np.asarray(edge_index, dtype=np.int64)
    e = np.arange(u.size, dtype=np.int64)
    node = np.concatenate([u, v])
    nbr = np.concatenate([v, u])
    eid = np.concatenate([e, e])
    order = np.lexsort((eid, nbr, node))
    node, nbr, eid = node[order], nbr[order], eid[order]
    rowptr = np.zeros(num_nodes + 1, dtype=np.int64)
    np.add.at(rowptr, node + 1, 1)
    return np.cumsum(rowptr), nbr, eid


# --------------------------------------------------------------------------
# A4+A5+A6 composed: the reference's build_hypergraph_knn_kmeans, vectorised
# --------------------------------------------------------------------------
def build_hypergraph_knn(wsi_features, tma_features, k=5, labels=None, num_hyperedges=0):
    """preprocess_hypergraph.py:374-433 with the KMeans labels as an INPUT
    (KMeans is not reproducible across implementations -- SURVEY.md section 7).
    Returns (edge_index [2,E] int64 sorted, edge_weights [E] float64)."""
    allf = np.concatenate([np.asarray(wsi_features), np.asarray(tma_features)], 0)
    idx, _ = knn_feature(allf, k, metric="euclidean")
    ei = edges_from_knn(idx, labels, num_hyperedges)
    return ei, edge_weights_cosine(allf, ei)


# --------------------------------------------------------------------------
# (f)3: what the consumer does with edge_index (parity unpinned: torch_geometric is absent here)
# --------------------------------------------------------------------------
def hypergraph_conv_propagate(x, edge_index, num_nodes=None):
    """Sparse part of torch_geometric.nn.HypergraphConv (2.x, use_attention=False, unit hyperedge
    weights) as the reference's consumer calls it -- ``conv(x, edge_index)`` at
    downstream_survival/models/cust_omics.py:58-63, 100-102 -- restated from its published
    formula  out = D^-1 H B^-1 H^T x :  row 0 of edge_index = node ids, row 1 = hyperedge ids;
    D_v = number of pairs with node v, B_j = number of pairs with hyperedge j, 1/0 -> 0;
    m_j = B_j^-1 sum_{(i, j)} x_i;  out_v = D_v^-1 sum_{(v, j)} m_j.  float64.
    Returns (out [N, F], m [num_hyperedges, F])."""
    x = np.asarray(x, dtype=np.float64)
    ei = np.asarray(edge_index, dtype=np.int64)
    n = x.shape[0] if num_nodes is None else num_nodes
    nodes, hedges = ei[0], ei[1]
    num_e = int(hedges.max()) + 1 if hedges.size else 0
    B = np.bincount(hedges, minlength=num_e).astype(np.float64)
    D = np.bincount(nodes, minlength=n).astype(np.float64)
    Binv = np.divide(1.0, B, out=np.zeros_like(B), where=B > 0)
    Dinv = np.divide(1.0, D, out=np.zeros_like(D), where=D > 0)
    m = np.zeros((num_e, x.shape[1]))
    np.add.at(m, hedges, x[nodes])
    m *= Binv[:, None]
    out = np.zeros((n, x.shape[1]))
    np.add.at(out, nodes, m[hedges])
    out *= Dinv[:, None]
    return out, m


# --------------------------------------------------------------------------
# (f)4: super-patch aggregation by cluster label
# --------------------------------------------------------------------------
def aggregate_super_patches(features, positions, labels, num_super_patches, K=None,
                            lambda_h=1.0, lambda_g=1.0, dtype=np.float64):
    """Mean feature / position per cluster and the intra-cluster similarity statistic.

    Follows build_hypergraph/preprocess_hypergraph.py:157-170 (boolean-mask mean per cluster id,
    ValueError on an empty cluster) and :175-190 (per cluster with more than one member: mean of
    the off-diagonal entries of K[idx][:, idx]; the statistic is the mean of those means, 0.0
    when no cluster qualifies).  ``K`` defaults to combined_similarity(features, positions).
    Returns (super_features, super_positions, counts, intra_per_cluster (NaN when m <= 1), avg)."""
    x = np.asarray(features, dtype=dtype)
    p = np.asarray(positions, dtype=dtype)
    lab = np.asarray(labels)
    if K is None:
        K = combined_similarity(x, p, lambda_h, lambda_g, dtype=dtype)
    sup_f, sup_p, counts = [], [], []
    intra = np.full(num_super_patches, np.nan, dtype=np.float64)
    for c in range(num_super_patches):
        mask = lab == c
        m = int(mask.sum())
        if m == 0:
            raise ValueError(f"Cluster {c} is empty")
        counts.append(m)
        sup_f.append(x[mask].mean(axis=0))
        sup_p.append(p[mask].mean(axis=0))
        if m > 1:
            idx = np.nonzero(mask)[0]
            blk = np.asarray(K, dtype=np.float64)[np.ix_(idx, idx)]
            off = ~np.eye(m, dtype=bool)
            intra[c] = blk[off].mean()
    defined = intra[~np.isnan(intra)]
    avg = float(defined.mean()) if defined.size else 0.0
    return np.stack(sup_f), np.stack(sup_p), np.asarray(counts, dtype=np.int64), intra, avg


# --------------------------------------------------------------------------
# (f)1: KMeans(n_clusters, random_state, n_init=10) -- preprocess_hypergraph.py:150-151, 299-300, 391-392
# --------------------------------------------------------------------------
_M64 = (1 << 64) - 1


def _km_mix(z: int) -> int:                       # splitmix64 finaliser (the device RNG of hg_kmeans_fit)
    z = (z + 0x9E3779B97F4A7C15) & _M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return z ^ (z >> 31)


def _km_rand(seed: int, run: int, step: int, t: int) -> int:
    return _km_mix(_km_mix((seed ^ (run << 40)) & _M64) ^ ((step << 8) & _M64) ^ t)


def kmeans_fit(features, n_clusters, n_init=10, max_iter=300, tol=1e-4, seed=42):
    """The estimator the reference constructs, as scikit-learn (1.3 / 1.9) defines it -- the library is
    not under /root/reference, so this restates its published algorithm (sklearn/cluster/_kmeans.py):
    X centred on its column means; tolerance = tol * mean(var(X, axis=0)); per restart greedy k-means++
    (`_kmeans_plusplus`: 2 + int(ln k) local trials, D^2 sampling by searchsorted on the cumulative sum,
    the candidate with the smallest potential wins) then Lloyd until no label moves or the squared centre
    shift is <= tolerance (`_kmeans_single_lloyd`), labels of the final centres, best restart by inertia.
    scikit-learn draws from numpy's MT19937; the CUDA path cannot share that stream, so the random numbers
    here are the counter-based ones of hg_kmeans_fit (kmeans.cu) -- which makes this restatement comparable
    label for label with the device result.  Empty clusters keep their centre here (scikit-learn relocates them, and so
    does hg_kmeans_fit -- km_relocate_kernel; it does not happen on the data this restatement is compared on).  Returns (labels int64 [n], centres float64 [k, d], inertia)."""
    x32 = np.ascontiguousarray(features, dtype=np.float32)
    n, d = x32.shape
    if n < n_clusters:
        raise ValueError(f"n_samples={n} should be >= n_clusters={n_clusters}.")
    mean = x32.astype(np.float64).mean(0).astype(np.float32)
    x = (x32 - mean).astype(np.float64)            # the device centres in FP32
    tol_abs = tol * x32.astype(np.float64).var(0).mean()
    trials = min(16, 2 + int(np.log(n_clusters)))
    sq = (x * x).sum(1)

    def d2_to(rows):                                # [len(rows), n]
        c = x[rows]
        return np.maximum(sq[None, :] + (c * c).sum(1)[:, None] - 2.0 * (c @ x.T), 0.0)

    best = None
    for run in range(n_init):
        idx = [_km_rand(seed, run, 0, 0) % n]
        closest = d2_to(idx)[0]
        for step in range(1, n_clusters):
            cum = np.cumsum(closest)
            us = np.array([(_km_rand(seed, run, step, t) >> 11) * (1.0 / 9007199254740992.0) for t in range(trials)])
            cand = np.minimum(np.searchsorted(cum, us * cum[-1], side="left"), n - 1)
            pot = np.minimum(closest[None, :], d2_to(cand))
            b = int(np.argmin(pot.sum(1)))
            closest = pot[b]
            idx.append(int(cand[b]))
        centers = x[idx].copy()
        prev = np.full(n, -1)
        for _ in range(max_iter):
            full = sq[:, None] + (centers * centers).sum(1)[None, :] - 2.0 * (x @ centers.T)
            labels = full.argmin(1)
            moved = int((labels != prev).sum())
            new = centers.copy()
            for c in range(n_clusters):
                m = labels == c
                if m.any():
                    new[c] = x[m].mean(0).astype(np.float32)       # the device stores FP32 centres
            shift = ((new - centers) ** 2).sum()
            centers = new
            prev = labels
            if shift <= tol_abs or moved == 0:
                break
        full = np.maximum(sq[:, None] + (centers * centers).sum(1)[None, :] - 2.0 * (x @ centers.T), 0.0)
        labels = full.argmin(1)
        inertia = float(full[np.arange(n), labels].sum())
        if best is None or inertia < best[2]:
            best = (labels.astype(np.int64), centers + mean.astype(np.float64), inertia)
    return best
