"""GPU parity tests: the CUDA path (through the C ABI) against the oracle and the golden
fixtures generated from the reference.  Bit-exact for indices / edges / CSR; 1e-5 relative
for similarity values (north_star tolerance); neighbour sets exact except reported near-ties
(oracle gap between the swapped candidates below 1e-5 relative)."""
import numpy as np
import pytest
import torch

import oracle_np as oracle
from fp64_check import assert_knn_fp64, check_knn_fp64
from hg_b200 import _lib, ops, synth
import hg_b200 as hg

pytestmark = pytest.mark.gpu
RTOL = 1e-5          # north_star: similarity values within 1e-5 relative


def _cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _oracle_score(feats, pos, metric, lambda_h, lambda_g, i, js):
    """float64 ranking score (ascending = better) of row i against columns js."""
    x = feats.astype(np.float64)
    if metric == "cosine":
        xn = x / np.maximum(np.linalg.norm(x, axis=1, keepdims=True), 1e-8)
        return -(xn[js] @ xn[i])
    d2 = ((x[js] - x[i]) ** 2).sum(1)
    if metric == "gaussian":
        d2 = lambda_h * d2
        if pos is not None and lambda_g:
            p = pos.astype(np.float64)
            d2 = d2 + lambda_g * ((p[js] - p[i]) ** 2).sum(1)
    return d2


def check_knn(idx_gpu, feats, k, metric, rows=None, pos=None, lambda_h=1.0, lambda_g=0.0, max_near_ties=None):
    """Sets equal per row, except rows whose differing members are near-ties of the oracle's
    k-th score (|gap| <= RTOL * scale).  Returns the number of near-tie rows (reported)."""
    rows = np.arange(feats.shape[0]) if rows is None else np.asarray(rows)
    idx_or, _ = oracle.knn_feature(feats, k, metric, positions=pos, lambda_h=lambda_h, lambda_g=lambda_g, rows=rows)
    idx_gpu = np.asarray(idx_gpu, dtype=np.int64)
    assert idx_gpu.shape == idx_or.shape
    assert (idx_gpu >= 0).all() and (idx_gpu != rows[:, None]).all()
    bad = np.nonzero((np.sort(idx_gpu, 1) != np.sort(idx_or, 1)).any(1))[0]
    for r in bad:
        i = rows[r]
        only_g = np.setdiff1d(idx_gpu[r], idx_or[r])
        only_o = np.setdiff1d(idx_or[r], idx_gpu[r])
        sg = _oracle_score(feats, pos, metric, lambda_h, lambda_g, i, only_g)
        so = _oracle_score(feats, pos, metric, lambda_h, lambda_g, i, only_o)
        scale = max(1.0, float(np.abs(so).max()))
        assert np.abs(sg.max() - so.min()) <= RTOL * scale, \
            f"row {i}: GPU picked {only_g} (scores {sg}) instead of {only_o} (scores {so})"
    limit = max(2, len(rows) // 500) if max_near_ties is None else max_near_ties
    assert len(bad) <= limit, f"{len(bad)} near-tie rows out of {len(rows)}"
    return len(bad)


# ------------------------------------------------------------------------------- operands
def test_split_operands():
    lib = _lib.load_library()
    n, d = 300, 100
    x = synth.make_features(n, d, 3, "gaussian")
    xc = _cuda(x)
    nbytes = lib.hg_operand_bytes(n, d)
    n_pad, d_pad = 512, 128
    assert nbytes == 2 * n_pad * d_pad * 2
    ops_buf = torch.empty(nbytes // 2, dtype=torch.float16, device="cuda")
    norms = torch.empty(n, dtype=torch.float32, device="cuda")
    _lib.check(lib.hg_split_operands(xc.data_ptr(), n, d, ops_buf.data_ptr(), norms.data_ptr(), 0))
    torch.cuda.synchronize()
    blocks = ops_buf.view(n_pad, d_pad // 32, 2, 32).float().cpu().numpy()      # [row][k-block][hi | lo][32]
    rec = (blocks[:, :, 0, :] + blocks[:, :, 1, :]).reshape(n_pad, d_pad)
    # every row was scaled by the power of two that puts its norm into [2^5, 2^6) (knn_tc.cuh op_scale)
    nrm = norms.cpu().numpy()
    e = (nrm.view(np.int32) >> 23) & 0xff
    scale = np.ldexp(np.float32(1.0), 6 - ((e - 125) >> 1)).astype(np.float32)
    scaled_norm = np.sqrt((rec[:n].astype(np.float64) ** 2).sum(1))
    assert (scaled_norm >= 2.0 ** 5 * (1 - 1e-6)).all() and (scaled_norm < 2.0 ** 6).all()
    # hi + lo carries 22 significant bits (2^-25 absolute on the subnormal grid)
    err = np.abs(rec[:n, :d].astype(np.float64) - x.astype(np.float64) * scale[:, None])
    assert (err <= 2.0 ** -22 * np.abs(x * scale[:, None]) + 2.0 ** -25).all()
    assert (np.abs(blocks[:n, :, 1, :]) <= 2.0 ** -11 * np.abs(blocks[:n, :, 0, :]) + 2.0 ** -25).all()   # lo is the rounding residue of hi
    assert (rec[n:] == 0).all() and (rec[:, d:] == 0).all()
    np.testing.assert_allclose(norms.cpu().numpy(), (x.astype(np.float64) ** 2).sum(1), rtol=1e-6)


# ------------------------------------------------------------------------ dense kernels (A1/A2/A7)
def test_dense_kernels_vs_reference_golden(golden):
    g = golden("dense_n192_d64")
    feats, pos = synth.make_slide(int(g["n"]), int(g["d"]), int(g["seed"]), "gaussian")
    lh, lg = float(g["lambda_h"]), float(g["lambda_g"])
    f, p = _cuda(feats), _cuda(pos)
    K_h = hg.compute_morphological_similarity(f, lh).cpu().numpy()
    K_g = hg.compute_spatial_similarity(p, lg).cpu().numpy()
    K = hg.compute_combined_similarity(f, p, lh, lg).cpu().numpy()
    np.testing.assert_allclose(K_h, g["K_h"], rtol=RTOL)
    np.testing.assert_allclose(K_g, g["K_g"], rtol=RTOL)
    np.testing.assert_allclose(K, g["K"], rtol=RTOL)
    np.testing.assert_allclose(K, oracle.combined_similarity(feats, pos, lh, lg), rtol=RTOL)
    tma = synth.make_features(int(g["n_tma"]), int(g["d"]), int(g["seed"]) + 1, "gaussian")
    sim, st = hg.compute_wsi_tma_similarity(f[:100], p[:100], _cuda(tma), lh, lg)
    np.testing.assert_allclose(sim.cpu().numpy(), g["wsi_tma"], rtol=RTOL)
    got = np.array([st[s] for s in ("mean", "std", "min", "max", "median")])
    np.testing.assert_allclose(got, g["wsi_tma_stats"], rtol=RTOL)
    pooled = hg.mean_pool_with_similarity(f).cpu().numpy()
    np.testing.assert_allclose(pooled, g["pooled"], rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("m,n,d", [(1, 1, 1), (65, 130, 33), (257, 64, 1024), (100, 80, 7)])
def test_dense_ragged_shapes(m, n, d):
    rng = np.random.default_rng(m * 1000 + n)
    a = (rng.standard_normal((m, d)) * 0.3).astype(np.float32)
    b = (rng.standard_normal((n, d)) * 0.3).astype(np.float32)
    for form in (_lib.FORM_EXPANSION, _lib.FORM_DIRECT):
        out = ops.similarity_dense(_cuda(a), _cuda(b), lambda_h=0.7, form=form).cpu().numpy()
        ref = np.exp(-0.7 * oracle.sq_dists_f64(a, b))
        np.testing.assert_allclose(out, ref, rtol=RTOL, atol=1e-30)


def test_stats_and_offdiag_median():
    rng = np.random.default_rng(5)
    for n in (2, 37, 300):
        K = rng.random((n, n)).astype(np.float32) - 0.3
        st = ops.similarity_stats(_cuda(K))
        ref = oracle.similarity_stats(K)
        for key in ref:
            assert np.isclose(st[key], ref[key], rtol=1e-6, atol=1e-7), (n, key, st[key], ref[key])
        med = ops.similarity_stats(_cuda(K), exclude_diagonal=True)["median"]
        assert med == oracle.lower_median(K[~np.eye(n, dtype=bool)])


@pytest.mark.parametrize("tag,ratio", [("thr", 1.0), ("thr2", 0.5)])
def test_threshold_builder_vs_reference_golden(golden, tag, ratio):
    g = golden("dense_n192_d64")
    feats, pos = synth.make_slide(int(g["n"]), int(g["d"]), int(g["seed"]), "gaussian")
    nt = int(g["n_thr"])
    ei, w = hg.build_weighted_hypergraph(_cuda(feats[:nt]), _cuda(pos[:nt]), float(g["lambda_h"]),
                                         float(g["lambda_g"]), threshold_median_ratio=ratio)
    ei, w = ei.cpu().numpy(), w.cpu().numpy()
    ref_ei, ref_w = g[tag + "_edge_index"], g[tag + "_edge_weights"]
    a, b = set(map(tuple, ei.T.tolist())), set(map(tuple, ref_ei.T.tolist()))
    # entries sitting on the threshold may flip with FP32 summation order: each flipped pair must be within 1e-5
    # relative of the threshold (the reference's own median * ratio), and there may be at most two of them
    K_ref = oracle.combined_similarity(feats[:nt], pos[:nt], float(g["lambda_h"]), float(g["lambda_g"]))
    thr = oracle.lower_median(K_ref[~np.eye(nt, dtype=bool)].astype(np.float32)) * ratio
    flipped = sorted(a ^ b)
    for (i, j) in flipped:
        dist = abs(K_ref[i, j] - thr) / thr
        print(f"threshold builder ({tag}): pair {(i, j)} flipped, K = {K_ref[i, j]:.9g}, threshold {thr:.9g}, {dist:.2e} relative")
        assert dist <= RTOL, f"pair {(i, j)} flipped {dist:.2e} away from the threshold"
    assert len(flipped) <= 2
    # order (row-major, diagonal and both directions kept) and weights on the common pairs, always
    keep = np.array([p in b for p in map(tuple, ei.T.tolist())])
    keep_ref = np.array([p in a for p in map(tuple, ref_ei.T.tolist())])
    assert (ei[:, keep] == ref_ei[:, keep_ref]).all()
    np.testing.assert_allclose(w[keep], ref_w[keep_ref], rtol=RTOL)
    if not flipped:
        assert (ei == ref_ei).all()
    # ratio None keeps everything (documented intent of the reference)
    ei_all, _ = hg.build_weighted_hypergraph(_cuda(feats[:nt]), _cuda(pos[:nt]), 1.0, 1.0)
    assert ei_all.shape[1] == nt * nt
    with pytest.raises(ValueError):
        hg.build_weighted_hypergraph(_cuda(feats[:1]), _cuda(pos[:1]), 1.0, 1.0, 1.0)


# --------------------------------------------------------------------------------- kNN (A1/A2/A4)
KNN_SHAPES = [
    # n, d, k, metric, seed
    (512, 128, 8, "cosine", 21),
    (1000, 100, 5, "euclidean", 31),
    (300, 70, 16, "gaussian", 32),
    (2500, 256, 16, "gaussian", 33),
    (4096, 1024, 8, "cosine", 1235),      # BASELINE config C1
    (129, 64, 32, "euclidean", 34),
    (257, 192, 1, "cosine", 35),
]


@pytest.mark.parametrize("precision", ["tc", "fp32"])
@pytest.mark.parametrize("n,d,k,metric,seed", KNN_SHAPES)
def test_knn_feature_vs_oracle(n, d, k, metric, seed, precision):
    mode = "cosine" if metric == "cosine" else "gaussian"
    feats = synth.make_features(n, d, seed, mode)
    idx, val = ops.knn_feature(_cuda(feats), k, metric, precision=precision)
    idx, val = idx.cpu().numpy(), val.cpu().numpy()
    check_knn(idx, feats, k, metric)
    # values: 1e-5 relative against float64 for the returned neighbours, best first
    rows = np.arange(n)
    x = feats.astype(np.float64)
    for r in rows[:: max(1, n // 64)]:
        s = _oracle_score(feats, None, metric, 1.0, 0.0, r, idx[r].astype(np.int64))
        ref = {"euclidean": s, "gaussian": np.exp(-s), "cosine": -s}[metric]
        # no absolute term for similarities; squared distances of (near-)duplicates may be ~0
        np.testing.assert_allclose(val[r], ref, rtol=RTOL, atol=1e-6 if metric == "euclidean" else 0.0)
        assert (np.diff(s) >= -RTOL * max(1.0, np.abs(s).max())).all()


@pytest.mark.parametrize("k,metric,groups", [(16, "euclidean", 50), (25, "euclidean", 50), (32, "euclidean", 50),
                                              (16, "euclidean", 6), (32, "euclidean", 6), (32, "cosine", 6), (16, "gaussian", 6)])
def test_knn_near_duplicates_stay_exact(k, metric, groups, monkeypatch, capfd):
    """Near-duplicate rows (WSI background patches): far more columns sit inside the 3xBF16 slack of the k-th
    best than the k + 8 a single selection round used to take.  With 50 groups (60 copies each, spread over the
    column splits) the candidate lists still hold every copy and the re-rank evaluates them exactly, 32 per
    round; with 6 groups (500 copies each) the lists themselves overflow with copies, cannot vouch for the exact
    top-k, and the rows go to the exact fallback against all columns (rerank.cu, exact_slices_kernel).
    Result = the float64 direct-difference ranking, gaps judged RELATIVE to the distance itself."""
    n, d = 3000, 256
    rng = np.random.default_rng(71)
    base = synth.make_features(groups, d, 71, "gaussian")
    feats = (base[rng.integers(0, groups, n)] + 1e-4 * rng.standard_normal((n, d))).astype(np.float32)
    monkeypatch.setenv("HG_B200_TC_DEBUG", "1")
    idx, val = ops.knn_feature(_cuda(feats), k, metric)
    torch.cuda.synchronize()
    err = capfd.readouterr().err
    flagged = int(err.split("hg_b200 rerank:")[1].split()[0])
    if groups == 6:
        assert flagged > n // 2, f"expected most rows to be uncertified, got {flagged}"
    idx, val = idx.cpu().numpy().astype(np.int64), val.cpu().numpy()
    x = feats.astype(np.float64)
    if metric == "cosine":
        x = x / np.linalg.norm(x, axis=1, keepdims=True)
    near = 0
    for r in range(0, n, 7):
        d2 = ((x - x[r]) ** 2).sum(1)                         # cosine on unit rows: d2 = 2 - 2 cos, same order
        d2[r] = np.inf
        ref = np.argsort(d2, kind="stable")[:k]
        if set(ref.tolist()) != set(idx[r].tolist()):
            only_g = np.setdiff1d(idx[r], ref)
            only_o = np.setdiff1d(ref, idx[r])
            # the tolerance is relative to the similarity the metric returns: d2 itself (euclidean), cos = 1 - d2 / 2
            # on unit rows, K = exp(-d2) (relative gap = difference of the exponents)
            gap = abs(d2[only_g].max() - d2[only_o].min()) / {"euclidean": d2[only_o].min(), "cosine": 2.0, "gaussian": 1.0}[metric]
            assert gap <= 1e-5, f"row {r}: picked {only_g} over {only_o}, relative gap {gap:.2e}"
            near += 1
        got = d2[idx[r]]
        refv = {"euclidean": got, "gaussian": np.exp(-got), "cosine": 1.0 - got / 2}[metric]
        np.testing.assert_allclose(val[r], refv, rtol=RTOL if metric != "euclidean" else 1e-4, atol=0)
    # FP32 cosines (and exp(-d2)) cannot tell near-duplicates 1e-8 apart: only squared distances resolve them
    assert near <= 3 or metric != "euclidean"


@pytest.mark.parametrize("sym", ["0", "8"])
def test_knn_fallback_rescans_only_the_flagged_splits(sym, monkeypatch, capfd):
    """40 groups of 500 near-duplicates, each group CONTIGUOUS in the row order: every row's in-slack candidates fall
    into one column split, whose list overflows with copies; the row is flagged with that split's bit only and the
    exact fallback rescans that split's column range, merging with what the re-rank already evaluated from the other
    splits (rerank.cu, exact_slices_kernel / exact_merge_kernel).  Classic pass (HG_B200_TC_SYM=0) and symmetric pass;
    result = the float64 direct-difference ranking up to near-ties relative to the distance itself."""
    n, d, k, groups = 20_000, 128, 16, 40
    rng = np.random.default_rng(72)
    base = synth.make_features(groups, d, 72, "gaussian")
    feats = (np.repeat(base, n // groups, axis=0) + 1e-4 * rng.standard_normal((n, d))).astype(np.float32)
    monkeypatch.setenv("HG_B200_TC_SYM", sym)
    monkeypatch.setenv("HG_B200_TC_DEBUG", "1")
    idx, val = ops.knn_feature(_cuda(feats), k, "euclidean")
    torch.cuda.synchronize()
    err = capfd.readouterr().err
    flagged = int(err.split("hg_b200 rerank:")[1].split()[0])
    assert flagged > n // 2, f"expected most rows to be uncertified, got {flagged}"
    idx, val = idx.cpu().numpy().astype(np.int64), val.cpu().numpy()
    x = feats.astype(np.float64)
    for r in range(0, n, 97):
        g0 = (r // 500) * 500
        d2 = ((x[g0:g0 + 500] - x[r]) ** 2).sum(1)            # every neighbour lies in the row's own group
        d2[r - g0] = np.inf
        ref = g0 + np.argsort(d2, kind="stable")[:k]
        assert ((idx[r] >= g0) & (idx[r] < g0 + 500)).all()
        if set(ref.tolist()) != set(idx[r].tolist()):
            only_g, only_o = np.setdiff1d(idx[r], ref), np.setdiff1d(ref, idx[r])
            gap = abs(d2[only_g - g0].max() - d2[only_o - g0].min()) / d2[only_o - g0].min()
            assert gap <= 1e-5, f"row {r}: picked {only_g} over {only_o}, relative gap {gap:.2e}"
        np.testing.assert_allclose(val[r], d2[idx[r] - g0], rtol=1e-4, atol=0)


@pytest.mark.parametrize("sym", ["0", "8"])
def test_knn_cosine_is_blind_to_row_scales(sym, monkeypatch):
    """The operands are FP16 halves of rows scaled by a per-row power of two (knn_tc.cuh op_scale): rows whose norms
    span 2^-20 .. 2^40 must give the cosine lists of the unscaled rows -- classic pass (three products) and
    symmetric pass (two products, forced at this size).  (Norms stay above the 1e-8 clamp of F.cosine_similarity,
    preprocess_hypergraph.py:419, below which the reference's cosine stops being scale-free.)"""
    n, d, k = 6000, 200, 12
    feats = synth.make_features(n, d, 81, "cosine")
    rng = np.random.default_rng(81)
    scaled = (feats * np.exp2(rng.integers(-20, 41, size=(n, 1))).astype(np.float32)).astype(np.float32)
    monkeypatch.setenv("HG_B200_TC_SYM", sym)
    idx0, val0 = ops.knn_feature(_cuda(feats), k, "cosine")
    idx1, val1 = ops.knn_feature(_cuda(scaled), k, "cosine")
    check_knn(idx1.cpu().numpy(), feats, k, "cosine")
    same = (idx0 == idx1).all(1).float().mean().item()
    assert same > 0.995, same                                      # the rest: near-ties that FP32 norms of scaled rows resolve differently
    np.testing.assert_allclose(val1.cpu().numpy(), val0.cpu().numpy(), rtol=0, atol=2e-6)


@pytest.mark.parametrize("sym", ["0", "8"])
def test_knn_wide_dynamic_range_inside_rows_and_tiny_rows(sym, monkeypatch):
    """Elements spanning twelve orders of magnitude inside a row (most of them land on FP16's subnormal grid after the
    row scale, or vanish), rows with tiny norms next to large ones, and an all-zero row: the exact re-rank decides, so
    the lists equal the float64 oracle's."""
    n, d, k = 5000, 96, 8
    rng = np.random.default_rng(82)
    feats = synth.make_features(n, d, 82, "gaussian")
    feats *= np.exp(rng.uniform(-14, 14, size=(1, d))).astype(np.float32)          # per-column scales e^-14 .. e^14
    feats[::7] *= 1e-6                                                             # tiny rows
    feats[11] = 0.0                                                                # a zero row
    feats = feats.astype(np.float32)
    monkeypatch.setenv("HG_B200_TC_SYM", sym)
    idx, val = ops.knn_feature(_cuda(feats), k, "euclidean")
    idx, val = idx.cpu().numpy(), val.cpu().numpy()
    check_knn(idx, feats, k, "euclidean", max_near_ties=25)
    x = feats.astype(np.float64)
    for r in (0, 7, 11, 12, 4999):
        d2 = ((x[idx[r].astype(np.int64)] - x[r]) ** 2).sum(1)
        np.testing.assert_allclose(val[r], d2, rtol=RTOL, atol=1e-30)


def test_knn_gaussian_with_positions():
    n, d, k = 3000, 128, 16
    feats, pos = synth.make_slide(n, d, 41, "gaussian")
    lam_g = 1.0 / 128.0
    idx, val = ops.knn_feature(_cuda(feats), k, "gaussian", positions=_cuda(pos), lambda_h=1.0, lambda_g=lam_g)
    idx = idx.cpu().numpy()
    check_knn(idx, feats, k, "gaussian", pos=pos, lambda_h=1.0, lambda_g=lam_g)
    K = oracle.combined_similarity(feats[:256], pos[:256], 1.0, lam_g)      # value spot check on a sub-block
    r = 17
    inside = idx[r] < 256
    np.testing.assert_allclose(val[r].cpu().numpy()[inside], K[r, idx[r][inside]], rtol=RTOL)


def test_knn_row_range_matches_full():
    n, d, k = 2000, 128, 8
    feats = synth.make_features(n, d, 51, "gaussian")
    f = _cuda(feats)
    full_i, full_v = ops.knn_feature(f, k, "euclidean")
    for r0, r1 in [(0, 700), (700, 1301), (1301, 2000), (5, 6)]:
        pi, pv = ops.knn_feature(f, k, "euclidean", rows=(r0, r1))
        assert torch.equal(pi, full_i[r0:r1])
        assert torch.equal(pv, full_v[r0:r1])


def test_knn_errors():
    f = torch.randn(1, 16, device="cuda")
    with pytest.raises(ValueError):
        ops.knn_feature(f, 1)                                    # N <= 1, reference similarity_kernel.py:176
    f = torch.randn(64, 16, device="cuda")
    with pytest.raises(ValueError):
        ops.knn_feature(f, 40)                                   # k > 32 on the tensor-core path
    with pytest.raises(ValueError):
        ops.knn_feature(f, 4, rows=(10, 5))
    idx, _ = ops.knn_feature(f[:5], 8, precision="fp32")         # fewer candidates than k -> padded with -1
    assert (idx[:, 4:] == -1).all() and (idx[:, :4] >= 0).all()
    idx, _ = ops.knn_feature(f[:5], 8)
    assert (idx[:, 4:] == -1).all() and (idx[:, :4] >= 0).all()


# ------------------------------------------------------------------------------- spatial kNN (A3)
@pytest.mark.parametrize("n,k,pd", [(500, 8, 2), (5000, 8, 2), (3000, 16, 3), (40, 32, 2)])
def test_knn_spatial_exact(n, k, pd):
    pos = synth.make_positions(n, n + k)
    if pd == 3:
        pos = np.concatenate([pos, (np.arange(n) % 3)[:, None].astype(np.float32)], 1)
    idx, d2 = ops.knn_spatial(_cuda(pos), k)
    ref_idx, ref_d2 = oracle.knn_spatial(pos, k)
    assert (idx.cpu().numpy() == ref_idx).all()          # grid ties resolved by (d2, index): bit-exact
    assert (d2.cpu().numpy() == ref_d2.astype(np.float32)).all()
    pi, pd2 = ops.knn_spatial(_cuda(pos), k, rows=(n // 3, n // 2))
    assert torch.equal(pi, idx[n // 3: n // 2]) and torch.equal(pd2, d2[n // 3: n // 2])


# ------------------------------------------------------------------- edges / weights / CSR (A5/A6)
KNN_GOLDEN = ["knn_n512_d128_k8", "knn_n768_d256_k5_cliques", "knn_n1024_d1024_k8", "knn_c1_n4096_d1024_k8"]


@pytest.mark.parametrize("name", KNN_GOLDEN)
def test_build_hypergraph_knn_kmeans_vs_reference_golden(golden, name):
    g = golden(name)
    n_wsi, n = int(g["n_wsi"]), int(g["n_wsi"]) + int(g["n_tma"])
    feats = synth.make_features(n, int(g["d"]), int(g["seed"]), str(g["mode"]))
    ei, w, stats = hg.build_hypergraph_knn_kmeans(
        _cuda(feats[:n_wsi]), _cuda(feats[n_wsi:]), np.zeros(n_wsi, dtype=np.int64), k=int(g["k"]),
        num_hyperedges=int(g["num_hyperedges"]), hyperedge_labels=g["labels"], return_csr=True)
    assert ei.dtype == torch.int64 and w.dtype == torch.float32
    assert stats["num_edges"] == int(g["num_edges"]) and stats["num_nodes"] == n
    assert (ei.cpu().numpy() == g["edge_index"]).all()                  # bit-exact incidence
    np.testing.assert_allclose(w.cpu().numpy(), g["edge_weights"], rtol=RTOL, atol=1e-6)
    rowptr, col, eid = (t.cpu().numpy() for t in stats["csr"])
    r_ref, c_ref, e_ref = oracle.csr_from_coo(g["edge_index"], n)
    assert (rowptr == r_ref).all() and (col == c_ref).all() and (eid == e_ref).all()


def test_edges_ragged_inputs():
    rng = np.random.default_rng(9)
    n = 777
    a = rng.integers(-1, n, size=(n, 7)).astype(np.int32)       # -1 = missing neighbour, self loops possible
    b = rng.integers(0, n, size=(n, 3)).astype(np.int32)
    labels = rng.integers(-2, 9, size=n).astype(np.int32)
    ei = ops.build_edges([_cuda(a), _cuda(b)], n, _cuda(labels), 5).cpu().numpy()
    ref = oracle.edges_from_knn(np.concatenate([a, b], 1), labels, 5)
    ref = ref[:, ref[0] != ref[1]]
    assert ei.shape == ref.shape and (ei == ref).all()
    empty = ops.build_edges([_cuda(np.full((n, 2), -1, np.int32))], n)
    assert empty.shape == (2, 0)
    rowptr, col, eid = ops.csr_from_coo(empty, n)
    assert (rowptr == 0).all() and col.numel() == 0
    w = ops.edge_weights_cosine(torch.randn(n, 8, device="cuda"), empty)
    assert w.numel() == 0


def test_radix_sort_large_keys():
    """2^20 random pairs over a 200k-node id space: 36-bit keys, 5 radix passes."""
    rng = np.random.default_rng(11)
    n = 200_000
    a = rng.integers(0, n, size=(n, 5)).astype(np.int32)
    ei = ops.build_edges([_cuda(a)], n).cpu().numpy()
    ref = oracle.edges_from_knn(a)
    ref = ref[:, ref[0] != ref[1]]
    assert ei.shape == ref.shape and (ei == ref).all()
    rowptr, col, eid = (t.cpu().numpy() for t in ops.csr_from_coo(_cuda(ei), n))
    r_ref, c_ref, e_ref = oracle.csr_from_coo(ei, n)
    assert (rowptr == r_ref).all() and (col == c_ref).all() and (eid == e_ref).all()


# ------------------------------------------------------------------------------ full-size properties
def test_c2_full_size_all_rows():
    """BASELINE config C2 (N=50k x 1024, Gaussian k=16): EVERY row against float64 scores (GPU checker,
    tests/fp64_check.py), 256 rows against the numpy oracle (which pins the checker), plus
    size-independent properties of the whole result."""
    n, d, k = 50_000, 1024, 16
    feats = synth.make_features(n, d, 1236, "gaussian")
    x = _cuda(feats)
    out = hg.knn_hyperedges(x, k=k, kernel="gaussian", lambda_h=1.0, return_csr=True)
    rep = check_knn_fp64(x, out["knn_idx"], out["knn_val"], torch.arange(n, device="cuda"), "gaussian", lambda_h=1.0)
    assert_knn_fp64(rep, "C2 all 50,000 rows")
    idx = out["knn_idx"].cpu().numpy()
    rows = np.random.default_rng(0).choice(n, 256, replace=False)
    check_knn(idx[rows], feats, k, "gaussian", rows=rows)
    val = out["knn_val"].cpu().numpy()
    assert (np.diff(val, axis=1) <= 0).all()                     # best first
    assert (idx != np.arange(n)[:, None]).all() and (idx >= 0).all() and (idx < n).all()
    assert all(len(set(r)) == k for r in idx[:: 97])
    ei = out["edge_index"].cpu().numpy()
    assert (ei[0] < ei[1]).all()
    key = ei[0] * n + ei[1]
    assert (np.diff(key) > 0).all()                              # sorted and unique
    ref = oracle.edges_from_knn(idx)
    assert ei.shape == ref.shape and (ei == ref).all()           # assembly bit-exact given the lists
    rowptr = out["csr_rowptr"].cpu().numpy()
    assert rowptr[0] == 0 and rowptr[-1] == 2 * ei.shape[1] and (np.diff(rowptr) >= 0).all()
    deg = np.bincount(np.concatenate([ei[0], ei[1]]), minlength=n)
    assert (np.diff(rowptr) == deg).all()
    w = out["edge_weights"].cpu().numpy()
    sel = np.random.default_rng(1).choice(ei.shape[1], 2000, replace=False)
    np.testing.assert_allclose(w[sel], oracle.edge_weights_cosine(feats, ei[:, sel]), rtol=RTOL, atol=1e-6)


def test_c3_combined_feature_and_spatial_sampled_rows():
    """BASELINE config C3 (N=100k x 1536, feature k=16 + spatial-grid k=8): 12.8 % of the rows vs float64."""
    n, d, k, ks = 100_000, 1536, 16, 8
    feats, pos = synth.make_slide(n, d, 1237, "gaussian")
    lam_g = 1.0 / 128.0
    out = hg.knn_hyperedges(_cuda(feats), _cuda(pos), k=k, kernel="gaussian", lambda_h=1.0, lambda_g=lam_g,
                            spatial_k=ks, return_csr=True)
    idx = out["knn_idx"].cpu().numpy()
    # 12.8 % of the rows against float64 scores on the GPU; 128 of them also against the numpy oracle
    big = torch.from_numpy(np.sort(np.random.default_rng(4).choice(n, 12_800, replace=False))).cuda()
    rep = check_knn_fp64(_cuda(feats), out["knn_idx"][big], out["knn_val"][big], big, "gaussian", lambda_h=1.0,
                         lambda_g=lam_g, pos=_cuda(pos), chunk=1024)
    assert_knn_fp64(rep, "C3 12,800 of 100,000 rows")
    rows = np.random.default_rng(3).choice(n, 128, replace=False)
    check_knn(idx[rows], feats, k, "gaussian", rows=rows, pos=pos, lambda_h=1.0, lambda_g=lam_g)
    sidx = out["spatial_idx"].cpu().numpy()
    ref_sidx, ref_sd2 = oracle.knn_spatial(pos, ks, rows=rows)
    assert (sidx[rows] == ref_sidx).all()
    assert (out["spatial_d2"].cpu().numpy()[rows] == ref_sd2.astype(np.float32)).all()
    ei = out["edge_index"].cpu().numpy()
    ref = oracle.edges_from_knn(np.concatenate([idx, sidx], 1))
    assert ei.shape == ref.shape and (ei == ref).all()
    rowptr = out["csr_rowptr"].cpu().numpy()
    assert rowptr[-1] == 2 * ei.shape[1]


def test_c5_row_block_of_a_giant_slide_sampled_rows():
    """BASELINE config C5 (N=250k x 1536, row-partitioned over 8 GPUs): the row block rank 3 owns, run
    against all N columns (what the rank holds after the all-gather), sampled rows vs the float64 oracle."""
    from hg_b200 import multigpu
    n, d, k = 250_000, 1536, 16
    feats = synth.make_features(n, d, 1239, "gaussian")
    r0, r1 = multigpu.row_range(n, 3, 8)
    idx, val = ops.knn_feature(_cuda(feats), k, "gaussian", lambda_h=1.0, rows=(r0, r1))
    idx, val = idx.cpu().numpy(), val.cpu().numpy()
    assert idx.shape == (r1 - r0, k)
    assert (np.diff(val, axis=1) <= 0).all() and (idx >= 0).all() and (idx < n).all()
    assert (idx != np.arange(r0, r1)[:, None]).all()
    x = _cuda(feats)
    big = torch.from_numpy(np.sort(r0 + np.random.default_rng(6).choice(r1 - r0, 4096, replace=False))).cuda()
    rep = check_knn_fp64(x, torch.from_numpy(idx).cuda()[big - r0], torch.from_numpy(val).cuda()[big - r0], big,
                         "gaussian", lambda_h=1.0, chunk=512)
    assert_knn_fp64(rep, "C5 row block of rank 3, 4,096 of 31,250 rows")
    del x
    rows = r0 + np.random.default_rng(5).choice(r1 - r0, 64, replace=False)
    check_knn(idx[rows - r0], feats, k, "gaussian", rows=rows)
    _, ref_val = oracle.knn_feature(feats, k, "gaussian", lambda_h=1.0, rows=rows[:16])
    np.testing.assert_allclose(val[rows[:16] - r0], ref_val, rtol=RTOL)


def test_streaming_stats_match_dense(monkeypatch):
    """A9 without the N x N matrix: chunked tensor-core sweep == statistics of the dense kernel."""
    n, d = 3000, 128
    feats, pos = synth.make_slide(n, d, 61, "gaussian")
    lam_g = 1.0 / 256
    ref = oracle.similarity_stats(oracle.combined_similarity(feats, pos, 1.0, lam_g))
    for chunk_mb in ("1", "4096"):                       # 24 chunks of 128 rows, then a single chunk
        monkeypatch.setenv("HG_B200_STATS_CHUNK_MB", chunk_mb)
        got = ops.similarity_stats_streaming(_cuda(feats), _cuda(pos), 1.0, lam_g)
        for key in ref:
            assert np.isclose(got[key], ref[key], rtol=RTOL, atol=1e-12), (chunk_mb, key, got[key], ref[key])
    got = ops.similarity_stats_streaming(_cuda(feats), None, 0.7, 0.0)
    ref = oracle.similarity_stats(oracle.morphological_similarity(feats, 0.7))
    for key in ref:
        assert np.isclose(got[key], ref[key], rtol=RTOL, atol=1e-12), (key, got[key], ref[key])


def test_aggregate_large_slide_streams_stats(monkeypatch):
    from hg_b200 import preprocess_hypergraph as pp
    monkeypatch.setattr(pp, "MAX_DENSE_N", 256)          # force the no-N x N branch on a small slide
    n, d = 700, 64
    feats, pos = synth.make_slide(n, d, 62, "gaussian")
    sf, sp, stats, K = pp.aggregate_wsi_super_patches(_cuda(feats), _cuda(pos), 6, 1.0, 1.0 / 128)
    assert K is None and sf.shape == (6, d) and sp.shape == (6, 2)
    Kref = oracle.combined_similarity(feats, pos, 1.0, 1.0 / 128)
    ref = oracle.similarity_stats(Kref)
    for key in ref:
        assert np.isclose(stats["wsi_similarity_matrix_stats"][key], ref[key], rtol=RTOL), key
    monkeypatch.setattr(pp, "MAX_DENSE_N", 32768)
    _, _, stats_dense, K2 = pp.aggregate_wsi_super_patches(_cuda(feats), _cuda(pos), 6, 1.0, 1.0 / 128)
    assert K2 is not None
    assert np.isclose(stats["avg_intra_cluster_similarity"], stats_dense["avg_intra_cluster_similarity"], rtol=1e-5)


@pytest.mark.parametrize("n,d,with_pos", [(4096, 1024, False), (5000, 256, True), (4100, 100, True), (6000, 1536, True)])
def test_dense_tc_default_routing_accuracy(n, d, with_pos, monkeypatch):
    """A1 / A2 of square inputs from 4,096 rows up run on the tensor cores by default (dense-output mode of the
    fused kernel: hi*hi in one TMEM accumulator, the small products in a second one, exact diagonal) and hold the
    same 1e-5 relative bar against float64 as the FP32 SIMT tiles they replace; HG_B200_DENSE_TC_MIN_N=0 (here:
    the module attribute) switches back."""
    feats, pos = synth.make_slide(n, d, 41, "gaussian")
    lam_h, lam_g = 1.0, (1.0 / 4096 if with_pos else 0.0)          # exponents stay below ~10: FP32 resolves them to 1e-6
    x, p = _cuda(feats), (_cuda(pos) if with_pos else None)
    assert ops.DENSE_TC_MIN_N == 4096 and ops._dense_on_tensor_cores(None, None, n, d, lam_h, _lib.FORM_EXPANSION)
    assert not ops._dense_on_tensor_cores(None, None, 4095, d, lam_h, _lib.FORM_EXPANSION)
    K = ops.similarity_dense(x, pos_a=p, pos_b=p, lambda_h=lam_h, lambda_g=lam_g)
    monkeypatch.setattr(ops, "DENSE_TC_MIN_N", 0)
    K_simt = ops.similarity_dense(x, pos_a=p, pos_b=p, lambda_h=lam_h, lambda_g=lam_g)
    monkeypatch.setattr(ops, "DENSE_TC_MIN_N", 4096)
    assert not torch.equal(K, K_simt)                                 # two different kernels ran
    assert bool((torch.diagonal(K) == 1.0).all())                     # d2(i, i) = 0 exactly
    rows = np.random.default_rng(2).choice(n, 512, replace=False)
    x64 = feats.astype(np.float64)
    d2 = oracle.sq_dists_f64(x64[rows], x64)
    if with_pos:
        p64 = pos.astype(np.float64)
        d2 = lam_h * d2 + lam_g * ((p64[rows, None, :] - p64[None, :, :]) ** 2).sum(-1)
    ref = np.exp(-d2)
    sel = torch.from_numpy(rows).cuda()
    np.testing.assert_allclose(K_simt[sel].cpu().numpy(), ref, rtol=RTOL, atol=1e-30)
    np.testing.assert_allclose(K[sel].cpu().numpy(), ref, rtol=RTOL, atol=1e-30)         # the same bar on the tensor cores
    if with_pos:
        Kc = hg.compute_combined_similarity(x, p, lam_h, lam_g)       # the reference-facing entry takes the same path
        assert torch.equal(Kc, K)


@pytest.mark.parametrize("m,n,d,with_pos", [(1000, 700, 96, True), (513, 2049, 1024, False), (640, 640, 33, True), (4096, 4096, 256, False)])
def test_dense_128_tile_kernel_equals_the_64_tile_kernel(m, n, d, with_pos):
    """outputs of at least 512 x 512 run on the 128 x 128-tile kernel; slices below that size run on the 64 x 64 one:
    same k-ordered FMA chains, so identical bits (and the float64 kernel within 1e-5)"""
    a = synth.make_features(m, d, 51, "gaussian")
    b = synth.make_features(n, d, 52, "gaussian")
    pa, pb = (synth.make_positions(m, 53), synth.make_positions(n, 54)) if with_pos else (None, None)
    kw = dict(lambda_h=1.0, lambda_g=1.0 / 4096 if with_pos else 0.0)
    A, B = _cuda(a), _cuda(b)
    PA, PB = (_cuda(pa), _cuda(pb)) if with_pos else (None, None)
    K = ops.similarity_dense(A, B, pos_a=PA, pos_b=PB, **kw)
    for r0 in range(0, m, 400):                                    # 400-row slices: the 64 x 64-tile kernel
        part = ops.similarity_dense(A[r0:r0 + 400].contiguous(), B, pos_a=PA[r0:r0 + 400].contiguous() if with_pos else None,
                                    pos_b=PB, **kw)
        assert torch.equal(K[r0:r0 + 400], part)
    rows = np.random.default_rng(6).choice(m, 64, replace=False)
    d2 = oracle.sq_dists_f64(a[rows].astype(np.float64), b.astype(np.float64))
    if with_pos:
        d2 = d2 + kw["lambda_g"] * ((pa[rows, None, :].astype(np.float64) - pb[None, :, :].astype(np.float64)) ** 2).sum(-1)
    np.testing.assert_allclose(K[torch.from_numpy(rows).cuda()].cpu().numpy(), np.exp(-d2), rtol=RTOL, atol=1e-30)
