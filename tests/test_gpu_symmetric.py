"""Symmetric mode of the fused similarity + top-k stage (knn_tc.cu): half of the tiles are computed,
each serves (i, j) and (j, i).  Its neighbour lists must equal the full pass bit for bit, and the
oracle on sampled rows; the bucket-overflow fallback must give the same answer too."""
import ctypes
import os

import numpy as np
import pytest
import torch

from hg_b200 import _lib, ops, synth
from test_gpu_parity import _cuda, check_knn

pytestmark = pytest.mark.gpu


class _env:
    def __init__(self, **kv):
        self.kv = {k: str(v) for k, v in kv.items()}

    def __enter__(self):
        self.old = {k: os.environ.get(k) for k in self.kv}
        os.environ.update(self.kv)

    def __exit__(self, *exc):
        for k, v in self.old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def _plan(n, d, k, r0=0, r1=None):
    lib = _lib.load_library()
    stride, flop = ctypes.c_int(0), ctypes.c_double(0)
    _lib.check(lib.hg_knn_feature_plan(n, d, k, r0, n if r1 is None else r1, ctypes.byref(stride), ctypes.byref(flop)))
    return stride.value, flop.value


def _both(x, k, metric, stride=8, **kw):
    with _env(HG_B200_TC_SYM=0):
        full = ops.knn_feature(x, k, metric, **kw)
    with _env(HG_B200_TC_SYM=stride):          # forced: small test shapes are below the automatic cut-over
        sym = ops.knn_feature(x, k, metric, **kw)
    torch.cuda.synchronize()
    return full, sym


@pytest.mark.parametrize("n,d,k,metric,with_pos", [
    (3000, 256, 8, "gaussian", False),          # tiny, forced
    (8192, 256, 8, "gaussian", False),
    (12345, 200, 16, "gaussian", True),         # ragged n and d, spatial term in the epilogue
    (20000, 512, 8, "cosine", False),
    (33000, 128, 5, "euclidean", False),        # stride 16
])
def test_symmetric_equals_full_pass(n, d, k, metric, with_pos):
    with _env(HG_B200_TC_SYM=8):
        stride, flop = _plan(n, d, k)
    assert stride == 8
    n_pad, d_pad = -(-n // 256) * 256, -(-d // 64) * 64
    if n >= 8192:                                                # tile granularity dominates below
        assert flop < 0.7 * 3 * 2.0 * n_pad * n_pad * d_pad      # about 0.5 + 1/8 of the full pass
    feats, pos = synth.make_slide(n, d, 11, "cosine" if metric == "cosine" else "gaussian")
    kw = dict(positions=_cuda(pos), lambda_h=1.0, lambda_g=1.0 / 128) if with_pos else {}
    (fi, fv), (si, sv) = _both(_cuda(feats), k, metric, **kw)
    assert torch.equal(fi, si)
    assert torch.equal(fv, sv)
    rows = np.random.default_rng(0).choice(n, 256, replace=False)
    check_knn(si.cpu().numpy()[rows], feats, k, metric, rows=rows, pos=pos if with_pos else None,
              lambda_h=1.0, lambda_g=1.0 / 128 if with_pos else 0.0)


@pytest.mark.parametrize("n,d,k,metric,with_pos", [(12_000, 1024, 16, "gaussian", True), (9_500, 1100, 8, "cosine", False)])
def test_symmetric_variants_give_the_same_lists(n, d, k, metric, with_pos):
    """The symmetric pass comes as single CTAs or CTA pairs (cta_group::2, the default from D = 1024 up), with two or
    three tensor-core products per k-step, and B tiles with or without lo halves: the exact re-rank decides, so every
    variant returns the lists of the classic full pass bit for bit."""
    feats, pos = synth.make_slide(n, d, 17, "cosine" if metric == "cosine" else "gaussian")
    kw = dict(positions=_cuda(pos), lambda_h=1.0, lambda_g=1.0 / 128) if with_pos else {}
    x = _cuda(feats)
    with _env(HG_B200_TC_SYM=0):
        ref_i, ref_v = ops.knn_feature(x, k, metric, **kw)
    for env in (dict(), dict(HG_B200_TC_SYM_CG=1), dict(HG_B200_TC_SYM_CG=2), dict(HG_B200_TC_SYM_PRODS=3),
                dict(HG_B200_TC_SYM_CG=1, HG_B200_TC_SYM_PRODS=3), dict(HG_B200_TC_B_HI=0), dict(HG_B200_TC_MC=1, HG_B200_TC_SYM_CG=1),
                dict(HG_B200_TC_PRE_HI=1)):
        with _env(**env):
            assert _plan(n, d, k)[0] == 8                          # symmetric by default at this size
            idx, val = ops.knn_feature(x, k, metric, **kw)
        assert torch.equal(idx, ref_i) and torch.equal(val, ref_v), env


@pytest.mark.parametrize("cc", [2, 4])
def test_prepass_with_column_splits(cc):
    """Large D: the threshold pre-pass runs 2 or 4 column splits side by side and thr_merge_kernel
    combines their lists (forced here on a small shape)."""
    n, d, k = 9000, 192, 8
    feats, pos = synth.make_slide(n, d, 13, "gaussian")
    with _env(HG_B200_TC_PRE_CC=cc):
        (fi, fv), (si, sv) = _both(_cuda(feats), k, "gaussian", stride=2, positions=_cuda(pos), lambda_h=1.0,
                                   lambda_g=1.0 / 128)
    assert torch.equal(fi, si) and torch.equal(fv, sv)


def test_row_ranges_and_small_slides_use_the_full_pass():
    assert _plan(50000, 1024, 16, 0, 25000)[0] == 0
    assert _plan(4096, 1024, 8)[0] == 0
    assert _plan(20000, 1024, 8)[0] == 8
    with _env(HG_B200_TC_SYM=0):
        assert _plan(50000, 1024, 16)[0] == 0
    stride, flop = _plan(50000, 1024, 16)
    lib = _lib.load_library()
    main, pre = ctypes.c_int(0), ctypes.c_int(0)
    _lib.check(lib.hg_knn_feature_products(50000, 1024, 16, 0, 50000, ctypes.byref(main), ctypes.byref(pre)))
    assert (main.value, pre.value) == (2, 2)                           # symmetric passes: hi*hi + hi*lo
    assert stride == 16 and 0.5 < flop / (main.value * 2.0 * 50000.0 ** 2 * 1024) < 0.6            # ~0.5 + 1/16
    _lib.check(lib.hg_knn_feature_products(4096, 1024, 8, 0, 4096, ctypes.byref(main), ctypes.byref(pre)))
    assert (main.value, pre.value) == (2, 0)                           # classic pass over a small slide: two as well
    _lib.check(lib.hg_knn_feature_products(50000, 1024, 16, 0, 25000, ctypes.byref(main), ctypes.byref(pre)))
    assert (main.value, pre.value) == (3, 0)                           # row block of a large slide: all three products
    with _env(HG_B200_TC_SYM_PRODS=3):
        _lib.check(lib.hg_knn_feature_products(50000, 1024, 16, 0, 50000, ctypes.byref(main), ctypes.byref(pre)))
        assert (main.value, pre.value) == (3, 3)


def test_bucket_overflow_falls_back_to_the_full_pass():
    n, d, k = 9000, 128, 8
    feats = synth.make_features(n, d, 5, "gaussian")
    x = _cuda(feats)
    with _env(HG_B200_TC_SYM=0):
        fi, fv = ops.knn_feature(x, k, "gaussian")
    with _env(HG_B200_TC_BKT_CAP=4, HG_B200_TC_SYM=8):      # every bucket overflows -> guarded full pass runs
        si, sv = ops.knn_feature(x, k, "gaussian")
    torch.cuda.synchronize()
    assert torch.equal(fi, si) and torch.equal(fv, sv)


def test_symmetric_with_duplicate_rows_and_outlier_norms():
    """Exact duplicates give score ties at the admission threshold; one huge-norm row stretches the
    error slack of every threshold.  Sets must still match the full pass."""
    n, d, k = 10000, 96, 8
    rng = np.random.default_rng(3)
    feats = synth.make_features(n, d, 9, "gaussian")
    dup = rng.choice(n, 2000, replace=False)
    feats[dup] = feats[rng.choice(n, 2000)]                  # many exact copies
    feats[17] *= 300.0
    (fi, fv), (si, sv) = _both(_cuda(feats), k, "gaussian")
    a, b = np.sort(fi.cpu().numpy(), 1), np.sort(si.cpu().numpy(), 1)
    differ = np.nonzero((a != b).any(1))[0]
    # only tie rows (equal k-th and (k+1)-th exact value) may choose a different copy
    fvn, svn = fv.cpu().numpy(), sv.cpu().numpy()
    np.testing.assert_array_equal(fvn[~np.isin(np.arange(n), differ)], svn[~np.isin(np.arange(n), differ)])
    for r in differ:
        np.testing.assert_array_equal(fvn[r], svn[r])        # same values, different duplicate picked


@pytest.mark.parametrize("n,d,k,metric", [(3000, 200, 8, "gaussian"), (9000, 256, 8, "cosine"), (2500, 130, 5, "euclidean")])
def test_build_weights_from_knn_lists_equal_direct_cosine(n, d, k, metric):
    """The one-call build takes edge weights from the re-rank's cosine by-product; edges that come
    from the spatial lists or from cliques fall back to the row gather.  Must equal
    hg_edge_weights_cosine bit for bit."""
    import hg_b200 as hg
    feats, pos = synth.make_slide(n, d, 21, "cosine" if metric == "cosine" else "gaussian")
    labels = np.random.default_rng(1).integers(0, 40, n).astype(np.int32)       # 40 clusters, 6 kept as cliques
    x = _cuda(feats)
    out = hg.knn_hyperedges(x, _cuda(pos), k=k, kernel=metric, lambda_h=1.0, lambda_g=1.0 / 128, spatial_k=4,
                            labels=labels, num_hyperedges=6)
    w_direct = ops.edge_weights_cosine(x, out["edge_index"])
    torch.cuda.synchronize()
    assert out["edge_index"].shape[1] > n * k // 2
    assert torch.equal(out["edge_weights"], w_direct)


@pytest.mark.parametrize("n,d,k,metric,with_pos,sym", [
    (3000, 256, 8, "gaussian", False, 0),          # classic full pass, odd number of row blocks
    (5000, 192, 16, "cosine", False, 0),
    (12345, 200, 16, "gaussian", True, 8),         # symmetric half pass, ragged n and d, spatial term
    (20000, 512, 8, "gaussian", False, 8),
])
def test_multicast_pairs_equal_single_ctas(n, d, k, metric, with_pos, sym):
    """HG_B200_TC_MC=1: two CTAs of a cluster share every B tile through TMA multicast.  Same lists."""
    feats, pos = synth.make_slide(n, d, 5 + n, "cosine" if metric == "cosine" else "gaussian")
    x = _cuda(feats)
    kw = dict(positions=_cuda(pos), lambda_h=1.0, lambda_g=1.0 / 128) if with_pos else {}
    with _env(HG_B200_TC_SYM=sym, HG_B200_TC_MC=0):
        ref = ops.knn_feature(x, k, metric, **kw)
    with _env(HG_B200_TC_SYM=sym, HG_B200_TC_MC=1):
        got = ops.knn_feature(x, k, metric, **kw)
        part = ops.knn_feature(x, k, metric, rows=(1000, 2111), **kw)      # row range: classic pass on pairs
    torch.cuda.synchronize()
    assert torch.equal(got[0], ref[0]) and torch.equal(got[1], ref[1])
    assert torch.equal(part[0], ref[0][1000:2111]) and torch.equal(part[1], ref[1][1000:2111])
