"""Staged kNN (hg_knn_staged_*): candidate columns handed over panel by panel, in any order, must give
exactly what the one-shot hg_knn_feature call gives for the same row range -- the single-GPU half of the
overlapped row-partitioned path (the exchange itself is covered by tests/test_multigpu_nccl.py)."""
import numpy as np
import pytest
import torch

import oracle_np as oracle
from hg_b200 import ops, synth
from test_gpu_parity import _cuda

pytestmark = pytest.mark.gpu


def _panels(n, width):
    return [(a, min(n, a + width)) for a in range(0, n, width)]


def _run_staged(x, k, metric, rows, panels, order, pos=None, lambda_h=1.0, lambda_g=0.0, return_cos=True, lanes=2):
    n, d = x.shape
    st = ops.StagedKnn(n, d, k, metric, rows, max(b - a for a, b in panels), positions=pos,
                       lambda_h=lambda_h, lambda_g=lambda_g, lanes=lanes, device=x.device)
    for p in order:
        a, b = panels[p]
        st.add_panel(x[a:b], a, b)
    return st.finish(x, return_cos=return_cos)


@pytest.mark.parametrize("n,d,k,metric,width,rows", [
    (3000, 128, 8, "euclidean", 768, (768, 1536)),
    (5000, 256, 16, "gaussian", 1280, (1280, 2560)),
    (4097, 96, 16, "cosine", 1024, (4096, 4097)),          # last panel: one row, owned by this "rank"
    (2600, 64, 32, "euclidean", 256, (512, 768)),           # k = 32: 32-entry lists, 11 panels
    (9000, 128, 5, "gaussian", 2304, (0, 2304)),
])
def test_staged_equals_one_shot(n, d, k, metric, width, rows):
    feats = synth.make_features(n, d, 100 + n, "cosine" if metric == "cosine" else "gaussian")
    x = _cuda(feats)
    panels = _panels(n, width)
    own = next(i for i, (a, b) in enumerate(panels) if a <= rows[0] and rows[1] <= b)
    rng = np.random.default_rng(n)
    others = [i for i in range(len(panels)) if i != own]
    ref = ops.knn_feature(x, k, metric, rows=rows, return_cos=True)
    for trial in range(3):
        order = [own] + list(rng.permutation(others))
        got = _run_staged(x, k, metric, rows, panels, order, lanes=(2, 1, 3)[trial])
        for g, r, name in zip(got, ref, ("idx", "val", "cos")):
            assert torch.equal(g, r), f"{name} differs (order {order})"
    # and against the float64 oracle on the row range
    oidx, _ = oracle.knn_feature(feats, k, metric, rows=np.arange(rows[0], rows[1]))
    same = (got[0].cpu().numpy() == oidx).all(axis=1).mean()
    assert same > 0.995                                      # near-ties aside (reported by the parity suite)


def test_staged_with_positions_and_missing_panel():
    n, d, k = 4000, 64, 8
    feats, pos = synth.make_slide(n, d, 7, "gaussian")
    x, p = _cuda(feats), _cuda(pos)
    panels = _panels(n, 1024)
    rows = (1024, 2048)
    lg = 1.0 / 256
    ref = ops.knn_feature(x, k, "gaussian", positions=p, lambda_h=1.0, lambda_g=lg, rows=rows, return_cos=True)
    got = _run_staged(x, k, "gaussian", rows, panels, [1, 3, 0, 2], pos=p, lambda_g=lg)
    assert all(torch.equal(g, r) for g, r in zip(got, ref))
    # a panel that never arrives contributes no candidates: same as kNN over the remaining columns
    got = _run_staged(x, k, "euclidean", rows, panels, [1, 0, 2])
    keep = np.r_[0:3072]
    oidx, _ = oracle.knn_feature(feats[keep], k, "euclidean", rows=np.arange(rows[0], rows[1]))
    assert (got[0].cpu().numpy() == oidx).all(axis=1).mean() > 0.995
    assert int(got[0].max()) < 3072


def test_staged_argument_checks():
    n, d = 3000, 64
    x = _cuda(synth.make_features(n, d, 3, "gaussian"))
    st = ops.StagedKnn(n, d, 8, "euclidean", (0, 1024), 1024, device=x.device)
    with pytest.raises(ValueError, match="first panel"):
        st.add_panel(x[1024:2048], 1024, 2048)
    st.add_panel(x[0:1024], 0, 1024)
    with pytest.raises(ValueError, match="multiples of 256"):
        st.add_panel(x[1000:2000], 1000, 2000)
    with pytest.raises(ValueError, match="multiples of 256"):
        st.add_panel(x[1024:2000], 1024, 2000)
    with pytest.raises(ValueError):
        st.add_panel(x[0:2048], 0, 2048)                     # wider than max_panel_cols
    with pytest.raises(ValueError):
        ops.StagedKnn(n, d, 33, "euclidean", (0, 1024), 1024, device=x.device)       # tensor-core path: k <= 32
