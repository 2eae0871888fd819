"""Uniform-grid path of the spatial kNN (knn_spatial.cu): same lists as the brute-force pass and as
the oracle, bit for bit, on patch lattices (massive ties), pixel-unit lattices, arbitrary float
coordinates, degenerate boxes, duplicate-heavy inputs (fallback) and row ranges."""
import numpy as np
import pytest
import torch

import oracle_np as oracle
from hg_b200 import ops, synth
from test_gpu_parity import _cuda
from test_gpu_symmetric import _env

pytestmark = pytest.mark.gpu


def _brute(pos, k, **kw):
    with _env(HG_B200_SP_GRID=0):
        out = ops.knn_spatial(_cuda(pos), k, **kw)
    torch.cuda.synchronize()
    return out


def _grid(pos, k, cell=None, **kw):
    env = dict(HG_B200_SP_GRID=1)
    if cell is not None:
        env["HG_B200_SP_CELL"] = cell
    with _env(**env):
        out = ops.knn_spatial(_cuda(pos), k, **kw)
    torch.cuda.synchronize()
    return out


def _same(a, b):
    assert torch.equal(a[0], b[0])
    va, vb = a[1].cpu().numpy(), b[1].cpu().numpy()
    assert ((va == vb) | (np.isnan(va) & np.isnan(vb))).all()


@pytest.mark.parametrize("n,k,stride", [(700, 8, 1.0), (5000, 8, 1.0), (20000, 8, 256.0), (30000, 16, 1.0), (9000, 32, 1.0),
                                        (12000, 5, 0.37), (12000, 5, 0.5)])
def test_grid_vs_oracle_on_patch_lattices(n, k, stride):
    pos = synth.make_positions(n, 3 * n + k, stride)
    idx, d2 = _grid(pos, k)
    _same((idx, d2), _brute(pos, k))
    if stride != 0.37:       # 0.37 is not a binary fraction: FP32 distances round, the float64 oracle's ties differ
        ref_idx, ref_d2 = oracle.knn_spatial(pos, k)
        assert (idx.cpu().numpy() == ref_idx).all()
        assert (d2.cpu().numpy() == ref_d2.astype(np.float32)).all()


def test_default_cut_over_uses_grid_and_matches():
    """n >= 8192 takes the grid by default; below it the brute-force pass.  Same answer either way."""
    for n in (8191, 8192, 50000):
        pos = synth.make_positions(n, n)
        out = ops.knn_spatial(_cuda(pos), 8)
        _same(out, _brute(pos, 8))


@pytest.mark.parametrize("cell", [0.5, 1, 4, 16])
def test_grid_cell_size_does_not_change_results(cell):
    pos = synth.make_positions(15000, 11)
    _same(_grid(pos, 8, cell=cell), _brute(pos, 8))


def test_grid_arbitrary_float_coordinates():
    rng = np.random.default_rng(5)
    n = 25000
    uni = (rng.random((n, 2)) * 1000.0 - 300.0).astype(np.float32)                     # uniform, negative coordinates
    blobs = np.concatenate([rng.normal(0, 1, (n // 2, 2)), rng.normal(500, 30, (n - n // 2, 2))]).astype(np.float32)
    for pos in (uni, blobs):
        _same(_grid(pos, 8), _brute(pos, 8))
        _same(_grid(pos, 16, cell=3), _brute(pos, 16))


def test_grid_degenerate_boxes():
    n = 10000
    line = np.stack([np.arange(n, dtype=np.float32) * 2.0, np.full(n, 7.0, np.float32)], 1)     # zero height
    _same(_grid(line, 8), _brute(line, 8))
    col = line[:, ::-1].copy()                                                                  # zero width
    _same(_grid(col, 8), _brute(col, 8))
    thin = np.stack([np.arange(n, dtype=np.float32), (np.arange(n) % 2).astype(np.float32)], 1)
    _same(_grid(thin, 8), _brute(thin, 8))
    same = np.full((9000, 2), 3.0, np.float32)                       # all points coincide: one cell -> fallback
    idx, d2 = _grid(same, 8)
    _same((idx, d2), _brute(same, 8))
    assert (d2 == 0).all() and (idx[1:, 0] == 0).all() and (idx[0, 0] == 1).all()


def test_grid_duplicates_and_fallback():
    rng = np.random.default_rng(9)
    n = 20000
    base = synth.make_positions(n, 77)
    dup = base.copy()
    dup[rng.integers(0, n, 4000)] = base[rng.integers(0, n, 4000)]    # scattered duplicates: ties at d2 = 0
    _same(_grid(dup, 8), _brute(dup, 8))
    heavy = base.copy()
    heavy[: n // 4] = base[0]                                         # one cell holds n/4 points: brute-force fallback
    _same(_grid(heavy, 8), _brute(heavy, 8))


def test_grid_row_range_and_small_n():
    pos = synth.make_positions(16000, 123)
    full = _grid(pos, 8)
    part = _grid(pos, 8, rows=(5000, 9001))
    assert torch.equal(part[0], full[0][5000:9001]) and torch.equal(part[1], full[1][5000:9001])
    tiny = synth.make_positions(6, 1)
    idx, d2 = _grid(tiny, 8)                                          # fewer candidates than k: padded with -1 / NaN
    assert (idx[:, 5:] == -1).all() and (idx[:, :5] >= 0).all() and torch.isnan(d2[:, 5:]).all()
    _same((idx, d2), _brute(tiny, 8))


def test_grid_nan_coordinates_do_not_poison_other_rows():
    pos = synth.make_positions(12000, 5)
    pos[100] = np.nan
    g = _grid(pos, 8)
    _same(g, _brute(pos, 8))
    assert (g[0][100] == -1).all()
