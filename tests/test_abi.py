"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol
that include/hg_b200.h declares, and the Python surface mirrors the reference's names.
No compute calls (there is no GPU here)."""
import ctypes
import inspect
import os
import re

import pytest
import torch

import hg_b200
from hg_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFERENCE_EXPORTS = [       # reference build_hypergraph/__init__.py:28-46
    "compute_morphological_similarity", "compute_spatial_similarity", "compute_combined_similarity",
    "build_weighted_hypergraph", "mean_pool_with_similarity", "process_single_file", "process_dataset",
    "load_wsi_data", "load_tma_data", "aggregate_wsi_super_patches", "compute_wsi_tma_similarity",
    "group_by_similarity", "build_hypergraph_knn_kmeans", "save_hypergraph_to_h5",
    "load_similarity_matrices", "rebuild_hypergraph_from_similarity", "batch_rebuild_hypergraph",
]


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "hg_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hg_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load_library()
    syms = _header_symbols()
    assert len(syms) >= 20
    for name in syms:
        assert hasattr(lib, name), f"{name} declared in hg_b200.h but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature"
    assert sorted(_lib.SIGNATURES) == syms
    assert lib.hg_version() >= 100


def test_error_convention_without_gpu():
    lib = _lib.load_library()
    out = ctypes.c_size_t(0)
    assert lib.hg_knn_feature_workspace_bytes(1, 16, 4, 0, 1, ctypes.byref(out)) == _lib.HG_ERR_BAD_ARG
    assert b"greater than 1" in lib.hg_last_error()
    with pytest.raises(ValueError):
        _lib.check(_lib.HG_ERR_BAD_ARG)
    assert lib.hg_knn_feature_workspace_bytes(50_000, 1024, 16, 0, 50_000, ctypes.byref(out)) == 0
    assert out.value > 2 * 50_000 * 1024 * 2        # operand planes + lists
    assert lib.hg_operand_bytes(4096, 1024) == 2 * 4096 * 1024 * 2


def test_python_surface_mirrors_reference():
    for name in REFERENCE_EXPORTS:
        assert name in hg_b200.__all__ and callable(getattr(hg_b200, name))
    import build_hypergraph                      # alias package with the reference's name
    assert build_hypergraph.build_hypergraph_knn_kmeans is hg_b200.build_hypergraph_knn_kmeans
    sig = inspect.signature(hg_b200.build_hypergraph_knn_kmeans)
    assert list(sig.parameters)[:6] == ["wsi_features", "tma_features", "group_labels", "k", "num_hyperedges", "device"]
    assert sig.parameters["k"].default == 5 and sig.parameters["num_hyperedges"].default == 10
    sig = inspect.signature(hg_b200.build_weighted_hypergraph)
    assert list(sig.parameters) == ["features", "positions", "lambda_h", "lambda_g", "threshold_median_ratio", "device"]
    sig = inspect.signature(hg_b200.process_single_file)
    params = list(sig.parameters.values())
    assert [p.default for p in params[1:8]] == [100, 10, 5, 10, 1.0, 1.0, None]
    assert all(p.kind is inspect.Parameter.KEYWORD_ONLY for p in params[8:])      # extras never shift the reference's arguments


def test_products_per_pass_by_shape(monkeypatch):
    """host-side plan logic (no GPU): two MMAs per product in the symmetric passes and in classic launches over small
    slides, three in classic launches over more than 16,384 columns; the issued-MMA count follows"""
    lib = _lib.load_library()
    main, pre = ctypes.c_int(-1), ctypes.c_int(-1)
    for n, d, k, r1, want in ((50_000, 1024, 16, 50_000, (2, 2)), (100_000, 1536, 16, 100_000, (2, 2)),
                              (250_000, 1536, 16, 250_000, (2, 2)), (4096, 1024, 8, 4096, (2, 0)),
                              (100_000, 1536, 16, 12_500, (3, 0))):
        assert lib.hg_knn_feature_products(n, d, k, 0, r1, ctypes.byref(main), ctypes.byref(pre)) == 0
        assert (main.value, pre.value) == want, (n, d, k, r1)
    stride, flop2, flop3 = ctypes.c_int(0), ctypes.c_double(0), ctypes.c_double(0)
    assert lib.hg_knn_feature_plan(50_000, 1024, 16, 0, 50_000, ctypes.byref(stride), ctypes.byref(flop2)) == 0
    monkeypatch.setenv("HG_B200_TC_SYM_PRODS", "3")
    assert lib.hg_knn_feature_plan(50_000, 1024, 16, 0, 50_000, ctypes.byref(stride), ctypes.byref(flop3)) == 0
    assert stride.value == 16 and abs(flop3.value / flop2.value - 1.5) < 1e-9
    assert lib.hg_knn_feature_products(1, 16, 4, 0, 1, ctypes.byref(main), ctypes.byref(pre)) == _lib.HG_ERR_BAD_ARG


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        hg_b200.compute_morphological_similarity(torch.randn(4, 8))
    with pytest.raises(RuntimeError):
        hg_b200.build_hypergraph_knn_kmeans(torch.randn(8, 4), torch.randn(8, 4), None)


def test_dense_tensor_core_routing_is_square_only(monkeypatch):
    """A1 / A2 dense outputs of 4,096 rows and more run on the tensor cores (1e-5 bar held, test_gpu_parity.py);
    HG_B200_DENSE_TC_MIN_N=0 keeps them on the FP32 kernels; only square expansion-form kernels of one matrix
    against itself qualify"""
    from hg_b200 import _lib, ops
    fake = object()
    assert ops.DENSE_TC_MIN_N == 4096
    monkeypatch.setattr(ops, "DENSE_TC_MIN_N", 0)
    assert not ops._dense_on_tensor_cores(None, None, 50000, 1024, 1.0, _lib.FORM_EXPANSION)
    monkeypatch.setattr(ops, "DENSE_TC_MIN_N", 4096)
    assert ops._dense_on_tensor_cores(None, None, 4096, 1024, 1.0, _lib.FORM_EXPANSION)
    assert not ops._dense_on_tensor_cores(None, None, 4095, 1024, 1.0, _lib.FORM_EXPANSION)     # below the cut-over
    assert not ops._dense_on_tensor_cores(fake, None, 8192, 1024, 1.0, _lib.FORM_EXPANSION)     # rectangular (A7)
    assert not ops._dense_on_tensor_cores(None, fake, 8192, 1024, 1.0, _lib.FORM_EXPANSION)     # other positions
    assert not ops._dense_on_tensor_cores(None, None, 8192, 1024, 1.0, _lib.FORM_DIRECT)        # direct-difference form
    assert not ops._dense_on_tensor_cores(None, None, 8192, 1024, 0.0, _lib.FORM_EXPANSION)     # lambda_h = 0
    assert not ops._dense_on_tensor_cores(None, None, 8192, 2, 1.0, _lib.FORM_EXPANSION)        # K_g alone (D = 2)
