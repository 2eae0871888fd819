"""Drop-in for the reference's ``build_hypergraph/similarity_kernel.py``.

Same names, argument order, defaults and return types; the arithmetic runs in
hand-written sm_100a CUDA behind the C ABI (no CPU fallback).

    kappa_h(i,j) = exp(-lambda_h |h_i - h_j|^2)        reference :17-54
    kappa_g(i,j) = exp(-lambda_g |g_i - g_j|^2)        reference :57-86
    kappa        = kappa_h * kappa_g                   reference :88-124

The dense-matrix functions are meant for small N (they return N x N); large
slides go through :func:`knn_hyperedges`, which never materialises N x N.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import numpy as np
import torch

from . import _lib as L
from . import ops


def _dev(*tensors, device=None):
    if device is not None:
        return L.require_cuda(device)
    for t in tensors:
        if t is not None and t.is_cuda:
            return t.device
    return L.require_cuda(None)


def compute_morphological_similarity(features: torch.Tensor, lambda_h: float = 1.0) -> torch.Tensor:
    """K_h [N, N] (reference similarity_kernel.py:17).  Expansion form, un-clamped, FP32."""
    return ops.similarity_dense(features, lambda_h=lambda_h, lambda_g=0.0, device=_dev(features))


def compute_spatial_similarity(positions: torch.Tensor, lambda_g: float = 1.0) -> torch.Tensor:
    """K_g [N, N] (reference similarity_kernel.py:57).  Positions are the 'features' of a
    2-/3-dimensional Gaussian kernel; direct-difference form (exact on integer grids)."""
    return ops.similarity_dense(positions, lambda_h=lambda_g, lambda_g=0.0, form=L.FORM_DIRECT,
                                device=_dev(positions))


def compute_combined_similarity(features: torch.Tensor, positions: torch.Tensor,
                                lambda_h: float = 1.0, lambda_g: float = 1.0) -> torch.Tensor:
    """K = K_h * K_g (reference similarity_kernel.py:88), fused into one exp."""
    return ops.similarity_dense(features, pos_a=positions, pos_b=positions, lambda_h=lambda_h,
                                lambda_g=lambda_g, device=_dev(features, positions))


def _ceil_f32(x: float) -> float:
    """Smallest float32 >= x: comparing float32 K against it equals the reference's
    double-precision comparison `K[i, j].item() < threshold` (similarity_kernel.py:195-198)."""
    t = np.float32(x)
    if float(t) < x:
        t = np.nextafter(t, np.float32(np.inf))
    return float(t)


def build_weighted_hypergraph(features: torch.Tensor, positions: torch.Tensor,
                              lambda_h: float = 1.0, lambda_g: float = 1.0,
                              threshold_median_ratio: float = None,
                              device: Optional[torch.device] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Dense-threshold builder (reference similarity_kernel.py:126-212).

    threshold = lower-median(off-diagonal K) * ratio; every ordered (i, j) -- diagonal and
    both directions included -- with K >= threshold is kept, row-major.  ratio=None keeps
    every pair (the reference's documented intent; its line :188 raises TypeError instead).
    Returns (edge_index int64 [2, E], edge_weights float32 [E]) on the CUDA device.
    """
    dev = _dev(features, positions, device=device)
    K = ops.similarity_dense(features, pos_a=positions, pos_b=positions, lambda_h=lambda_h,
                             lambda_g=lambda_g, device=dev)
    n = K.shape[0]
    if n <= 1:
        raise ValueError(f"Number of nodes must be greater than 1, got N={n}. "
                         f"Hypergraph construction requires at least 2 nodes.")
    if threshold_median_ratio is None:
        threshold = float("-inf")
    else:
        median_sim = ops.similarity_stats(K, exclude_diagonal=True)["median"]
        threshold = _ceil_f32(median_sim * threshold_median_ratio)
    return ops.threshold_edges(K, threshold)


def mean_pool_with_similarity(features: torch.Tensor) -> torch.Tensor:
    """Global mean feature [1, D] (reference similarity_kernel.py:214-238; plain torch there too)."""
    return torch.mean(features, dim=0, keepdim=True)


def build_hypergraph_data(features: torch.Tensor, positions: torch.Tensor,
                          lambda_h: float = 1.0, lambda_g: float = 1.0,
                          threshold_median_ratio: float = None, use_pooling: bool = True,
                          device: Optional[torch.device] = None) -> dict:
    """Dict packer (reference similarity_kernel.py:240-306)."""
    dev = _dev(features, positions, device=device)
    features = features.to(dev)
    positions = positions.to(dev)
    edge_index, edge_weights = build_weighted_hypergraph(
        features, positions, lambda_h, lambda_g, threshold_median_ratio, dev)
    result = {"x": features, "edge_index": edge_index, "edge_attr": edge_weights, "pos": positions}
    if use_pooling:
        result["pooled_feature"] = mean_pool_with_similarity(features)
    return result


# ------------------------------------------------------------------------------------------
# The fused path (new capability): kNN hyperedges straight from embeddings / coordinates.
# ------------------------------------------------------------------------------------------
def knn_hyperedges(features: torch.Tensor, positions: Optional[torch.Tensor] = None, *,
                   k: int = 16, kernel: str = "gaussian", lambda_h: float = 1.0, lambda_g: float = 0.0,
                   spatial_k: int = 0, labels=None, num_hyperedges: int = 0,
                   return_csr: bool = False, rows: Optional[Tuple[int, int]] = None,
                   device: Optional[torch.device] = None) -> dict:
    """Feature-space kNN (+ optional spatial-grid kNN and cluster cliques) -> de-duplicated
    ``edge_index`` / cosine ``edge_weights`` (+ CSR incidence), without the N x N matrix.

    kernel: 'gaussian' (top-k of K_h * K_g), 'cosine', or 'euclidean' (the reference's
    sklearn call, preprocess_hypergraph.py:379-382).  The neighbour lists are also returned
    (``knn_idx``/``knn_val``, ``spatial_idx``/``spatial_d2``).
    """
    dev = _dev(features, positions, device=device)
    x = L.f32c(features, dev)
    n = x.shape[0]
    pos = L.f32c(positions, dev) if positions is not None else None
    full = rows is None or (rows[0] == 0 and rows[1] == n)
    if full and k <= 32:
        # whole slide: one C call, one stream synchronisation
        return ops.build_knn_hypergraph(x, k, kernel, positions=pos, lambda_h=lambda_h,
                                        lambda_g=lambda_g if kernel == "gaussian" else 0.0, spatial_k=spatial_k,
                                        labels=labels, num_hyperedges=num_hyperedges, return_csr=return_csr, device=dev)
    out = {}
    idx, val = ops.knn_feature(x, k, kernel, positions=pos if (kernel == "gaussian" and lambda_g) else None,
                               lambda_h=lambda_h, lambda_g=lambda_g, rows=rows, device=dev)
    out["knn_idx"], out["knn_val"] = idx, val
    lists = [idx]
    if spatial_k and pos is not None:
        sidx, sd2 = ops.knn_spatial(pos, spatial_k, rows=rows, device=dev)
        out["spatial_idx"], out["spatial_d2"] = sidx, sd2
        lists.append(sidx)
    if rows is not None and (rows[0] != 0 or rows[1] != n):
        return out                              # row-partitioned call: assembly happens after the gather
    ei = ops.build_edges(lists, n, labels, num_hyperedges, device=dev)
    out["edge_index"] = ei
    out["edge_weights"] = ops.edge_weights_cosine(x, ei)
    if return_csr:
        out["csr_rowptr"], out["csr_col"], out["csr_eid"] = ops.csr_from_coo(ei, n)
    return out
