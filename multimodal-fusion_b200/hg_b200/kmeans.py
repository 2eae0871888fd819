"""KMeans on the GPU (scope row (f)1, "next"): `kmeans` is one C-ABI call (hg_kmeans_fit: seeding,
Lloyd iterations and the choice of the best restart stay on the device, deterministic for a seed);
`kmeans_stepwise` is the same estimator driven from Python over the building blocks
hg_kmeans_assign / hg_kmeans_update / hg_kmeans_candidate_d2 (kept for tests and tuning).

Mirrors the estimator the reference constructs at preprocess_hypergraph.py:150, :299, :391 --
``KMeans(n_clusters, random_state=42, n_init=10)`` with scikit-learn's defaults: greedy
k-means++ seeding (2 + log(k) local trials), Lloyd iterations until the summed squared centre
shift falls below ``tol * mean(var(X))`` or ``max_iter`` = 300, best of ``n_init`` runs by
inertia.  The random stream differs from scikit-learn's, so labels are not comparable one to one;
the parity bar is the inertia (tests/test_gpu_kmeans.py).  torch is used for the O(n) vector
plumbing (cumsum / searchsorted for the D^2 sampling); every O(n*d) step is a CUDA kernel of ours.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Tuple

import numpy as np
import torch

from . import _lib as L


def _kmeanspp(lib, x, n_clusters: int, gen: torch.Generator, st) -> torch.Tensor:
    n, d = x.shape
    dev = x.device
    trials = min(16, 2 + int(math.log(n_clusters)))
    centers = torch.empty(n_clusters, dtype=torch.int64, device=dev)
    first = torch.randint(n, (1,), generator=gen, device=dev)
    centers[0:1] = first
    cand_d2 = torch.empty((16, n), dtype=torch.float32, device=dev)
    L.check(lib.hg_kmeans_candidate_d2(L.ptr(x), n, d, L.ptr(first), 1, L.ptr(cand_d2), st))
    closest = cand_d2[0].clone()
    for c in range(1, n_clusters):
        cum = torch.cumsum(closest.double(), 0)
        r = torch.rand(trials, generator=gen, device=dev, dtype=torch.float64) * cum[-1]
        cand = torch.searchsorted(cum, r).clamp_(max=n - 1)
        L.check(lib.hg_kmeans_candidate_d2(L.ptr(x), n, d, L.ptr(cand), trials, L.ptr(cand_d2), st))
        pot = torch.minimum(cand_d2[:trials], closest.unsqueeze(0))
        best = torch.argmin(pot.double().sum(1))
        closest = pot[best].clone()
        centers[c] = cand[best]
    return x.index_select(0, centers).contiguous()


LAST_N_ITER = 0          # Lloyd iterations of the winning restart of the most recent `kmeans` call (tests, tuning)


def kmeans(features: torch.Tensor, n_clusters: int, n_init: int = 10, max_iter: int = 300, tol: float = 1e-4,
           seed: int = 42, device=None) -> Tuple[np.ndarray, torch.Tensor, float]:
    """Returns (labels int64 numpy [n], centers float32 CUDA [k, d], inertia)."""
    global LAST_N_ITER
    lib = L.load_library()
    dev = L.require_cuda(device if device is not None else (features.device if features.is_cuda else None))
    x = L.f32c(features, dev)
    n, d = x.shape
    with torch.cuda.device(dev):
        nbytes = C.c_size_t(0)
        L.check(lib.hg_kmeans_fit_workspace_bytes(n, d, int(n_clusters), C.byref(nbytes)))
        ws = L.Workspace.get(dev, nbytes.value, "kmeans_fit")
        labels = torch.empty(n, dtype=torch.int32, device=dev)
        centers = torch.empty((int(n_clusters), d), dtype=torch.float32, device=dev)
        inertia = torch.zeros(1, dtype=torch.float64, device=dev)
        n_iter = torch.zeros(1, dtype=torch.int32, device=dev)
        L.check(lib.hg_kmeans_fit(L.ptr(x), n, d, int(n_clusters), int(n_init), int(max_iter), float(tol),
                                  int(seed) & 0xFFFFFFFFFFFFFFFF, L.ptr(labels), L.ptr(centers), L.ptr(inertia),
                                  L.ptr(n_iter), L.ptr(ws), nbytes.value, L.stream_ptr(dev)))
        out = labels.cpu().numpy().astype(np.int64)
        LAST_N_ITER = int(n_iter.item())
    return out, centers, float(inertia.item())


def kmeans_stepwise(features: torch.Tensor, n_clusters: int, n_init: int = 10, max_iter: int = 300, tol: float = 1e-4,
                    seed: int = 42, device=None) -> Tuple[np.ndarray, torch.Tensor, float]:
    """The same estimator as a Python loop over the building blocks (torch RNG; one read-back per iteration)."""
    lib = L.load_library()
    dev = L.require_cuda(device if device is not None else (features.device if features.is_cuda else None))
    x = L.f32c(features, dev)
    n, d = x.shape
    if n < n_clusters:
        raise ValueError(f"n_samples={n} should be >= n_clusters={n_clusters}.")     # scikit-learn's message
    with torch.cuda.device(dev):
        st = L.stream_ptr(dev)
        tol_abs = float(x.var(dim=0, unbiased=False).mean()) * tol
        nbytes = C.c_size_t(0)
        L.check(lib.hg_kmeans_workspace_bytes(d, n_clusters, C.byref(nbytes)))
        ws = L.Workspace.get(dev, nbytes.value, "kmeans")
        labels = torch.empty(n, dtype=torch.int32, device=dev)
        prev = torch.empty(n, dtype=torch.int32, device=dev)
        d2 = torch.empty(n, dtype=torch.float32, device=dev)
        scal = torch.zeros(2, dtype=torch.float64, device=dev)          # [shift^2, inertia]
        best = None
        for run in range(n_init):
            gen = torch.Generator(device=dev)
            gen.manual_seed(seed + 1000003 * run)
            centers = _kmeanspp(lib, x, n_clusters, gen, st)
            prev.fill_(-1)
            for _ in range(max_iter):
                L.check(lib.hg_kmeans_assign(L.ptr(x), n, d, L.ptr(centers), n_clusters, L.ptr(labels), L.ptr(d2), st))
                L.check(lib.hg_kmeans_update(L.ptr(x), n, d, L.ptr(labels), n_clusters, L.ptr(centers),
                                             scal.data_ptr(), None, L.ptr(ws), nbytes.value, st))
                # one read-back per iteration: centre shift (scikit-learn's tol test) and whether any label
                # moved (its strict-convergence test, which is what ends runs on nearly constant data)
                shift, moved = torch.stack([scal[0], (labels != prev).sum().double()]).tolist()
                if shift <= tol_abs or moved == 0:
                    break
                prev.copy_(labels)
            L.check(lib.hg_kmeans_assign(L.ptr(x), n, d, L.ptr(centers), n_clusters, L.ptr(labels), L.ptr(d2), st))
            L.check(lib.hg_kmeans_inertia(L.ptr(d2), n, scal.data_ptr() + 8, st))
            inertia = float(scal[1])
            if best is None or inertia < best[2]:
                best = (labels.clone(), centers.clone(), inertia)
    return best[0].cpu().numpy().astype(np.int64), best[1], best[2]


class KMeans:
    """The slice of scikit-learn's estimator API the reference uses -- KMeans(...).fit_predict(X) -- plus
    fit / predict and the fitted attributes (labels_, cluster_centers_, inertia_, n_iter_)."""

    def __init__(self, n_clusters: int = 8, random_state: int = 42, n_init: int = 10, max_iter: int = 300,
                 tol: float = 1e-4):
        self.n_clusters, self.random_state, self.n_init, self.max_iter, self.tol = \
            n_clusters, random_state, n_init, max_iter, tol

    @staticmethod
    def _tensor(X) -> torch.Tensor:
        return X if isinstance(X, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(X, dtype=np.float32))

    def fit(self, X) -> "KMeans":
        self.labels_, self.cluster_centers_, self.inertia_ = kmeans(
            self._tensor(X), self.n_clusters, self.n_init, self.max_iter, self.tol, seed=int(self.random_state or 0))
        self.n_iter_ = LAST_N_ITER
        return self

    def fit_predict(self, X) -> np.ndarray:
        return self.fit(X).labels_

    def predict(self, X) -> np.ndarray:
        """nearest fitted centre of every row (hg_kmeans_assign)"""
        lib = L.load_library()
        cen = self.cluster_centers_
        x = L.f32c(self._tensor(X), cen.device)
        n, d = x.shape
        labels = torch.empty(n, dtype=torch.int32, device=cen.device)
        d2 = torch.empty(n, dtype=torch.float32, device=cen.device)
        with torch.cuda.device(cen.device):
            L.check(lib.hg_kmeans_assign(L.ptr(x), n, d, L.ptr(cen), cen.shape[0], L.ptr(labels), L.ptr(d2),
                                         L.stream_ptr(cen.device)))
        return labels.cpu().numpy().astype(np.int64)
