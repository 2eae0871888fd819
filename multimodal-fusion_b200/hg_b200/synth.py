"""Deterministic synthetic slides.

Patch embeddings are a mixture of 32 Gaussian clusters in D dimensions, half of
the slides passed through abs() to mimic post-activation embeddings, then either
L2-normalised rows (cosine configs) or scaled so that mean |h|^2 = 1 (Gaussian
configs, so lambda_h = 1 gives lambda*d^2 = O(1) instead of the underflow the
reference defaults produce on raw embeddings).  Positions are unique integer
grid cells drawn without replacement from a ceil(1.3 sqrt(N))^2 lattice.

numpy's PCG64 stream is stable across numpy versions, so (seed, n, d) fully
names a slide on any machine -- tests, bench and the oracle all regenerate the
same host arrays from it.
"""
from __future__ import annotations

import numpy as np

N_CLUSTERS = 32


def make_features(n: int, d: int, seed: int, mode: str = "gaussian") -> np.ndarray:
    """mode 'cosine': unit rows; 'gaussian': mean squared norm 1; 'raw': neither."""
    rng = np.random.default_rng(seed)
    centres = rng.standard_normal((N_CLUSTERS, d), dtype=np.float32)
    assign = rng.integers(0, N_CLUSTERS, size=n)
    x = centres[assign]
    # chunked so a 100k x 1536 slide never holds two float64 temporaries
    for s in range(0, n, 16384):
        x[s:s + 16384] += 0.5 * rng.standard_normal((min(16384, n - s), d), dtype=np.float32)
    if seed % 2 == 0:
        np.abs(x, out=x)
    if mode == "cosine":
        x /= np.maximum(np.linalg.norm(x, axis=1, keepdims=True), 1e-12)
    elif mode == "gaussian":
        x *= np.float32(1.0 / np.sqrt(np.mean(np.sum(x.astype(np.float64) ** 2, axis=1))))
    elif mode != "raw":
        raise ValueError(f"unknown mode {mode!r}")
    return np.ascontiguousarray(x, dtype=np.float32)


def make_positions(n: int, seed: int, stride: float = 1.0) -> np.ndarray:
    """Unique (x, y) grid cells, ~60% occupancy, float32 in units of `stride`."""
    rng = np.random.default_rng(seed + 7919)
    side = int(np.ceil(1.3 * np.sqrt(n)))
    cells = rng.choice(side * side, size=n, replace=False)
    pos = np.stack([cells % side, cells // side], axis=1).astype(np.float32)
    return np.ascontiguousarray(pos * np.float32(stride))


def make_slide(n: int, d: int, seed: int, mode: str = "gaussian"):
    return make_features(n, d, seed, mode), make_positions(n, seed)


def slide_sizes(num_slides: int, seed: int = 1238, lo: int = 2048, hi: int = 8192) -> np.ndarray:
    """Config C4: N_s ~ U{lo..hi}."""
    return np.random.default_rng(seed).integers(lo, hi + 1, size=num_slides)
