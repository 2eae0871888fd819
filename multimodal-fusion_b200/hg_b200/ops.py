"""Torch-tensor front end of the C ABI: allocates outputs / workspaces with torch and
passes raw pointers to libhg_b200.so.  One function per kernel family."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib as L


def _size(fn, *args) -> int:
    out = C.c_size_t(0)
    L.check(fn(*args, C.byref(out)))
    return int(out.value)


# --------------------------------------------------------------------------- kNN
def knn_feature(features: torch.Tensor, k: int, metric: str = "euclidean",
                positions: Optional[torch.Tensor] = None, lambda_h: float = 1.0, lambda_g: float = 0.0,
                rows: Optional[Tuple[int, int]] = None, exclude_self: bool = True,
                precision: str = "tc", device=None, return_cos: bool = False):
    """Fused similarity + top-k (A1/A2/A4).  Returns (idx int32 [R,k], val float32 [R,k]);
    val = squared distance / Gaussian K / cosine depending on `metric`.  return_cos=True adds
    cos(f_row, f_idx) [R,k], the by-product of the exact re-rank that edge_weights_from_knn reads."""
    lib = L.load_library()
    dev = L.require_cuda(device if device is not None else (features.device if features.is_cuda else None))
    x = L.f32c(features, dev)
    n, d = x.shape
    pos = L.f32c(positions, dev) if positions is not None else None
    r0, r1 = (0, n) if rows is None else rows
    if n <= 1:
        raise ValueError(f"Number of nodes must be greater than 1, got N={n}.")
    out_idx = torch.empty((max(r1 - r0, 0), k), dtype=torch.int32, device=dev)
    out_val = torch.empty((max(r1 - r0, 0), k), dtype=torch.float32, device=dev)
    prec = {"tc": 0, "fp32": 1}[precision]
    with torch.cuda.device(dev):
        ws = None
        nbytes = 0
        if prec == 0:
            nbytes = _size(lib.hg_knn_feature_workspace_bytes, n, d, k, r0, r1)
            ws = L.Workspace.get(dev, nbytes)
        out_cos = torch.empty_like(out_val) if return_cos else None
        L.check(lib.hg_knn_feature_cos(
            L.ptr(x), n, d, L.ptr(pos), 0 if pos is None else pos.shape[1], L.METRIC[metric],
            float(lambda_h), float(lambda_g), k, r0, r1, int(exclude_self), prec,
            L.ptr(out_idx), L.ptr(out_val), L.ptr(out_cos), L.ptr(ws), nbytes, L.stream_ptr(dev)))
    return (out_idx, out_val, out_cos) if return_cos else (out_idx, out_val)


def knn_spatial(positions: torch.Tensor, k: int, rows: Optional[Tuple[int, int]] = None,
                device=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """A3: k smallest squared spatial distances, ties by (d2, index).  (idx int32, d2 float32)."""
    lib = L.load_library()
    dev = L.require_cuda(device if device is not None else (positions.device if positions.is_cuda else None))
    p = L.f32c(positions, dev)
    n, pd = p.shape
    r0, r1 = (0, n) if rows is None else rows
    out_idx = torch.empty((r1 - r0, k), dtype=torch.int32, device=dev)
    out_d2 = torch.empty((r1 - r0, k), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        nbytes = _size(lib.hg_knn_spatial_workspace_bytes, n, k, r0, r1)
        ws = L.Workspace.get(dev, nbytes)
        L.check(lib.hg_knn_spatial(L.ptr(p), n, pd, k, r0, r1, L.ptr(out_idx), L.ptr(out_d2),
                                   L.ptr(ws), nbytes, L.stream_ptr(dev)))
    return out_idx, out_d2


# ------------------------------------------------------------------- dense kernels
# Rows from which square expansion-form kernels (A1 / A2 as the reference returns them, similarity_kernel.py:43-52,
# 79-84, 122) go through the tcgen05 kernel's dense-output mode; below that the FP32 SIMT tiles are as fast as the
# operand split + launch.  The mode is 6-7x faster than the SIMT tiles (N = 16,384 x 1024: 11.6 -> 1.8 ms) and holds
# the 1e-5 bar of the scope contract: hi*hi accumulates in one TMEM accumulator, the three small products
# (hi*lo, lo*hi, lo*lo) in a second one, added in FP32 in the epilogue, and the diagonal is exact -- measured maximum
# over ALL entries of a 16,384^2 output against float64: 4.4e-6 relative at D = 1024, 6.0e-6 at D = 1536 (the SIMT
# tiles: 4.8e-6 / 6.6e-6; profiles/r02_dense_tc_accuracy.jsonl).  HG_B200_DENSE_TC_MIN_N=0 turns the routing off.
DENSE_TC_MIN_N = int(os.environ.get("HG_B200_DENSE_TC_MIN_N", "4096"))


def _dense_on_tensor_cores(xb, pb, n: int, d: int, lambda_h: float, form: int) -> bool:
    return (DENSE_TC_MIN_N > 0 and xb is None and pb is None and form == L.FORM_EXPANSION and lambda_h > 0
            and n >= DENSE_TC_MIN_N and d >= 64)


def similarity_dense(a: torch.Tensor, b: Optional[torch.Tensor] = None,
                     pos_a: Optional[torch.Tensor] = None, pos_b: Optional[torch.Tensor] = None,
                     lambda_h: float = 1.0, lambda_g: float = 0.0, form: int = L.FORM_EXPANSION,
                     device=None) -> torch.Tensor:
    """A1/A2/A7: dense exp(-lambda_h d2_h - lambda_g d2_g), [M,N] float32."""
    lib = L.load_library()
    dev = L.require_cuda(device if device is not None else (a.device if a.is_cuda else None))
    xa = L.f32c(a, dev)
    xb = L.f32c(b, dev) if b is not None else None
    pa = L.f32c(pos_a, dev) if pos_a is not None else None
    pb = L.f32c(pos_b, dev) if pos_b is not None else None
    m, d = xa.shape
    n = m if xb is None else xb.shape[0]
    out = torch.empty((m, n), dtype=torch.float32, device=dev)
    if m == 0 or n == 0:
        return out
    if _dense_on_tensor_cores(xb, None if (pos_b is None or pos_b is pos_a) else pb, m, d, lambda_h, form):
        with torch.cuda.device(dev):
            nbytes = _size(lib.hg_similarity_dense_tc_workspace_bytes, m, d)
            ws = L.Workspace.get(dev, nbytes, "dense_tc")
            L.check(lib.hg_similarity_dense_tc(L.ptr(xa), m, d, L.ptr(pa), 0 if pa is None else pa.shape[1],
                                               float(lambda_h), float(lambda_g if pa is not None else 0.0), L.ptr(out), n,
                                               L.ptr(ws), nbytes, L.stream_ptr(dev)))
        return out
    with torch.cuda.device(dev):
        L.check(lib.hg_similarity_dense(L.ptr(xa), m, L.ptr(xb), n, d, L.ptr(pa), L.ptr(pb),
                                        0 if pa is None else pa.shape[1], float(lambda_h), float(lambda_g),
                                        form, L.ptr(out), n, L.stream_ptr(dev)))
    return out


def similarity_stats(mat: torch.Tensor, exclude_diagonal: bool = False) -> dict:
    """A9: mean / std / min / max / lower-median of a CUDA float32 matrix (one D2H of 5 doubles)."""
    lib = L.load_library()
    dev = L.require_cuda(mat.device if mat.is_cuda else None)
    m = L.f32c(mat, dev)
    count = m.numel()
    out = torch.empty(5, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        nbytes = _size(lib.hg_stats_workspace_bytes, count)
        ws = L.Workspace.get(dev, nbytes, "stats")
        L.check(lib.hg_similarity_stats(L.ptr(m), count, m.shape[0] if exclude_diagonal else 0, L.ptr(out),
                                        L.ptr(ws), nbytes, L.stream_ptr(dev)))
    v = out.cpu().tolist()
    return {"mean": v[0], "std": v[1], "min": v[2], "max": v[3], "median": v[4]}


def similarity_stats_streaming(features: torch.Tensor, positions: Optional[torch.Tensor] = None,
                               lambda_h: float = 1.0, lambda_g: float = 0.0, device=None) -> dict:
    """A9 for large N: statistics of K = K_h * K_g over all N^2 entries without storing the matrix."""
    lib = L.load_library()
    dev = L.require_cuda(device if device is not None else (features.device if features.is_cuda else None))
    x = L.f32c(features, dev)
    pos = L.f32c(positions, dev) if positions is not None else None
    n, d = x.shape
    out = torch.empty(5, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        nbytes = _size(lib.hg_similarity_stats_streaming_workspace_bytes, n, d)
        ws = L.Workspace.get(dev, nbytes)
        L.check(lib.hg_similarity_stats_streaming(L.ptr(x), n, d, L.ptr(pos), 0 if pos is None else pos.shape[1],
                                                  float(lambda_h), float(lambda_g), L.ptr(out), L.ptr(ws), nbytes,
                                                  L.stream_ptr(dev)))
    v = out.cpu().tolist()
    return {"mean": v[0], "std": v[1], "min": v[2], "max": v[3], "median": v[4]}


# ------------------------------------------------------------------ edge assembly
def build_edges(knn_lists: Sequence[torch.Tensor], n: int, labels: Optional[torch.Tensor] = None,
                num_hyperedges: int = 0, device=None) -> torch.Tensor:
    """A5: kNN pairs (+ intra-cluster cliques) -> unique (min,max) edge_index int64 [2,E], sorted."""
    lib = L.load_library()
    dev = L.require_cuda(device if device is not None else (knn_lists[0].device if knn_lists else None))
    lists = [t.to(device=dev, dtype=torch.int32).contiguous() for t in knn_lists]
    for t in lists:
        if t.shape[0] != n:
            raise ValueError("every neighbour list must have n rows")
    ks = [int(t.shape[1]) for t in lists]
    lab = None
    pairs = C.c_int64(0)
    with torch.cuda.device(dev):
        st = L.stream_ptr(dev)
        if labels is not None and num_hyperedges > 0:
            lab = torch.as_tensor(labels).to(device=dev, dtype=torch.int32).contiguous()
            if lab.numel() != n:
                raise ValueError("labels must have n entries")
            L.check(lib.hg_clique_pairs(L.ptr(lab), n, num_hyperedges, C.byref(pairs), st))
        else:
            num_hyperedges = 0
        slots = sum(n * k for k in ks)
        nbytes = _size(lib.hg_edges_workspace_bytes, n, slots, num_hyperedges, pairs.value)
        ws = L.Workspace.get(dev, nbytes, "edges")
        arr_p = (C.c_void_p * max(len(lists), 1))(*[t.data_ptr() for t in lists])
        arr_k = (C.c_int * max(len(lists), 1))(*ks)
        e = C.c_int64(0)
        L.check(lib.hg_edges_count(arr_p, arr_k, len(lists), n, L.ptr(lab), num_hyperedges, pairs.value,
                                   C.byref(e), L.ptr(ws), nbytes, st))
        edge_index = torch.empty((2, e.value), dtype=torch.int64, device=dev)
        L.check(lib.hg_edges_fill(e.value, L.ptr(edge_index), L.ptr(ws), nbytes, st))
    return edge_index


def edge_weights_cosine(features: torch.Tensor, edge_index: torch.Tensor) -> torch.Tensor:
    """A6: max(0, cos(f_i, f_j)) per edge, float32 [E]."""
    lib = L.load_library()
    dev = L.require_cuda(edge_index.device if edge_index.is_cuda else None)
    x = L.f32c(features, dev)
    ei = edge_index.to(device=dev, dtype=torch.int64).contiguous()
    e = ei.shape[1]
    w = torch.empty(e, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        L.check(lib.hg_edge_weights_cosine(L.ptr(x), x.shape[0], x.shape[1], L.ptr(ei), e, L.ptr(w), L.stream_ptr(dev)))
    return w


def edge_weights_from_knn(features: torch.Tensor, knn_idx: torch.Tensor, knn_cos: torch.Tensor,
                          edge_index: torch.Tensor) -> torch.Tensor:
    """A6 without gathering feature rows: w_e looked up in the full-range kNN lists (knn_idx / knn_cos
    [n, k] from knn_feature(..., return_cos=True)); edges in neither endpoint's list are computed from
    the rows.  Bit-identical to edge_weights_cosine."""
    lib = L.load_library()
    dev = L.require_cuda(edge_index.device if edge_index.is_cuda else None)
    x = L.f32c(features, dev)
    ei = edge_index.to(device=dev, dtype=torch.int64).contiguous()
    idx = knn_idx.to(device=dev, dtype=torch.int32).contiguous()
    cs = L.f32c(knn_cos, dev)
    if idx.shape != cs.shape or idx.shape[0] != x.shape[0]:
        raise ValueError("knn_idx / knn_cos must be [n, k] lists of the full row range")
    e = ei.shape[1]
    w = torch.empty(e, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        L.check(lib.hg_edge_weights_from_knn(L.ptr(x), x.shape[0], x.shape[1], L.ptr(idx), L.ptr(cs), idx.shape[1],
                                             L.ptr(ei), e, L.ptr(w), L.stream_ptr(dev)))
    return w


def csr_from_coo(edge_index: torch.Tensor, n: int):
    """CSR incidence: (rowptr [n+1], col [2E], eid [2E]) int64."""
    lib = L.load_library()
    dev = L.require_cuda(edge_index.device if edge_index.is_cuda else None)
    ei = edge_index.to(device=dev, dtype=torch.int64).contiguous()
    e = ei.shape[1]
    rowptr = torch.empty(n + 1, dtype=torch.int64, device=dev)
    col = torch.empty(2 * e, dtype=torch.int64, device=dev)
    eid = torch.empty(2 * e, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        nbytes = _size(lib.hg_csr_workspace_bytes, n, e)
        ws = L.Workspace.get(dev, nbytes, "csr")
        L.check(lib.hg_csr_from_coo(L.ptr(ei), e, n, L.ptr(rowptr), L.ptr(col), L.ptr(eid), L.ptr(ws), nbytes,
                                    L.stream_ptr(dev)))
    return rowptr, col, eid


def hypergraph_propagate(x: torch.Tensor, csr_rowptr: torch.Tensor, csr_col: torch.Tensor,
                         return_hyperedge_features: bool = False):
    """Row (f)3: out = D^-1 H B^-1 H^T x over the stored pair list read as (node, hyperedge) pairs
    -- the sparse part of torch_geometric's HypergraphConv as the reference's consumer calls it
    (cust_omics.py:58-63, 100-102).  x [n, F]; CSR from csr_from_coo / the build's return_csr."""
    lib = L.load_library()
    dev = L.require_cuda(csr_rowptr.device if csr_rowptr.is_cuda else None)
    xx = L.f32c(x, dev)
    n, f = xx.shape
    rp = csr_rowptr.to(device=dev, dtype=torch.int64).contiguous()
    cl = csr_col.to(device=dev, dtype=torch.int64).contiguous()
    if rp.numel() != n + 1:
        raise ValueError("csr_rowptr must have n + 1 entries")
    m = torch.empty_like(xx)
    out = torch.empty_like(xx)
    with torch.cuda.device(dev):
        L.check(lib.hg_hypergraph_propagate(L.ptr(xx), n, f, L.ptr(rp), L.ptr(cl), L.ptr(m), L.ptr(out), L.stream_ptr(dev)))
    return (out, m) if return_hyperedge_features else out


def aggregate_super_patches(features: torch.Tensor, positions: Optional[torch.Tensor], labels, num_clusters: int,
                            lambda_h: float = 1.0, lambda_g: float = 0.0, dense_k: Optional[torch.Tensor] = None,
                            want_intra: bool = True, return_order: bool = False, device=None) -> dict:
    """Row (f)4: mean feature / position per cluster label and the intra-cluster similarity means
    (reference preprocess_hypergraph.py:157-184) in one stream-ordered call.  Everything in the
    returned dict stays on the device; `summary` = [avg_intra, clusters with m > 1, first empty
    cluster or -1, rows with a label in range] is the only thing a caller needs to read back."""
    lib = L.load_library()
    dev = L.require_cuda(device if device is not None else (features.device if features.is_cuda else None))
    x = L.f32c(features, dev)
    n, d = x.shape
    pos = L.f32c(positions, dev) if positions is not None else None
    pd = 0 if pos is None else pos.shape[1]
    lab = torch.as_tensor(np.asarray(labels) if not torch.is_tensor(labels) else labels).to(device=dev, dtype=torch.int64).contiguous()
    if lab.numel() != n:
        raise ValueError("labels must have one entry per row")
    kd = L.f32c(dense_k, dev) if dense_k is not None else None
    if kd is not None and (kd.dim() != 2 or kd.shape[0] != n or kd.shape[1] != n):
        raise ValueError("dense_k must be [N, N]")
    s = int(num_clusters)
    out = {
        "features": torch.empty((s, d), dtype=torch.float32, device=dev),
        "positions": torch.empty((s, pd), dtype=torch.float32, device=dev) if pos is not None else None,
        "counts": torch.empty(s, dtype=torch.int64, device=dev),
        "intra": torch.empty(s, dtype=torch.float32, device=dev),
        "summary": torch.empty(4, dtype=torch.float32, device=dev),
        "order": torch.empty(n, dtype=torch.int32, device=dev) if return_order else None,
    }
    with torch.cuda.device(dev):
        nbytes = _size(lib.hg_aggregate_workspace_bytes, n, d, s)
        ws = L.Workspace.get(dev, nbytes, "agg")
        L.check(lib.hg_aggregate_super_patches(
            L.ptr(x), n, d, L.ptr(pos), pd, L.ptr(lab), s, float(lambda_h), float(lambda_g),
            L.ptr(kd), 0 if kd is None else kd.stride(0), int(want_intra),
            L.ptr(out["features"]), L.ptr(out["positions"]), L.ptr(out["counts"]), L.ptr(out["intra"]),
            L.ptr(out["summary"]), L.ptr(out["order"]), L.ptr(ws), nbytes, L.stream_ptr(dev)))
    return out


def threshold_edges(K: torch.Tensor, threshold: float):
    """A8: every ordered (i, j) with not (K[i,j] < threshold), row-major; (edge_index, weights)."""
    lib = L.load_library()
    dev = L.require_cuda(K.device if K.is_cuda else None)
    Kc = L.f32c(K, dev)
    n = Kc.shape[0]
    with torch.cuda.device(dev):
        st = L.stream_ptr(dev)
        nbytes = _size(lib.hg_threshold_workspace_bytes, n)
        ws = L.Workspace.get(dev, nbytes, "thr")
        e = C.c_int64(0)
        L.check(lib.hg_threshold_edges_count(L.ptr(Kc), n, float(threshold), C.byref(e), L.ptr(ws), nbytes, st))
        ei = torch.empty((2, e.value), dtype=torch.int64, device=dev)
        w = torch.empty(e.value, dtype=torch.float32, device=dev)
        L.check(lib.hg_threshold_edges_fill(L.ptr(Kc), n, float(threshold), e.value, L.ptr(ei), L.ptr(w),
                                            L.ptr(ws), nbytes, st))
    return ei, w


# ------------------------------------------------------------------ one call per slide
def _pack_build(b: dict, ne: int, return_csr: bool, spatial_k: int) -> dict:
    """Views of the capacity-sized output buffers for an edge count that is now known."""
    out = {"knn_idx": b["knn_idx"], "knn_val": b["knn_val"]}
    if spatial_k:
        out["spatial_idx"], out["spatial_d2"] = b["sidx"], b["sd2"]
    out["edge_index"] = b["ei"][: 2 * ne].view(2, ne)
    out["edge_weights"] = b["w"][:ne]
    if return_csr:
        out["csr_rowptr"], out["csr_col"], out["csr_eid"] = b["rowptr"], b["col"][: 2 * ne], b["eid"][: 2 * ne]
    return out


class PendingBuild:
    """A slide whose build has been enqueued with hg_build_knn_hypergraph_enqueue: nothing has
    synchronised yet.  `result()` waits for the slide's stream event, reads the edge count and
    returns the same dict as `build_knn_hypergraph` (views of the capacity-sized buffers)."""

    def __init__(self, bufs: dict, e_host: torch.Tensor, event, return_csr: bool, spatial_k: int, keep):
        self._b, self._e_host, self.event = bufs, e_host, event
        self._csr, self._sk, self._keep, self._out = return_csr, spatial_k, keep, None

    def result(self) -> dict:
        if self._out is None:
            self.event.synchronize()
            self._out = _pack_build(self._b, int(self._e_host.item()), self._csr, self._sk)
            self._keep = None                    # inputs were only kept alive for the enqueued kernels
        return self._out


def build_knn_hypergraph(features: torch.Tensor, k: int, metric: str = "euclidean",
                         positions: Optional[torch.Tensor] = None, lambda_h: float = 1.0, lambda_g: float = 0.0,
                         spatial_k: int = 0, labels: Optional[torch.Tensor] = None, num_hyperedges: int = 0,
                         return_csr: bool = False, device=None, enqueue_only: bool = False):
    """hg_build_knn_hypergraph: kNN lists, edge_index, weights (and CSR) of one slide with a single
    host call and a single stream synchronisation.  enqueue_only=True uses the stream-ordered entry
    point instead (no synchronisation at all) and returns a PendingBuild."""
    lib = L.load_library()
    dev = L.require_cuda(device if device is not None else (features.device if features.is_cuda else None))
    x = L.f32c(features, dev)
    n, d = x.shape
    if n <= 1:
        raise ValueError(f"Number of nodes must be greater than 1, got N={n}.")
    pos = L.f32c(positions, dev) if positions is not None else None
    if pos is None:
        spatial_k = 0
    with torch.cuda.device(dev):
        st = L.stream_ptr(dev)
        lab, pairs = None, C.c_int64(0)
        if labels is not None and num_hyperedges > 0:
            lab = torch.as_tensor(labels).to(device=dev, dtype=torch.int32).contiguous()
            if lab.numel() != n:
                raise ValueError("labels must have n entries")
            L.check(lib.hg_clique_pairs(L.ptr(lab), n, num_hyperedges, C.byref(pairs), st))
        else:
            num_hyperedges = 0
        max_e = n * (k + spatial_k) + pairs.value
        b = {"knn_idx": torch.empty((n, k), dtype=torch.int32, device=dev),
             "knn_val": torch.empty((n, k), dtype=torch.float32, device=dev),
             "sidx": torch.empty((n, spatial_k), dtype=torch.int32, device=dev) if spatial_k else None,
             "sd2": torch.empty((n, spatial_k), dtype=torch.float32, device=dev) if spatial_k else None,
             "ei": torch.empty(2 * max_e, dtype=torch.int64, device=dev),
             "w": torch.empty(max_e, dtype=torch.float32, device=dev),
             "rowptr": torch.empty(n + 1, dtype=torch.int64, device=dev) if return_csr else None,
             "col": torch.empty(2 * max_e, dtype=torch.int64, device=dev) if return_csr else None,
             "eid": torch.empty(2 * max_e, dtype=torch.int64, device=dev) if return_csr else None}
        nbytes = _size(lib.hg_build_workspace_bytes, n, d, k, spatial_k, num_hyperedges, pairs.value, max_e,
                       int(return_csr))
        ws = L.Workspace.get(dev, nbytes)
        args = (L.ptr(x), n, d, L.ptr(pos), 0 if pos is None else pos.shape[1], L.METRIC[metric], float(lambda_h),
                float(lambda_g), k, spatial_k, L.ptr(lab), num_hyperedges, pairs.value,
                L.ptr(b["knn_idx"]), L.ptr(b["knn_val"]), L.ptr(b["sidx"]), L.ptr(b["sd2"]), max_e, L.ptr(b["ei"]),
                L.ptr(b["w"]), L.ptr(b["rowptr"]), L.ptr(b["col"]), L.ptr(b["eid"]))
        if enqueue_only:
            e_dev = torch.empty(1, dtype=torch.int64, device=dev)
            e_host = torch.empty(1, dtype=torch.int64, pin_memory=True)
            L.check(lib.hg_build_knn_hypergraph_enqueue(*args, L.ptr(e_dev), L.ptr(ws), nbytes, st))
            e_host.copy_(e_dev, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(dev))
            return PendingBuild(b, e_host, ev, return_csr, spatial_k, keep=(x, pos, lab, e_dev))
        e = C.c_int64(0)
        L.check(lib.hg_build_knn_hypergraph(*args, C.byref(e), L.ptr(ws), nbytes, st))
    return _pack_build(b, e.value, return_csr, spatial_k)
