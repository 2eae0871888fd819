"""Stream-pipelined per-slide builds for dataset loops (the reference's `process_dataset`
loop, preprocess_hypergraph.py:651-670, is strictly sequential: load -> build -> save).

Streams per device:
  * copy-in  : H2D of the next slides from pinned host memory,
  * `lanes` compute streams: slide i is built on lane i % lanes with the stream-ordered entry
    point (hg_build_knn_hypergraph_enqueue) -- no host synchronisation, so the host runs ahead and
    the small kernels of one slide (re-rank, radix passes, CSR) fill the SMs next to another
    slide's tensor-core pass.  This is what batches of small slides (BASELINE config C4) need;
  * copy-out : D2H of a slide's outputs into pinned host buffers once its edge count is known.
PCIe is full duplex, so a steady-state step costs max(H2D, build, D2H) instead of their sum.
Every slide's H2D and D2H still happen -- they are overlapped, not skipped.
"""
from __future__ import annotations

import os
from collections import deque
from typing import Callable, Deque, Dict, Iterable, Iterator, List, Optional, Tuple

import torch

from . import _lib as L
from . import ops

_OUT_KEYS = ("edge_index", "edge_weights", "csr_rowptr", "csr_col", "csr_eid", "knn_idx", "knn_val",
             "spatial_idx", "spatial_d2")
_METRIC_OF = {"gaussian": "gaussian", "cosine": "cosine", "euclidean": "euclidean"}


def _enqueue(x, p, dev, *, k=16, kernel="gaussian", lambda_h=1.0, lambda_g=0.0, spatial_k=0, labels=None,
             num_hyperedges=0, return_csr=False) -> ops.PendingBuild:
    return ops.build_knn_hypergraph(x, k, _METRIC_OF[kernel], positions=p, lambda_h=lambda_h,
                                    lambda_g=lambda_g if kernel == "gaussian" else 0.0, spatial_k=spatial_k,
                                    labels=labels, num_hyperedges=num_hyperedges, return_csr=return_csr, device=dev,
                                    enqueue_only=True)


class _Eager:
    """Adapter for a user-supplied synchronous build function."""

    def __init__(self, out):
        self._out = out

    def result(self):
        return self._out


class SlidePipeline:
    """submit(features[, positions]) enqueues one slide; results() yields dicts of pinned host
    tensors in submission order.  `keys` selects which outputs are copied back.  The pinned result
    buffers rotate over two slots: consume (or copy) a result before two more slides finish.
    `lanes` compute streams build consecutive slides concurrently; `depth` slides may be in
    flight on the device before the host waits for the oldest one."""

    def __init__(self, device=None, keys: Tuple[str, ...] = ("edge_index", "edge_weights", "csr_rowptr", "csr_col",
                                                              "csr_eid"),
                 build_fn: Optional[Callable] = None, lanes: Optional[int] = None, depth: Optional[int] = None,
                 **build_kwargs):
        if lanes is None:                      # HG_B200_PIPE_LANES: tuning override of the default
            lanes = int(os.environ.get("HG_B200_PIPE_LANES", "2"))
        self.dev = L.require_cuda(device)
        self.keys = keys
        self.kw = build_kwargs
        self.build_fn = build_fn
        self.lanes = max(1, int(lanes))
        self.depth = max(1, int(depth)) if depth is not None else 2 * self.lanes
        self.s_in = torch.cuda.Stream(self.dev)
        self.s_out = torch.cuda.Stream(self.dev)
        self.s_cmp = [torch.cuda.Stream(self.dev) for _ in range(self.lanes)]
        self._inflight: Deque = deque()        # (pending build, its compute stream, built event)
        self._done: Deque = deque()            # (host_dict, done_event)
        self._host_pool: Dict = {}
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self._submitted = 0
        self._collected = 0

    # -- H2D on the copy-in stream, build enqueued on the slide's lane ---------------------------
    def submit(self, features: torch.Tensor, positions: Optional[torch.Tensor] = None):
        with torch.cuda.stream(self.s_in):
            x = features.to(self.dev, dtype=torch.float32, non_blocking=True)
            p = positions.to(self.dev, dtype=torch.float32, non_blocking=True) if positions is not None else None
            ev = torch.cuda.Event()
            ev.record(self.s_in)
        self.h2d_bytes += features.numel() * 4 + (positions.numel() * 4 if positions is not None else 0)
        s = self.s_cmp[self._submitted % self.lanes]
        self._submitted += 1
        s.wait_event(ev)
        with torch.cuda.stream(s):
            x.record_stream(s)
            if p is not None:
                p.record_stream(s)
            if self.build_fn is None:
                pend = _enqueue(x, p, self.dev, **self.kw)
            else:
                pend = _Eager(self.build_fn(x, p, device=self.dev, **self.kw))
            built = torch.cuda.Event()
            built.record(s)
        self._inflight.append((pend, s, built))
        while len(self._inflight) > self.depth:
            self._collect_one()

    # -- D2H on the copy-out stream once the slide's edge count has arrived ------------------------
    def _collect_one(self):
        pend, s, built = self._inflight.popleft()
        out = pend.result()                    # waits for this slide only; later slides keep the GPU busy
        host = {}
        self.s_out.wait_event(built)
        with torch.cuda.stream(self.s_out):
            for name in self.keys:
                t = out.get(name)
                if t is None:
                    continue
                t.record_stream(self.s_out)
                buf = self._host_buffer(name, t)
                buf.copy_(t, non_blocking=True)
                host[name] = buf
                self.d2h_bytes += t.numel() * t.element_size()
            done = torch.cuda.Event()
            done.record(self.s_out)
        self._done.append((host, done))
        self._collected += 1

    def _host_buffer(self, name: str, t: torch.Tensor) -> torch.Tensor:
        # two rotating pinned buffers per output so slide i's copy never lands on slide i-1's result
        slot = (name, self._collected & 1)
        need = t.numel()
        buf = self._host_pool.get(slot)
        if buf is None or buf.dtype != t.dtype or buf.numel() < need:
            buf = torch.empty(max(need, 1), dtype=t.dtype, pin_memory=True)
            self._host_pool[slot] = buf
        return buf[:need].view(t.shape)

    def ready(self) -> int:
        """Slides whose D2H has been enqueued and that results() would yield."""
        return len(self._done)

    def results(self, flush: bool = False, limit: Optional[int] = None) -> Iterator[dict]:
        """Yield finished slides in order (at most `limit`); with flush=True also collect what is
        still in flight -- then consume each result before asking for the next (two host slots)."""
        n = 0
        while limit is None or n < limit:
            if not self._done:
                if not (flush and self._inflight):
                    return
                self._collect_one()
            host, ev = self._done.popleft()
            ev.synchronize()
            n += 1
            yield host


def build_many(slides: Iterable, device=None, pipeline: Optional[SlidePipeline] = None, **build_kwargs) -> Iterator[dict]:
    """slides: iterable of `features` or `(features, positions)` host tensors (pinned memory makes
    the H2D asynchronous).  Yields one result dict (pinned host tensors) per slide, in order.
    Pass a long-lived `pipeline` to keep its streams, device staging and pinned buffers warm
    across calls (a fresh pipeline pays cudaMalloc / cudaHostAlloc on its first slides)."""
    pipe = pipeline if pipeline is not None else SlidePipeline(device, **build_kwargs)
    for item in slides:
        feats, pos = item if isinstance(item, (tuple, list)) else (item, None)
        pipe.submit(feats, pos)
        # the host pool rotates two slots per output: never let two collected results pile up
        while pipe.ready() >= 1:
            yield from pipe.results(limit=1)
    yield from pipe.results(flush=True)


LAST_BATCH_STATS: Dict = {}          # host-side timing of the most recent build_batch call (tuning aid)

# hg_slide (include/hg_b200.h): 14 fields of 8 bytes -> one int64 row per slide
_F_FEATS, _F_N, _F_POS, _F_KIDX, _F_KVAL, _F_SIDX, _F_SD2, _F_MAXE, _F_EI, _F_W, _F_ROWPTR, _F_COL, _F_EID, _F_NE = range(14)


class BatchResults:
    """Sequence of per-slide result dicts (device tensors, views of one output arena); the dicts are
    made on access.  Same keys as `knn_hyperedges`."""

    def __init__(self, arena, table, offs, counts, k, spatial_k, return_csr):
        self._arena, self._t, self._offs, self._ne = arena, table, offs, counts
        self._k, self._sk, self._csr = k, spatial_k, return_csr

    def __len__(self):
        return len(self._ne)

    def _view(self, off, count, dtype, shape):
        nbytes = count * (8 if dtype is torch.int64 else 4)
        return self._arena[off:off + nbytes].view(dtype).view(shape)

    def __getitem__(self, i: int) -> dict:
        if i < 0:
            i += len(self)
        n, ne, o = int(self._t[i, _F_N]), int(self._ne[i]), self._offs[i]
        out = {"knn_idx": self._view(o[0], n * self._k, torch.int32, (n, self._k)),
               "knn_val": self._view(o[1], n * self._k, torch.float32, (n, self._k))}
        if self._sk and self._t[i, _F_SIDX]:
            out["spatial_idx"] = self._view(o[2], n * self._sk, torch.int32, (n, self._sk))
            out["spatial_d2"] = self._view(o[3], n * self._sk, torch.float32, (n, self._sk))
        out["edge_index"] = self._view(o[4], 2 * ne, torch.int64, (2, ne))
        out["edge_weights"] = self._view(o[5], ne, torch.float32, (ne,))
        if self._csr:
            out["csr_rowptr"] = self._view(o[6], n + 1, torch.int64, (n + 1,))
            out["csr_col"] = self._view(o[7], 2 * ne, torch.int64, (2 * ne,))
            out["csr_eid"] = self._view(o[8], 2 * ne, torch.int64, (2 * ne,))
        return out

    def __iter__(self):
        return (self[i] for i in range(len(self)))


def build_batch(slides: Iterable, device=None, lanes: int = 4, threads: Optional[int] = None, *, k: int = 16,
                kernel: str = "gaussian", lambda_h: float = 1.0, lambda_g: float = 0.0, spatial_k: int = 0,
                return_csr: bool = False, grouped: Optional[bool] = None, group_rows: Optional[int] = None) -> BatchResults:
    """Device-resident batch of slides (`features` or `(features, positions)` CUDA tensors, same D):
    ONE C call enqueues every build stream-ordered with no synchronisation; outputs live in one arena.
    Returns after the edge counts have arrived (one small D2H).  The calling stream is ordered before
    and after the batch.

    grouped (default on, HG_B200_BATCH_GROUPED=0 turns it off; needs k <= 32): hg_build_batch_grouped_enqueue --
    the feature kNN of up to `group_rows` (HG_B200_GROUP_ROWS, default 131,072) operand rows of consecutive
    slides is ONE persistent launch of the fused kernel, the back halves run on `lanes` streams behind it.
    Otherwise hg_build_batch_enqueue: slide i entirely on lane i % lanes; threads (HG_B200_BATCH_THREADS) > 1
    drives the lanes from that many host threads.  Both give the same results bit for bit."""
    import ctypes as C
    import time

    import numpy as np
    lib = L.load_library()
    dev = L.require_cuda(device)
    if threads is None:
        threads = int(os.environ.get("HG_B200_BATCH_THREADS", "1"))
    items = [item if isinstance(item, (tuple, list)) else (item, None) for item in slides]
    num = len(items)
    if num == 0:
        return BatchResults(None, np.zeros((0, 14), np.int64), [], [], k, spatial_k, return_csr)
    t0 = time.perf_counter()
    with torch.cuda.device(dev):
        xs = [L.f32c(x, dev) for x, _ in items]
        ps = [L.f32c(p, dev) if (p is not None and spatial_k + (lambda_g != 0)) else None for _, p in items]
        d = xs[0].shape[1]
        if any(x.shape[1] != d for x in xs):
            raise ValueError("build_batch: all slides of a batch must have the same feature dimension")
        if any(x.shape[0] <= 1 for x in xs):
            raise ValueError("Number of nodes must be greater than 1.")
        pds = {p.shape[1] for p in ps if p is not None}
        if len(pds) > 1:
            raise ValueError("build_batch: all slides of a batch must have the same position dimension")
        pos_dim = pds.pop() if pds else 0
        ns = np.array([x.shape[0] for x in xs], dtype=np.int64)
        sk = np.array([spatial_k if p is not None else 0 for p in ps], dtype=np.int64)
        cap = ns * (k + sk)
        al = lambda b: (b + 255) // 256 * 256                                          # noqa: E731
        sizes = np.stack([al(ns * k * 4), al(ns * k * 4), al(ns * sk * 4), al(ns * sk * 4), al(2 * cap * 8), al(cap * 4),
                          al((ns + 1) * 8) * int(return_csr), al(2 * cap * 8) * int(return_csr),
                          al(2 * cap * 8) * int(return_csr)], axis=1)
        flat = np.concatenate([[0], np.cumsum(sizes.ravel())])
        offs = flat[:-1].reshape(num, 9)
        arena = torch.empty(int(flat[-1]) + 256, dtype=torch.uint8, device=dev)
        base = arena.data_ptr() + (-arena.data_ptr()) % 256
        arena = arena[base - arena.data_ptr():]
        e_dev = torch.empty(num, dtype=torch.int64, device=dev)
        table = np.zeros((num, 14), dtype=np.int64)
        table[:, _F_FEATS] = [x.data_ptr() for x in xs]
        table[:, _F_N] = ns
        table[:, _F_POS] = [0 if p is None else p.data_ptr() for p in ps]
        ptrs = base + offs
        table[:, _F_KIDX], table[:, _F_KVAL] = ptrs[:, 0], ptrs[:, 1]
        has_sp = sk > 0
        table[:, _F_SIDX], table[:, _F_SD2] = np.where(has_sp, ptrs[:, 2], 0), np.where(has_sp, ptrs[:, 3], 0)
        table[:, _F_MAXE], table[:, _F_EI], table[:, _F_W] = cap, ptrs[:, 4], ptrs[:, 5]
        if return_csr:
            table[:, _F_ROWPTR], table[:, _F_COL], table[:, _F_EID] = ptrs[:, 6], ptrs[:, 7], ptrs[:, 8]
        table[:, _F_NE] = e_dev.data_ptr() + 8 * np.arange(num)
        tptr = table.ctypes.data
        if grouped is None:
            grouped = os.environ.get("HG_B200_BATCH_GROUPED", "1") != "0"
        grouped = bool(grouped) and k <= 32
        if group_rows is None:
            group_rows = int(os.environ.get("HG_B200_GROUP_ROWS", str(128 * 1024)))
        nbytes, fbytes = C.c_size_t(0), C.c_size_t(0)
        if grouped:
            L.check(lib.hg_build_batch_grouped_workspace_bytes(tptr, num, d, k, spatial_k, int(return_csr), group_rows,
                                                               C.byref(fbytes), C.byref(nbytes)))
        else:
            L.check(lib.hg_build_batch_workspace_bytes(tptr, num, d, k, spatial_k, int(return_csr), C.byref(nbytes)))
        cur = torch.cuda.current_stream(dev)
        # grouped: stream 0 is the front stream (fused tensor-core kernel), the others are lanes.  (A high-priority
        # front stream, so that the next group's persistent CTAs get onto the SMs while the previous group's re-rank
        # is still draining, was measured and is WORSE: 117 vs 112 ms for the 512 slides -- the fused kernel runs at
        # the power cap, and memory-bound work beside it only takes power away from it.)
        streams = _lane_streams(dev, min(lanes, num) + (1 if grouped else 0))
        start = torch.cuda.Event()
        start.record(cur)
        wss = []
        for i, s in enumerate(streams):
            s.wait_event(start)
            with torch.cuda.stream(s):
                if grouped and i == 0:         # the front stream owns the two alternating group workspaces
                    wss.append((L.Workspace.get(dev, max(fbytes.value, 1), "batch_front0"),
                                L.Workspace.get(dev, max(fbytes.value, 1), "batch_front1")))
                else:
                    wss.append(L.Workspace.get(dev, max(nbytes.value, 1), "batch"))
        t1 = time.perf_counter()
        if grouped:
            front, lane_st, lane_ws = streams[0], streams[1:], wss[1:]
            arr_f = (C.c_void_p * 2)(wss[0][0].data_ptr(), wss[0][1].data_ptr())
            arr_ws = (C.c_void_p * len(lane_st))(*[w.data_ptr() for w in lane_ws])
            arr_st = (C.c_void_p * len(lane_st))(*[s.cuda_stream for s in lane_st])
            L.check(lib.hg_build_batch_grouped_enqueue(tptr, num, d, pos_dim, L.METRIC[kernel], float(lambda_h),
                                                       float(lambda_g if kernel == "gaussian" else 0.0), k, spatial_k,
                                                       group_rows, arr_f, fbytes.value, front.cuda_stream,
                                                       arr_ws, nbytes.value, arr_st, len(lane_st)))
        else:
            arr_ws = (C.c_void_p * len(streams))(*[w.data_ptr() for w in wss])
            arr_st = (C.c_void_p * len(streams))(*[s.cuda_stream for s in streams])
            L.check(lib.hg_build_batch_enqueue(tptr, num, d, pos_dim, L.METRIC[kernel], float(lambda_h),
                                               float(lambda_g if kernel == "gaussian" else 0.0), k, spatial_k,
                                               arr_ws, nbytes.value, arr_st, len(streams), max(1, int(threads))))
        t2 = time.perf_counter()
        for s in streams:
            ev = torch.cuda.Event()
            ev.record(s)
            cur.wait_event(ev)
            arena.record_stream(s)
            e_dev.record_stream(s)
        for t in xs + [p for p in ps if p is not None]:
            for s in streams:
                t.record_stream(s)
        counts = e_dev.cpu().numpy()           # waits for the batch (the calling stream is ordered after every lane)
    LAST_BATCH_STATS.update(prepare_ms=(t1 - t0) * 1e3, enqueue_ms=(t2 - t1) * 1e3, total_ms=(time.perf_counter() - t0) * 1e3,
                            lanes=len(streams) - int(grouped), threads=int(threads), slides=num, grouped=bool(grouped))
    return BatchResults(arena, table, offs, counts, k, spatial_k, return_csr)


_LANES: Dict = {}


def _lane_streams(dev, lanes: int):
    have = _LANES.setdefault(dev.index, [])
    while len(have) < lanes:
        have.append(torch.cuda.Stream(dev))
    return have[:max(1, lanes)]

