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


def build_batch(slides: Iterable, device=None, lanes: int = 4, threads: Optional[int] = None, **build_kwargs) -> List[dict]:
    """Device-resident batch (no host copies): every slide is enqueued round-robin on `lanes`
    streams without any synchronisation, then the results (device tensors) are gathered.  The
    calling stream is ordered before and after the batch.  threads > 1 (HG_B200_BATCH_THREADS)
    enqueues from that many host threads, lane l driven by thread l % threads -- the C calls
    release the GIL, so the ~30 kernel launches of different slides are issued in parallel."""
    import time
    dev = L.require_cuda(device)
    cur = torch.cuda.current_stream(dev)
    streams = _lane_streams(dev, lanes)
    if threads is None:
        threads = int(os.environ.get("HG_B200_BATCH_THREADS", "1"))
    threads = max(1, min(int(threads), len(streams)))
    items = [item if isinstance(item, (tuple, list)) else (item, None) for item in slides]
    start = torch.cuda.Event()
    start.record(cur)
    for s in streams:
        s.wait_event(start)
    pend: List = [None] * len(items)
    t0 = time.perf_counter()

    def drive(tid: int):
        with torch.cuda.device(dev):
            for i in range(len(items)):
                lane = i % len(streams)
                if lane % threads != tid:
                    continue
                with torch.cuda.stream(streams[lane]):
                    pend[i] = _enqueue(items[i][0], items[i][1], dev, **build_kwargs)

    if threads == 1:
        drive(0)
    else:
        import threading
        errs: List = []

        def guarded(tid):
            try:
                drive(tid)
            except BaseException as e:           # re-raised on the calling thread
                errs.append(e)

        ts = [threading.Thread(target=guarded, args=(t,)) for t in range(threads)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        if errs:
            raise errs[0]
    t1 = time.perf_counter()
    for s in streams:
        ev = torch.cuda.Event()
        ev.record(s)
        cur.wait_event(ev)
    outs = [p.result() for p in pend]
    for out in outs:                           # allocated on a lane stream, consumed on the caller's
        for t in out.values():
            t.record_stream(cur)
    LAST_BATCH_STATS.update(enqueue_ms=(t1 - t0) * 1e3, total_ms=(time.perf_counter() - t0) * 1e3, lanes=len(streams),
                            threads=threads, slides=len(items))
    return outs


_LANES: Dict = {}


def _lane_streams(dev, lanes: int):
    have = _LANES.setdefault(dev.index, [])
    while len(have) < lanes:
        have.append(torch.cuda.Stream(dev))
    return have[:max(1, lanes)]
