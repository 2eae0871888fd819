"""Stream-pipelined per-slide builds for dataset loops (the reference's `process_dataset`
loop, preprocess_hypergraph.py:651-670, is strictly sequential: load -> build -> save).

Three CUDA streams per device:
  * copy-in  : H2D of slide i+1 from pinned host memory (double-buffered device staging),
  * compute  : the hypergraph build of slide i (every C-ABI call is stream-ordered),
  * copy-out : D2H of slide i's outputs into pinned host buffers while slide i+1 computes.
PCIe is full duplex, so a steady-state step costs max(H2D, build, D2H) instead of their sum.
Every slide's H2D and D2H still happen -- they are overlapped, not skipped.
"""
from __future__ import annotations

from collections import deque
from typing import Callable, Deque, Dict, Iterable, Iterator, Optional, Tuple

import torch

from . import _lib as L
from .similarity_kernel import knn_hyperedges

_OUT_KEYS = ("edge_index", "edge_weights", "csr_rowptr", "csr_col", "csr_eid", "knn_idx", "knn_val",
             "spatial_idx", "spatial_d2")


class SlidePipeline:
    """submit(features[, positions]) enqueues one slide; results() yields dicts of pinned host
    tensors in submission order.  `keys` selects which outputs are copied back.  The pinned result
    buffers rotate over two slots: consume (or copy) a result before two more slides finish."""

    def __init__(self, device=None, keys: Tuple[str, ...] = ("edge_index", "edge_weights", "csr_rowptr", "csr_col",
                                                              "csr_eid"),
                 build_fn: Optional[Callable] = None, **build_kwargs):
        self.dev = L.require_cuda(device)
        self.keys = keys
        self.kw = build_kwargs
        self.build_fn = build_fn or knn_hyperedges
        self.s_in = torch.cuda.Stream(self.dev)
        self.s_out = torch.cuda.Stream(self.dev)
        self.s_cmp = torch.cuda.Stream(self.dev)
        self._staged: Deque = deque()          # (x_dev, pos_dev, ready_event)
        self._done: Deque = deque()            # (host_dict, done_event)
        self._host_pool: Dict = {}
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self._built = 0

    # -- stage 1: H2D on the copy-in stream --------------------------------------------------
    def _stage(self, features: torch.Tensor, positions: Optional[torch.Tensor]):
        with torch.cuda.stream(self.s_in):
            x = features.to(self.dev, dtype=torch.float32, non_blocking=True)
            p = positions.to(self.dev, dtype=torch.float32, non_blocking=True) if positions is not None else None
            ev = torch.cuda.Event()
            ev.record(self.s_in)
        self.h2d_bytes += features.numel() * 4 + (positions.numel() * 4 if positions is not None else 0)
        self._staged.append((x, p, ev))

    # -- stage 2 + 3: build on the compute stream, D2H on the copy-out stream ------------------
    def _run_one(self):
        x, p, ev = self._staged.popleft()
        self.s_cmp.wait_event(ev)
        with torch.cuda.stream(self.s_cmp):
            x.record_stream(self.s_cmp)
            if p is not None:
                p.record_stream(self.s_cmp)
            out = self.build_fn(x, p, device=self.dev, **self.kw)
            built = torch.cuda.Event()
            built.record(self.s_cmp)
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
        self._built += 1

    def _host_buffer(self, name: str, t: torch.Tensor) -> torch.Tensor:
        # two rotating pinned buffers per output so slide i's copy never lands on slide i-1's result
        slot = (name, self._built & 1)
        need = t.numel()
        buf = self._host_pool.get(slot)
        if buf is None or buf.dtype != t.dtype or buf.numel() < need:
            buf = torch.empty(max(need, 1), dtype=t.dtype, pin_memory=True)
            self._host_pool[slot] = buf
        return buf[:need].view(t.shape)

    def submit(self, features: torch.Tensor, positions: Optional[torch.Tensor] = None):
        """Enqueue a slide.  The previous slide (if any) is built now, so its compute overlaps
        this slide's H2D."""
        self._stage(features, positions)
        if len(self._staged) > 1:
            self._run_one()

    def results(self, flush: bool = False) -> Iterator[dict]:
        """Yield finished slides in order; with flush=True also build what is still staged."""
        if flush:
            while self._staged:
                self._run_one()
        while self._done:
            host, ev = self._done.popleft()
            ev.synchronize()
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
        # keep at most two results buffered (the host pool rotates two slots per output)
        if len(pipe._done) >= 2:
            yield next(pipe.results())
    yield from pipe.results(flush=True)
