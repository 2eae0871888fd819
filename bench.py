#!/usr/bin/env python
"""Benchmark of the hypergraph-construction hot path (BASELINE.json metric: patches/sec and
slides/sec hypergraph build at 1/2/4/8 B200; % of the FP32-emulation tensor-core roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Headline workload = BASELINE config C3, the largest single-GPU configuration: one synthetic slide
of N=100,000 patches x D=1536, feature-space Gaussian kNN (k=16, spatial factor lambda_g on) +
spatial-grid kNN (k=8) -> de-duplicated edge_index, cosine edge weights, CSR incidence.  A step is
one full per-slide build.  With N > 1 GPUs every rank builds its own slide of that shape
(slide-per-GPU, no data-path collective; weak scaling).

`value`    : patches/s with the slide resident in HBM (CUDA events, max over ranks).
`e2e`      : the same build through hg_b200.pipeline.build_many with pinned HOST inputs: H2D of the
             features and positions and D2H of every output inside the timed region, every step.
`roofline` : the fused similarity + top-k stage (knn_tc_kernel: threshold pre-pass + symmetric half
             pass), algorithmic 2*N^2*D FLOP per slide (full N x N, symmetry NOT credited, SURVEY.md
             8(d)) over its CUDA-event duration, against the measured 16-bit tensor peak / MMAs per
             logical product (2 in the symmetric passes, 3 elsewhere: split-FP16 operands).  The
             symmetric pass issues ~0.57 of the tiles, so `frac` can exceed 1; `executed_frac` = MMA
             work actually issued / the raw tensor peak.
`secondary`: C2 (N=50k x 1024, Gaussian kNN k=16 + CSR), last round's headline, for continuity.
`c4`       : 512 slides (N in [2048, 8192], D=1024) sharded slide-per-GPU by LPT: slides/s resident
             (one grouped build per rank) and end to end from pinned host memory.
`c5` (N>1) : one slide of 250k x 1536 row-partitioned over the N ranks (NCCL all-gather of the
             feature row blocks, per-rank fused kernel on its rows, all-gather of the lists).
`parity`   : after the timed region, sampled rows of the last build are checked on the GPU against a
             float64 evaluation of the same scores (torch, checker only).
`cpu_baseline` / `--impl reference`: the UNMODIFIED reference (baseline/_ref, installed by
             baseline/install_ref.py) -- its own build_hypergraph_knn_kmeans
             (preprocess_hypergraph.py:335-433), sklearn KMeans replaced by a stand-in that adds no
             cliques -- on a bounded sample of the workload: a 4,096-row sub-slide of the C3 slide
             (the reference's kNN is O(N) per patch and its N x N temporaries do not fit at 100k, so
             the sub-slide FAVOURS the reference; it also has no spatial-kNN stage).
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))

C3 = {"name": "C3: single slide N=100k x D=1536, Gaussian feature kNN k=16 (lambda_g on) + spatial-grid kNN k=8 "
              "+ de-dup + cosine weights + CSR incidence",
      "n": 100_000, "d": 1536, "k": 16, "spatial_k": 8, "kernel": "gaussian", "lambda_h": 1.0, "lambda_g": 1.0 / 128,
      "seed": 1237}
C2 = {"name": "C2: single slide N=50k x D=1024, Gaussian kNN k=16 + CSR incidence",
      "n": 50_000, "d": 1024, "k": 16, "spatial_k": 0, "kernel": "gaussian", "lambda_h": 1.0, "lambda_g": 0.0,
      "seed": 1236}
C5 = {"n": 250_000, "d": 1536, "k": 16, "seed": 1239}
METRIC = "patches/sec hypergraph build (C3: N=100k x 1536, feature kNN k=16 + spatial kNN k=8 + CSR)"
REF_SAMPLE_ROWS = 4096


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"bf16": p["bf16_tflops"], "bf16_sustained": p["bf16_tflops_sustained"], "hbm": p["hbm_gbs"],
                "source": "MEASURED_PEAKS.json"}
    return {"bf16": 1590.0, "bf16_sustained": 1400.0, "hbm": 6650.0, "source": "B200_PROFILING.md fallback"}


# --------------------------------------------------------------------------- clocks sampling
class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int, period_s: float = 0.004):
        self.gpu = gpu_index
        self.period = period_s
        self.rows = []
        self.proc = None

    # NVML is polled every few milliseconds (the timed region is a fraction of a second: nvidia-smi's
    # 100 ms loop would see it a handful of times); nvidia-smi -lms is the fallback without the binding.
    _REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def start(self):
        self.nvml = None
        self._stop = threading.Event()
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            uuid = torch.cuda.get_device_properties(self.gpu).uuid
            try:
                handle = pynvml.nvmlDeviceGetHandleByUUID("GPU-" + str(uuid))
            except Exception:                                   # noqa: BLE001 -- older bindings / odd UUID form
                handle = pynvml.nvmlDeviceGetHandleByIndex(self.gpu)
            self.nvml = (pynvml, handle)
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception as e:                                  # noqa: BLE001
            print(f"[bench] NVML polling unavailable ({type(e).__name__}: {e}); falling back to nvidia-smi", file=sys.stderr)
            self.nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _poll(self):
        pynvml, h = self.nvml
        mx = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
        while not self._stop.is_set():
            try:
                sm = float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                mask = int(pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)) if hasattr(
                    pynvml, "nvmlDeviceGetCurrentClocksEventReasons") else int(pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                row = ["", sm, mx, "", ""] + [("Active" if mask & bit else "Not Active") for _, bit in self._REASONS]
                self.rows.append(row)
            except Exception:                                   # noqa: BLE001
                pass
            self._stop.wait(self.period)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.nvml is not None:
            self._stop.set()
            self.thread.join(1.0)
        elif self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        else:
            time.sleep(0.15)
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7),
                              ("sw_power_cap", 8)):
                if len(r) > col and str(r[col]).lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "source": "nvml" if self.nvml is not None else "nvidia-smi", "reasons": sorted(reasons)}


# ------------------------------------------------------------------------ reference CPU path
def use_all_host_threads() -> int:
    """torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU arm must run on all the host threads
    it can use, so the BLAS / OpenMP pools are widened again at run time.  Returns the thread count."""
    cores = os.cpu_count() or 1
    for var in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[var] = str(cores)
    try:
        import sklearn.neighbors                   # noqa: F401 - load its OpenMP runtime before widening the pools
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=cores)            # process-wide when not used as a context manager
    except Exception:                              # noqa: BLE001 - best effort, the count is reported either way
        pass
    try:
        import torch
        torch.set_num_threads(cores)
    except Exception:                              # noqa: BLE001
        pass
    return cores


def load_reference():
    """The unmodified reference from baseline/_ref, or None (then the CPU arm falls back to the port)."""
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import ref_loader
    if not ref_loader.available():
        return None
    pkg, pp, _ = ref_loader.load()
    return pkg, pp, ref_loader


def reference_substep(ref, feats: np.ndarray, k: int, cliques: bool = False, num_hyperedges: int = 10):
    """ONE call of the reference's own build_hypergraph_knn_kmeans (preprocess_hypergraph.py:335-433) on the
    given rows: sklearn NearestNeighbors(k+1, 'euclidean') + its Python pair / set / cosine loops.
    cliques=False installs the no-clique KMeans stand-in (kNN + de-dup + weights only)."""
    import torch
    from sklearn.cluster import KMeans
    pkg, pp, loader = ref
    pp.KMeans = KMeans if cliques else loader.NoCliqueKMeans
    n = feats.shape[0]
    n_tma = min(256, n // 4)                       # the function concatenates its two inputs first (:374)
    t0 = time.perf_counter()
    ei, w, stats = pkg.build_hypergraph_knn_kmeans(torch.from_numpy(feats[: n - n_tma]), torch.from_numpy(feats[n - n_tma:]),
                                                   np.zeros(n - n_tma, dtype=np.int64), k=k, num_hyperedges=num_hyperedges)
    return time.perf_counter() - t0, int(ei.shape[1])


def port_step(feats: np.ndarray, k: int, rows: np.ndarray, nn=None):
    """The reference's kNN call verbatim (NearestNeighbors(k+1,'euclidean').fit(X).kneighbors,
    preprocess_hypergraph.py:379-382) for the sampled query rows against ALL candidates, then the oracle's
    vectorised restatement of :385-422 (drop self, canonicalise, de-duplicate, cosine weights)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_np as oracle
    from sklearn.neighbors import NearestNeighbors
    if nn is None:
        nn = NearestNeighbors(n_neighbors=k + 1, metric="euclidean").fit(feats)
    _, idx = nn.kneighbors(feats[rows])
    lists = np.full((feats.shape[0], k), -1, dtype=np.int64)
    lists[rows] = idx[:, 1:]
    ei = oracle.edges_from_knn(lists)
    oracle.edge_weights_cosine(feats, ei, dtype=np.float32)
    return nn


def time_port(feats, k, target_s=8.0):
    """Port on a bounded sample of query rows at the FULL candidate count: rows/s."""
    n = feats.shape[0]
    rng = np.random.default_rng(0)
    nn = port_step(feats, k, rng.choice(n, 128, replace=False))        # fit + thread-pool warm-up
    t0 = time.perf_counter()
    port_step(feats, k, rng.choice(n, 256, replace=False), nn)
    probe = time.perf_counter() - t0
    sample = int(min(n, max(256, 256 * target_s / max(probe, 1e-3)))) // 256 * 256
    rows = rng.choice(n, sample, replace=False)
    t0 = time.perf_counter()
    port_step(feats, k, rows, nn)
    dt = time.perf_counter() - t0
    return sample / dt, dt, sample


def cpu_baseline_block(feats_c3: np.ndarray, steps: int, warmup: int, with_port: bool, with_c1: bool,
                       c1_untouched: bool = True):
    """The CPU arm.  value = patches/s of the unmodified reference on a REF_SAMPLE_ROWS-row sub-slide of C3."""
    cores = use_all_host_threads()
    from hg_b200 import synth
    w = C3
    ref = load_reference()
    rng = np.random.default_rng(0)
    sub = np.ascontiguousarray(feats_c3[np.sort(rng.choice(w["n"], REF_SAMPLE_ROWS, replace=False))])
    out = {"unit": "patches/s", "cores": cores}
    if ref is not None:
        for _ in range(warmup):
            reference_substep(ref, sub, w["k"])
        times, edges = [], 0
        for _ in range(steps):
            dt, edges = reference_substep(ref, sub, w["k"])
            times.append(dt)
        dt = float(np.mean(times))
        out.update(value=REF_SAMPLE_ROWS / dt, kind="reference", ms_per_step=dt * 1e3,
                   sample=f"unmodified build_hypergraph_knn_kmeans (baseline/_ref, KMeans stand-in without cliques) on a "
                          f"{REF_SAMPLE_ROWS}-row sub-slide of the C3 slide (D={w['d']}, k={w['k']}, E={edges}), {dt:.2f} s per "
                          f"step; the reference has no spatial-kNN stage and its kNN costs O(N) per patch, so a sub-slide "
                          f"of N/24 rows favours it")
    if with_port or ref is None:
        v, dt, sample = time_port(feats_c3, w["k"])
        port = {"value": v, "unit": "patches/s", "kind": "port",
                "sample": f"{sample} of {w['n']} query rows against all {w['n']} candidates: sklearn brute kNN as "
                          f"preprocess_hypergraph.py:379-382 + vectorised de-dup / cosine weights (oracle), {dt:.1f} s"}
        if ref is None:
            out.update(value=v, kind="port", ms_per_step=dt * 1e3, sample=port["sample"])
        else:
            out["port_full_n"] = port
    if with_c1 and ref is not None:
        # the reference's own CPU-runnable case, C1 (N=4,096 x 1024, cosine kNN k=8), untouched code:
        # once with the KMeans stand-in (kNN + de-dup + weights) and once fully unmodified (num_hyperedges=10 cliques)
        f1 = synth.make_features(4096, 1024, 1235, "cosine")
        t_knn, e_knn = reference_substep(ref, f1, 8)
        c1 = {"n": 4096, "d": 1024, "k": 8, "knn_only_s": t_knn, "knn_only_edges": e_knn,
              "knn_only_patches_per_s": 4096 / t_knn, "kind": "reference"}
        if c1_untouched and os.environ.get("HG_BENCH_C1_CLIQUES", "1") == "1":
            t_full, e_full = reference_substep(ref, f1, 8, cliques=True)
            c1.update(untouched_s=t_full, untouched_edges=e_full, untouched_patches_per_s=4096 / t_full)
        out["c1_reference"] = c1
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from hg_b200 import synth
    w = C3
    feats = synth.make_features(w["n"], w["d"], w["seed"], "gaussian")
    cpu = cpu_baseline_block(feats, steps=args.steps, warmup=min(args.warmup, 1), with_port=True,
                             with_c1=args.gpus == 1)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": cpu["value"], "unit": "patches/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": cpu["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": w["name"], "n": w["n"], "d": w["d"], "k": w["k"], "spatial_k": w["spatial_k"],
                   "sample": cpu["sample"]},
        "cpu_baseline": cpu,
        "e2e": {"value": cpu["value"], "unit": "patches/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ----------------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    import hg_b200
    from hg_b200 import _lib, multigpu, synth
    from hg_b200.pipeline import SlidePipeline, build_batch, build_many

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback for the product path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load_library()
    peaks = measured_peaks()
    steps, warmup = args.steps, max(args.warmup, 3)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v: float) -> float:
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def min_over_ranks(v: float) -> float:
        return -max_over_ranks(-v)

    def timed_resident(fn, nsteps, nwarm):
        """CUDA events around nsteps calls, max over ranks; also the fused kernel's own time per step."""
        for _ in range(nwarm):
            fn()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        _lib.check(lib.hg_profile_enable(1))
        l0 = lib.hg_launch_count()
        ev0.record()
        for _ in range(nsteps):
            fn()
        ev1.record()
        torch.cuda.synchronize()
        launches = (lib.hg_launch_count() - l0) // nsteps
        k_ms, k_n = ctypes.c_float(0), ctypes.c_int(0)
        _lib.check(lib.hg_profile_collect(ctypes.byref(k_ms), ctypes.byref(k_n)))
        _lib.check(lib.hg_profile_enable(0))
        barrier()
        return max_over_ranks(ev0.elapsed_time(ev1)) / nsteps, k_ms.value / nsteps, int(launches)

    def slide_bench(w, nsteps, nwarm, sample_clocks):
        """Resident + end-to-end timing of one single-slide workload; every rank builds its own slide."""
        n, d, k = w["n"], w["d"], w["k"]
        feats = synth.make_features(n, d, w["seed"] + rank, "gaussian")
        pos = synth.make_positions(n, w["seed"] + rank) if w["spatial_k"] or w["lambda_g"] else None
        x_host = torch.from_numpy(feats).pin_memory()
        p_host = torch.from_numpy(pos).pin_memory() if pos is not None else None
        x_dev = x_host.to(dev)
        p_dev = p_host.to(dev) if p_host is not None else None
        kw = dict(k=k, kernel=w["kernel"], lambda_h=w["lambda_h"], lambda_g=w["lambda_g"], spatial_k=w["spatial_k"],
                  return_csr=True)
        last = {}

        def build():
            last["out"] = hg_b200.knn_hyperedges(x_dev, p_dev, device=dev, **kw)

        sampler = ClockSampler(local) if sample_clocks else None
        for _ in range(nwarm):
            build()
        if sampler:
            sampler.start()
        ms_step, kern_ms, launches = timed_resident(build, nsteps, 0)
        clocks = sampler.stop() if sampler else None
        out = last["out"]
        num_edges = int(out["edge_index"].shape[1])

        # ---- end to end: pinned host inputs -> H2D -> build -> D2H of every output, every step ----------
        keys = ("edge_index", "edge_weights", "csr_rowptr", "csr_col", "csr_eid")
        pipe = SlidePipeline(dev, keys=keys, **kw)

        def e2e_pipelined(m):
            got = 0
            items = ((x_host, p_host) if p_host is not None else x_host for _ in range(m))
            for res in build_many(items, pipeline=pipe):
                got += int(res["edge_index"].shape[1] == num_edges)
            torch.cuda.synchronize()
            assert got == m, "pipelined build returned a different edge count"

        e2e_pipelined(3)
        d2h = sum(out[name].numel() * out[name].element_size() for name in keys)
        h2d = x_host.numel() * 4 + (p_host.numel() * 4 if p_host is not None else 0)
        e2e_ms = None
        for _ in range(3):                               # host-side jitter: best of three K-step regions
            barrier()
            t0 = time.perf_counter()
            e2e_pipelined(nsteps)
            wall_ms = (time.perf_counter() - t0) * 1e3
            barrier()
            rep = max_over_ranks(wall_ms) / nsteps
            e2e_ms = rep if e2e_ms is None else min(e2e_ms, rep)

        # ---- the host feed by itself: pinned H2D of the same buffers, all ranks at once ------------------
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(4):
            x_dev.copy_(x_host, non_blocking=True)
        ev1.record()
        torch.cuda.synchronize()
        h2d_gbs = min_over_ranks(4 * x_host.numel() * 4 / (ev0.elapsed_time(ev1) * 1e-3) / 1e9)
        barrier()
        return dict(feats=feats, pos=pos, x_dev=x_dev, p_dev=p_dev, out=out, ms_step=ms_step, kern_ms=kern_ms,
                    launches=launches, clocks=clocks, num_edges=num_edges, e2e_ms=e2e_ms, h2d=h2d, d2h=d2h,
                    h2d_gbs=h2d_gbs, lanes=pipe.lanes)

    def roofline_of(w, kern_ms, ms_step, sustained: bool):
        n, d, k = w["n"], w["d"], w["k"]
        flop = 2.0 * n * n * d                                   # algorithmic, per launch
        achieved = flop / (kern_ms * 1e-3) / 1e12 if kern_ms > 0 else None
        bf16 = peaks["bf16_sustained"] if sustained else peaks["bf16"]
        sym_stride, mma_flop = ctypes.c_int(0), ctypes.c_double(0.0)
        _lib.check(lib.hg_knn_feature_plan(n, d, k, 0, n, ctypes.byref(sym_stride), ctypes.byref(mma_flop)))
        prods, pre_prods = ctypes.c_int(0), ctypes.c_int(0)
        _lib.check(lib.hg_knn_feature_products(n, d, k, 0, n, ctypes.byref(prods), ctypes.byref(pre_prods)))
        per = float(prods.value)                                  # MMAs per logical product of the main pass
        executed = mma_flop.value / (kern_ms * 1e-3) / 1e12 if kern_ms > 0 else None
        mode = (f"symmetric half pass + 1/{sym_stride.value} threshold pre-pass" if sym_stride.value > 0 else "full pass")
        traffic = None
        prof = os.path.join(ROOT, "profiles", "knn_tc_traffic.json")       # from the committed ncu --set full captures
        if os.path.exists(prof):
            t = json.load(open(prof))
            traffic = t.get(f"dram_bytes_per_launch_n{n}_d{d}", t.get("dram_bytes_per_launch") if n == 50_000 else None)
        return {"bound": "tensor", "kernel": f"knn_tc_kernel (fused split-FP16 similarity + top-k, {prods.value} MMAs per product, {mode})",
                "achieved": achieved, "peak": bf16 / per, "unit": "TFLOP/s",
                "frac": (achieved / (bf16 / per)) if achieved else None, "traffic": traffic,
                "kernel_ms": kern_ms, "kernel_share_of_step": kern_ms / ms_step, "mma_per_product": prods.value,
                # continuity with the lines of rounds 1-2 before the two-product pass (peak / 3, SURVEY 8(d)'s 3xBF16 figure)
                "frac_vs_3_product_roofline": (achieved / (bf16 / 3.0)) if achieved else None,
                "peak_source": (f"{peaks['source']} bf16_tflops{'_sustained' if sustained else ''} (kind::f16 MMAs: FP16 operands "
                                f"run at the BF16 rate) / {prods.value} MMAs per product "
                                f"(burst {peaks['bf16']:.0f}, sustained {peaks['bf16_sustained']:.0f} TFLOP/s); "
                                f"{'sustained: the timed region is many long launches back to back' if sustained else 'burst: 5 ms launches'}"),
                "frac_note": "frac = algorithmic 2*N^2*D FLOP (full N x N, symmetry not credited) / time / peak: above 1 is "
                             "possible because the symmetric pass issues ~0.57 of the tiles; executed_frac = issued 16-bit MMA "
                             "FLOP / time / raw 16-bit tensor peak is the hardware-utilisation figure",
                "algorithmic_flop": flop, "executed_mma_flop": mma_flop.value,
                "executed_tflops": executed, "executed_frac": (executed / bf16) if executed else None,
                "executed_frac_of_burst": (executed / peaks["bf16"]) if executed else None}

    def parity_check(w, r, rows_n=256):
        """Sampled rows of the last build against float64 scores evaluated with torch on the GPU (checker)."""
        n, k = w["n"], w["k"]
        g = torch.Generator(device="cpu").manual_seed(5)
        rows = torch.randperm(n, generator=g)[:rows_n].to(dev)
        x64 = r["x_dev"].double()
        sq = (x64 * x64).sum(1)
        d2 = (sq[rows, None] + sq[None, :] - 2.0 * (x64[rows] @ x64.T)).clamp_min_(0.0)
        score = w["lambda_h"] * d2
        if w["lambda_g"]:
            p64 = r["p_dev"].double()
            score += w["lambda_g"] * ((p64[rows, None, :] - p64[None, :, :]) ** 2).sum(-1)
        score[torch.arange(rows_n, device=dev), rows] = float("inf")
        ref_val, ref_idx = torch.topk(score, k + 1, dim=1, largest=False)
        got_idx = r["out"]["knn_idx"][rows].long()
        got_val = r["out"]["knn_val"][rows].double()
        same_set = (torch.sort(got_idx, 1).values == torch.sort(ref_idx[:, :k], 1).values).all(1)
        # a differing row is a near-tie iff the float64 gap between rank k and k+1 is below 1e-5 relative
        gap = (ref_val[:, k] - ref_val[:, k - 1]) / ref_val[:, k - 1].clamp_min(1e-30)
        bad = (~same_set) & (gap > 1e-5)
        exact = (got_idx == ref_idx[:, :k]).all(1)
        rel = ((got_val - torch.exp(-ref_val[:, :k])).abs() / torch.exp(-ref_val[:, :k]))[exact]
        sp_ok = None
        if w["spatial_k"]:
            p64 = r["p_dev"].double()
            dg = ((p64[rows, None, :] - p64[None, :, :]) ** 2).sum(-1)
            dg[torch.arange(rows_n, device=dev), rows] = float("inf")
            key = dg * (2.0 * n) + torch.arange(n, device=dev, dtype=torch.float64)[None, :]     # (d2, index) order; d2 integral
            sref = torch.topk(key, w["spatial_k"], dim=1, largest=False).indices
            sp_ok = bool((sref == r["out"]["spatial_idx"][rows].long()).all())
        return {"checker": "torch float64 on the GPU, after the timed region", "rows_checked": rows_n,
                "rows_set_equal": int(same_set.sum()), "rows_near_tie": int(((~same_set) & (gap <= 1e-5)).sum()),
                "rows_wrong": int(bad.sum()), "max_rel_err_values": float(rel.max()) if rel.numel() else None,
                "spatial_lists_equal": sp_ok, "ok": bool(bad.sum() == 0 and (rel.numel() == 0 or rel.max() <= 1e-5)
                                                         and sp_ok is not False)}

    # ======================================= C3: the headline ========================================
    r3 = slide_bench(C3, steps, warmup, sample_clocks=True)
    n = C3["n"]
    value = world * n / (r3["ms_step"] * 1e-3)
    e2e_value = world * n / (r3["e2e_ms"] * 1e-3)
    roofline = roofline_of(C3, r3["kern_ms"], r3["ms_step"], sustained=True)
    parity = parity_check(C3, r3) if rank == 0 else None
    feats_c3 = r3["feats"]
    for key in ("x_dev", "p_dev", "out"):
        r3[key] = None
    _lib.Workspace.clear()
    torch.cuda.empty_cache()

    # ======================================= C2: continuity =========================================
    s2 = max(3, min(steps, 10))
    r2 = slide_bench(C2, s2, 3, sample_clocks=False)
    roof2 = roofline_of(C2, r2["kern_ms"], r2["ms_step"], sustained=False)
    secondary = {"workload": C2["name"], "value": world * C2["n"] / (r2["ms_step"] * 1e-3), "unit": "patches/s",
                 "ms_per_step": r2["ms_step"], "steps": s2, "edges": r2["num_edges"],
                 "e2e": {"value": world * C2["n"] / (r2["e2e_ms"] * 1e-3), "ms_per_step": r2["e2e_ms"],
                         "h2d_bytes_per_step": r2["h2d"], "d2h_bytes_per_step": r2["d2h"]},
                 "gpu_launches": r2["launches"],
                 "roofline": {key: roof2[key] for key in ("achieved", "peak", "frac", "kernel_ms", "executed_tflops",
                                                          "executed_frac", "peak_source")}}
    r2 = None
    _lib.Workspace.clear()
    torch.cuda.empty_cache()

    # ============================ C4: 512 small slides, slide-per-GPU (LPT) ===========================
    sizes = synth.slide_sizes(512)
    mine = multigpu.shard_slides(sizes, world)[rank]
    pools = [torch.from_numpy(synth.make_features(8192, 1024, 1238 + s, "gaussian")).pin_memory() for s in range(4)]
    dpools = [q.to(dev) for q in pools]
    kw4 = dict(k=16, kernel="gaussian", lambda_h=1.0, return_csr=True)
    lanes4 = int(os.environ.get("HG_B200_PIPE_LANES", "4"))

    def c4_resident():
        build_batch((dpools[i % 4][: int(sizes[i])] for i in mine), device=dev, lanes=lanes4, **kw4)

    ms4, kms4, launches4 = timed_resident(c4_resident, 3, 2)
    pipe4 = SlidePipeline(dev, lanes=lanes4, **kw4)

    def c4_from_host():
        for _ in build_many((pools[i % 4][: int(sizes[i])] for i in mine), pipeline=pipe4):
            pass
        torch.cuda.synchronize()

    c4_from_host()
    e2e4 = None
    for _ in range(2):
        barrier()
        t0 = time.perf_counter()
        c4_from_host()
        wall = max_over_ranks((time.perf_counter() - t0) * 1e3)
        e2e4 = wall if e2e4 is None else min(e2e4, wall)
    _p4, _q4 = ctypes.c_int(0), ctypes.c_int(0)
    _lib.check(lib.hg_knn_feature_products(4096, 1024, 16, 0, 4096, ctypes.byref(_p4), ctypes.byref(_q4)))
    c4_prods = float(_p4.value)                     # MMAs per product of the grouped launches (small slides)
    tot4 = int(sizes.sum())
    flop4 = float(sum(2.0 * float(sizes[i]) ** 2 * 1024 for i in mine))
    c4 = {"workload": "C4: 512 slides, N ~ U{2048..8192} x D=1024, Gaussian kNN k=16 + CSR, LPT slide-per-GPU",
          "slides": 512, "patches": tot4, "gpus": world, "lanes": lanes4, "ms_total": ms4,
          "slides_per_s": 512 / (ms4 * 1e-3), "patches_per_s": tot4 / (ms4 * 1e-3),
          "tc_kernel_ms_rank0": kms4, "tc_tflops_algorithmic_rank0": flop4 / (kms4 * 1e-3) / 1e12 if kms4 > 0 else None,
          "tc_mma_per_product": c4_prods,
          "tc_frac_of_burst_roofline": (flop4 / (kms4 * 1e-3) / 1e12) / (peaks["bf16"] / c4_prods) if kms4 > 0 else None,
          "tc_frac_of_burst_peak_over_3": (flop4 / (kms4 * 1e-3) / 1e12) / (peaks["bf16"] / 3.0) if kms4 > 0 else None,
          "whole_batch_tflops_algorithmic_rank0": flop4 / (ms4 * 1e-3) / 1e12,
          "gpu_launches": launches4,
          "e2e": {"ms_total": e2e4, "slides_per_s": 512 / (e2e4 * 1e-3), "patches_per_s": tot4 / (e2e4 * 1e-3),
                  "h2d_bytes": int(sum(int(sizes[i]) for i in mine)) * 1024 * 4}}
    pools = dpools = pipe4 = None
    _lib.Workspace.clear()
    torch.cuda.empty_cache()

    # ===================== C5 (N > 1): one 250k x 1536 slide row-partitioned over the ranks =================
    c5 = None
    if world > 1 and not args.no_c5:
        n5, d5, k5 = int(os.environ.get("HG_BENCH_C5_N", C5["n"])), C5["d"], C5["k"]
        r0, r1 = multigpu.row_range_aligned(n5, rank, world)
        blk = torch.from_numpy(synth.make_features(r1 - r0, d5, C5["seed"] + rank, "gaussian")).to(dev)
        last5 = {}

        def c5_step():
            last5["out"] = multigpu.knn_hyperedges_row_partitioned(blk, n5, k=k5, kernel="gaussian", lambda_h=1.0,
                                                                   return_csr=True, aligned=True)

        ms5, kms5, launches5 = timed_resident(c5_step, 3, 2)
        # all-gather alone (the one big collective), timed the same way
        def gather_only():
            multigpu.all_gather_rows(blk, n5, None, multigpu.row_range_aligned)
        ag_ms, _, _ = timed_resident(gather_only, 3, 1)
        # bit-equality of the distributed lists with rank 0's single-GPU lists for a row block
        feats_all = multigpu.all_gather_rows(blk, n5, None, multigpu.row_range_aligned)
        equal = None
        if rank == 0:
            from hg_b200 import ops
            a = (n5 // 3) // 256 * 256
            b = min(n5, a + 2048)
            idx1, val1 = ops.knn_feature(feats_all, k5, "gaussian", lambda_h=1.0, rows=(a, b))
            equal = bool((idx1 == last5["out"]["knn_idx"][a:b]).all() and (val1 == last5["out"]["knn_val"][a:b]).all())
        flop5 = 2.0 * (r1 - r0) * n5 * d5
        c5 = {"workload": f"C5: single slide N={n5} x D={d5}, Gaussian kNN k={k5} + CSR, row-partitioned over {world} GPUs",
              "n": n5, "d": d5, "k": k5, "gpus": world, "ms_per_slide": ms5, "patches_per_s": n5 / (ms5 * 1e-3),
              "all_gather_ms": ag_ms, "fused_kernel_ms_rank0": kms5, "mode": last5["out"].get("mode", "classic row blocks"),
              "tc_tflops_algorithmic_per_gpu": flop5 / (kms5 * 1e-3) / 1e12 if kms5 > 0 else None,
              "edges": int(last5["out"]["edge_index"].shape[1]), "gpu_launches": launches5,
              "lists_bit_equal_to_single_gpu_rows": equal}
        blk = feats_all = None
        last5.clear()

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cpu = cpu_baseline_block(feats_c3, steps=3, warmup=1, with_port=True, with_c1=True, c1_untouched=False)
        line = {
            "metric": METRIC, "value": value, "unit": "patches/s", "n_gpus": world, "steps": steps,
            "warmup": warmup, "ms_per_step": r3["ms_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f16-split+f32", "data": "synthetic",
            "config": {"workload": C3["name"], "n": n, "d": C3["d"], "k": C3["k"], "spatial_k": C3["spatial_k"],
                       "kernel": C3["kernel"], "lambda_h": C3["lambda_h"], "lambda_g": C3["lambda_g"],
                       "edges": r3["num_edges"], "slides_per_s": world / (r3["ms_step"] * 1e-3),
                       "parallelism": f"slide-per-gpu x{world}",
                       "l2": "inputs larger than L2 (614 MB FP32 features + 614 MB FP16 hi/lo operands per step)"},
            "e2e": {"value": e2e_value, "unit": "patches/s", "ms_per_step": r3["e2e_ms"],
                    "api": f"hg_b200.pipeline.build_many (H2D stream / {r3['lanes']} stream-ordered build lanes / D2H stream, "
                           "wall clock, best of 3 K-step regions)",
                    "h2d_bytes_per_step": r3["h2d"], "d2h_bytes_per_step": r3["d2h"],
                    "h2d_ceiling_gbs": r3["h2d_gbs"],
                    "h2d_ceiling_note": "pinned-host -> device cudaMemcpyAsync of the feature buffer alone, all ranks at "
                                        "once, slowest rank; the feed a step needs is h2d_bytes_per_step / ms_per_step"},
            "gpu_launches": r3["launches"], "roofline": roofline, "parity": parity, "cpu_baseline": cpu,
            "clocks": r3["clocks"], "secondary": secondary, "c4": c4,
        }
        if c5 is not None:
            line["c5"] = c5
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--no-cpu-baseline", dest="no_cpu_baseline", action="store_true")
    ap.add_argument("--no-c5", dest="no_c5", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
