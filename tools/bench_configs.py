#!/usr/bin/env python
"""Side benchmark for the BASELINE.json configs that bench.py does not time by default.

    python tools/bench_configs.py c1 c3 c4            # single GPU
    torchrun --nproc-per-node G tools/bench_configs.py c4 c5

c1  N=4,096 x 1024, cosine kNN k=8 (+ the reference CPU path beside it, this one fits the CPU)
c3  N=100k x 1536, feature k=16 + spatial-grid k=8 hyperedges
c4  512 slides N in [2048, 8192] x 1024 sharded slide-per-GPU (LPT), Gaussian k=16 + CSR
c5  N=250k x 1536 row-partitioned across the ranks, NCCL all-gather of the feature row blocks
One JSON line per config (rank 0).  Device timing: CUDA events, max over ranks.
"""
from __future__ import annotations

import ctypes
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
import hg_b200                                             # noqa: E402
from hg_b200 import _lib, multigpu, synth                  # noqa: E402
from hg_b200 import pipeline as hg_pipeline               # noqa: E402
from hg_b200.pipeline import SlidePipeline, build_batch, build_many     # noqa: E402


def main():
    which = [a for a in sys.argv[1:] if a.startswith("c")] or ["c1", "c3", "c4"]
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load_library()

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def rmax(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(fn, steps, warmup=3):
        for _ in range(warmup):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sync()
        _lib.check(lib.hg_profile_enable(1))
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        sync()
        ms, cnt = ctypes.c_float(0), ctypes.c_int(0)
        _lib.check(lib.hg_profile_collect(ctypes.byref(ms), ctypes.byref(cnt)))
        _lib.check(lib.hg_profile_enable(0))
        return rmax(e0.elapsed_time(e1)) / steps, ms.value / steps

    def emit(**kw):
        if rank == 0:
            print(json.dumps(kw), flush=True)

    for cfg in which:
        if cfg == "c1":
            n, d, k = 4096, 1024, 8
            feats = synth.make_features(n, d, 1235, "cosine")
            x = torch.from_numpy(feats).to(dev)
            ms, kms = timed(lambda: hg_b200.knn_hyperedges(x, k=k, kernel="cosine", return_csr=True), 50)
            sys.path.insert(0, ROOT)
            import bench
            t0 = time.perf_counter()
            bench.cpu_reference_step(feats, k, np.arange(n))
            cpu_s = time.perf_counter() - t0
            emit(config="c1", n=n, d=d, k=k, ms_per_slide=ms, tc_kernel_ms=kms, patches_per_s=n / ms * 1e3,
                 tc_tflops_algorithmic=2.0 * n * n * d / (kms * 1e-3) / 1e12,
                 cpu_reference_port_s=cpu_s, cpu_cores=os.cpu_count(), speedup_vs_cpu=cpu_s * 1e3 / ms)
        elif cfg == "c3":
            n, d, k, ks = 100_000, 1536, 16, 8
            feats, pos = synth.make_slide(n, d, 1237, "gaussian")
            x, p = torch.from_numpy(feats).to(dev), torch.from_numpy(pos).to(dev)
            fn = lambda: hg_b200.knn_hyperedges(x, p, k=k, kernel="gaussian", lambda_h=1.0, lambda_g=1.0 / 128,   # noqa: E731
                                                spatial_k=ks, return_csr=True)
            ms, kms = timed(fn, 3, warmup=2)
            emit(config="c3", n=n, d=d, k=k, spatial_k=ks, ms_per_slide=ms, tc_kernel_ms=kms,
                 patches_per_s=n / ms * 1e3, tc_tflops_algorithmic=2.0 * n * n * d / (kms * 1e-3) / 1e12)
        elif cfg == "c4":
            sizes = synth.slide_sizes(512)
            mine = multigpu.shard_slides(sizes, world)[rank]
            pools = [torch.from_numpy(synth.make_features(8192, 1024, 1238 + s, "gaussian")).pin_memory() for s in range(4)]
            dpools = [q.to(dev) for q in pools]
            kw = dict(k=16, kernel="gaussian", lambda_h=1.0, return_csr=True)

            def resident():
                for i in mine:
                    hg_b200.knn_hyperedges(dpools[i % 4][: int(sizes[i])], **kw)

            lanes = int(os.environ.get("HG_B200_PIPE_LANES", "4"))

            def batched():                                 # stream-ordered builds on `lanes` streams, no host sync per slide
                build_batch((dpools[i % 4][: int(sizes[i])] for i in mine), device=dev, lanes=lanes, **kw)

            pipe = SlidePipeline(dev, lanes=lanes, **kw)

            def from_host():
                for _ in build_many((pools[i % 4][: int(sizes[i])] for i in mine), pipeline=pipe):
                    pass
                torch.cuda.synchronize()

            from_host()                                    # warm the pipeline's buffers

            ms_seq, kms = timed(resident, 1, warmup=1)
            ms, _ = timed(batched, 2, warmup=1)
            sync()
            t0 = time.perf_counter()
            from_host()
            e2e_ms = rmax((time.perf_counter() - t0) * 1e3)
            tot = int(sizes.sum())
            emit(config="c4", slides=512, gpus=world, patches=tot, lanes=lanes, ms_total=ms, ms_total_one_by_one=ms_seq,
                 tc_kernel_ms_total=kms, batch_host=dict(hg_pipeline.LAST_BATCH_STATS),
                 slides_per_s=512 / ms * 1e3, patches_per_s=tot / ms * 1e3,
                 e2e_ms_total=e2e_ms, e2e_slides_per_s=512 / e2e_ms * 1e3, e2e_patches_per_s=tot / e2e_ms * 1e3)
        elif cfg == "c5":
            n, d, k = int(os.environ.get("HG_BENCH_C5_N", 250_000)), 1536, 16
            if world == 1:
                local_feats = torch.from_numpy(synth.make_features(n, d, 1239, "gaussian")).to(dev)
                fn = lambda: hg_b200.knn_hyperedges(local_feats, k=k, kernel="gaussian", lambda_h=1.0, return_csr=True)   # noqa: E731
                ms, kms = timed(fn, 2, warmup=1)
                emit(config="c5", n=n, d=d, k=k, gpus=1, ms_per_slide=ms, tc_kernel_ms=kms, patches_per_s=n / ms * 1e3)
                continue
            # every rank generates only its own row block (same generator, block-wise seeds); two partitions:
            # near-equal blocks with gather-then-compute, 256-aligned blocks with the overlapped exchange.
            # The modes alternate round by round on the same box (clocks drift under the power cap).
            modes = {}
            for mode, ranges, kwm in (("classic_row_blocks", multigpu.row_range, dict(aligned=False, symmetric=False)),
                                      ("symmetric_tile_pairs", multigpu.row_range_aligned, dict(aligned=True, symmetric=True)),
                                      ("classic_aligned_blocks", multigpu.row_range_aligned, dict(aligned=True, symmetric=False))):
                r0, r1 = ranges(n, rank, world)
                modes[mode] = dict(rows=r1 - r0, kw=kwm, ms=[], kms=[], out=None,
                                   feats=torch.from_numpy(synth.make_features(r1 - r0, d, 1239 + rank, "gaussian")).to(dev))
            for rnd in range(int(os.environ.get("HG_BENCH_C5_ROUNDS", 3))):
                for mode, m in modes.items():
                    def fn():
                        m["out"] = multigpu.knn_hyperedges_row_partitioned(
                            m["feats"], n, k=k, kernel="gaussian", lambda_h=1.0, return_csr=True, **m["kw"])
                    ms, kms = timed(fn, 3, warmup=2 if rnd == 0 else 1)
                    m["ms"].append(ms)
                    m["kms"].append(kms)
            for mode, m in modes.items():
                ms, kms = float(np.median(m["ms"])), float(np.median(m["kms"]))
                emit(config="c5", mode=mode, ran_as=m["out"]["mode"], n=n, d=d, k=k, gpus=world,
                     rows_rank0=m["rows"], ms_per_slide=ms, ms_rounds=m["ms"], patches_per_s=n / ms * 1e3,
                     tc_kernel_ms_sum=kms, tc_tflops_algorithmic_per_gpu=2.0 * m["rows"] * n * d / (kms * 1e-3) / 1e12,
                     edges=int(m["out"]["edge_index"].shape[1]),
                     note="tc_kernel_ms_sum adds up the fused kernel's launches on rank 0 (pre-pass + sweep + guarded pass)")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
