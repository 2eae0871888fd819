"""The reference-facing per-slide pipeline (process_single_file, preprocess_hypergraph.py:514-603) end to end on
synthetic .npz slide containers: wall time per stage, with the reference's scikit-learn KMeans (default) and
with the GPU Lloyd (HG_B200_KMEANS=cuda).
    python tools/bench_process_file.py [N_wsi ...]"""
import json, os, sys, tempfile, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
import hg_b200
from hg_b200 import preprocess_hypergraph as pp, store, synth

sizes = [int(a) for a in sys.argv[1:]] or [4096, 16384, 50000]
d, n_tma = 1024, 256
STAGES = ["load_wsi_data", "load_tma_data", "aggregate_wsi_super_patches", "compute_wsi_tma_similarity",
          "group_by_similarity", "build_hypergraph_knn_kmeans", "save_hypergraph_to_h5"]
clock = {}


def timed(name, fn):
    def wrapper(*a, **kw):
        torch.cuda.synchronize()
        t = time.perf_counter()
        out = fn(*a, **kw)
        torch.cuda.synchronize()
        clock[name] = clock.get(name, 0.0) + time.perf_counter() - t
        return out
    return wrapper


for name in STAGES:
    setattr(pp, name, timed(name, getattr(pp, name)))
kmeans_inner = pp._cluster
pp._cluster = timed("kmeans (inside the stages)", kmeans_inner)

tmp = tempfile.mkdtemp(prefix="hg_b200_bench_")
for n in sizes:
    feats, pos = synth.make_slide(n, d, 40 + n % 97, "gaussian")
    tma = synth.make_features(n_tma, d, 41, "gaussian")
    for backend in ("sklearn", "cuda"):
        if backend == "sklearn" and n > 20000 and not os.environ.get("HG_BENCH_SKLEARN_LARGE"):
            continue                                         # minutes of host KMeans: not a GPU-box measurement
        pp.KMEANS_BACKEND = backend
        path = os.path.join(tmp, f"slide_{n}_{backend}.npz")
        with store.open_store(path, "w") as f:
            g = f.create_group("wsi")
            g["features"] = feats
            g["positions"] = pos
            f.create_group("tma")["features"] = tma
        runs = []
        for rep in range(2):                                 # second run: warm library, workspaces, page cache
            clock.clear()
            torch.cuda.synchronize()
            t = time.perf_counter()
            stats = hg_b200.process_single_file(path, num_wsi_super_patches=100, num_groups=10, hypergraph_k=5,
                                                num_hyperedges=10, lambda_h=1.0 / d, lambda_g=1.0 / 4096)
            torch.cuda.synchronize()
            runs.append((time.perf_counter() - t, dict(clock)))
        total, stages = runs[-1]
        print(json.dumps({"n_wsi": n, "d": d, "n_tma": n_tma, "kmeans": backend, "total_s": total,
                          "stages_s": {k: round(v, 4) for k, v in stages.items()},
                          "first_run_s": runs[0][0], "edges": stats["hypergraph"]["num_edges"],
                          "file_mb": os.path.getsize(path) / 1e6}), flush=True)
        os.unlink(path)
