"""GPU KMeans (hg_kmeans_fit, one C call) against the Python-driven loop over the building blocks and
against scikit-learn on the host cores.  HG_BENCH_NO_SKLEARN=1 skips the host arm.
    python tools/bench_kmeans.py"""
import os, sys, time, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import synth
from hg_b200.kmeans import kmeans, kmeans_stepwise
from sklearn.cluster import KMeans as Sk


def timed(fn, *a, **kw):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out = fn(*a, **kw)
    torch.cuda.synchronize()
    return time.perf_counter() - t0, out


# the three KMeans calls of process_single_file (grouping, hyperedges, super-patches) and two larger slides
for n, d, c in [(100, 256, 10), (356, 1024, 10), (4096, 1024, 10), (4096, 1024, 100), (16384, 1024, 100), (50000, 1024, 100)]:
    x = synth.make_features(n, d, 80, "gaussian")
    xd = torch.from_numpy(x).cuda()
    kmeans(xd, c, n_init=1); kmeans_stepwise(xd, c, n_init=1)
    fit_s, (_, _, inertia) = timed(kmeans, xd, c, n_init=10)
    step_s, (_, _, inertia_step) = timed(kmeans_stepwise, xd, c, n_init=10)
    row = {"n": n, "d": d, "clusters": c, "fit_s_n_init10": fit_s, "stepwise_s_n_init10": step_s,
           "inertia_fit": inertia, "inertia_stepwise": inertia_step}
    if not os.environ.get("HG_BENCH_NO_SKLEARN"):
        sk_init = 10 if n * c <= 4096 * 10 else 1
        t0 = time.perf_counter(); sk = Sk(n_clusters=c, random_state=42, n_init=sk_init).fit(x)
        sk_s = (time.perf_counter() - t0) * 10 / sk_init
        row.update(sklearn_s_n_init10=sk_s, sklearn_measured_n_init=sk_init, cores=os.cpu_count(),
                   inertia_sklearn=float(sk.inertia_), speedup_vs_sklearn=sk_s / fit_s)
    print(json.dumps(row), flush=True)
