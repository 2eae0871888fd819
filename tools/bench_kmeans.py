import os, sys, time, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import synth
from hg_b200.kmeans import kmeans
from sklearn.cluster import KMeans as Sk
for n, d, c in [(4096, 1024, 10), (50000, 1024, 100)]:
    x = synth.make_features(n, d, 80, "gaussian")
    xd = torch.from_numpy(x).cuda()
    kmeans(xd, c, n_init=1)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    _, _, inertia = kmeans(xd, c, n_init=10)
    torch.cuda.synchronize(); gpu_s = time.perf_counter() - t0
    sk_init = 10 if n <= 8192 else 1
    t0 = time.perf_counter(); sk = Sk(n_clusters=c, random_state=42, n_init=sk_init).fit(x); sk_s = (time.perf_counter() - t0) * 10 / sk_init
    print(json.dumps({"n": n, "d": d, "clusters": c, "gpu_s_n_init10": gpu_s, "sklearn_s_n_init10": sk_s,
                      "sklearn_measured_n_init": sk_init, "cores": os.cpu_count(),
                      "inertia_gpu": inertia, "inertia_sklearn": float(sk.inertia_), "speedup": sk_s / gpu_s}), flush=True)
