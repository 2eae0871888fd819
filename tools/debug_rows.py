import os, sys, time
os.environ["HG_B200_TC_DEBUG"] = "1"
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import ops, synth
n, d, k = 250_000, 1536, 16
x = torch.cat([torch.from_numpy(synth.make_features(31250, d, 1239 + r, "gaussian")) for r in range(8)]).cuda()
for rows in ((0, 31250), (31250, 62500), (218750, 250000)):
    for rep in range(2):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        ops.knn_feature(x, k, "gaussian", lambda_h=1.0, rows=rows, return_cos=True)
        torch.cuda.synchronize(); print(rows, f"{1e3 * (time.perf_counter() - t0):.1f} ms", flush=True)
