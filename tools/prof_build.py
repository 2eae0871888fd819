"""Profiling target: a few full per-slide builds of one configuration (for ncu launch lists).
    python tools/prof_build.py N D K [SPATIAL_K] [REPS]"""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
import hg_b200
from hg_b200 import synth
n, d, k = (int(a) for a in sys.argv[1:4])
ks = int(sys.argv[4]) if len(sys.argv) > 4 else 0
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 2
feats, pos = synth.make_slide(n, d, 1, "gaussian")
x, p = torch.from_numpy(feats).cuda(), torch.from_numpy(pos).cuda()
for _ in range(reps):
    hg_b200.knn_hyperedges(x, p if ks else None, k=k, kernel="gaussian", lambda_h=1.0, lambda_g=(1.0 / 128 if ks else 0.0),
                           spatial_k=ks, return_csr=True)
torch.cuda.synchronize()
