"""Profiling target: a few hg_knn_feature calls on one configuration (for ncu launch lists / captures)."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import ops, synth
n, d, k = (int(a) for a in sys.argv[1:4])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 2
with_pos = len(sys.argv) > 5 and sys.argv[5] == "pos"
feats, pos = synth.make_slide(n, d, 1, "gaussian")
x = torch.from_numpy(feats).cuda()
kw = dict(positions=torch.from_numpy(pos).cuda(), lambda_h=1.0, lambda_g=1.0 / 128) if with_pos else {}
for _ in range(reps):
    ops.knn_feature(x, k, "gaussian", **kw)
torch.cuda.synchronize()
