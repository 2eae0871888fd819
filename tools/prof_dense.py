"""Profiling target: the FP32 dense similarity kernel on one square input.   python tools/prof_dense.py N D"""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import ops, synth
n, d = int(sys.argv[1]), int(sys.argv[2])
x = torch.from_numpy(synth.make_features(n, d, 1, "gaussian")).cuda()
for _ in range(2):
    K = ops.similarity_dense(x, lambda_h=1.0)
torch.cuda.synchronize()
