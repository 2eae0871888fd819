#!/usr/bin/env python
"""One grouped C4 batch (512 slides) under the profiler: warm-up outside the capture range.
   ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python tools/prof_c4.py"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import synth
from hg_b200.pipeline import build_batch
num = int(sys.argv[1]) if len(sys.argv) > 1 else 512
sizes = synth.slide_sizes(512)[:num]
pools = [torch.from_numpy(synth.make_features(8192, 1024, 1238 + s, "gaussian")).cuda() for s in range(4)]
kw = dict(k=16, kernel="gaussian", lambda_h=1.0, return_csr=True)
run = lambda: build_batch((pools[i % 4][: int(sizes[i])] for i in range(num)), lanes=4, **kw)
run(); torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
run(); torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
