#!/usr/bin/env python
"""How many rows the re-rank sends to the exact fallback (HG_B200_TC_DEBUG prints it), per config."""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
os.environ["HG_B200_TC_DEBUG"] = "1"
from hg_b200 import ops, synth          # noqa: E402
for (n, d, k, seed, pos) in [(50000, 1024, 16, 1236, False), (100000, 1536, 16, 1237, True), (4096, 1024, 16, 1238, False),
                             (2048, 1024, 16, 1238, False), (8192, 1024, 16, 1238, False), (4096, 1024, 8, 1235, False)]:
    f, p = synth.make_slide(n, d, seed, "gaussian")
    x = torch.from_numpy(f).cuda()
    pp = torch.from_numpy(p).cuda() if pos else None
    print(f"--- n={n} d={d} k={k}", file=sys.stderr, flush=True)
    for _ in range(2):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        ops.knn_feature(x, k, "gaussian", positions=pp, lambda_h=1.0, lambda_g=1.0 / 128 if pos else 0.0, return_cos=True)
        torch.cuda.synchronize(); print(f"    {1e3 * (time.perf_counter() - t0):.2f} ms", file=sys.stderr, flush=True)
