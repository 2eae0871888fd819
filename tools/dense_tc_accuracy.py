#!/usr/bin/env python
"""Accuracy and speed of the tensor-core dense-output mode (hg_similarity_dense_tc) against a float64 kernel
evaluated with torch on the GPU, over ALL entries:  python tools/dense_tc_accuracy.py [n] [d ...]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import ops, synth          # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    dims = [int(a) for a in sys.argv[2:]] or [1024, 1536]
    for d in dims:
        for with_pos in (False, True):
            feats, pos = synth.make_slide(n, d, 41, "gaussian")
            x = torch.from_numpy(feats).cuda()
            p = torch.from_numpy(pos).cuda() if with_pos else None
            lam_g = 1.0 / 4096 if with_pos else 0.0
            x64 = x.double()
            sq = (x64 * x64).sum(1)
            ref = (sq[:, None] + sq[None, :] - 2.0 * (x64 @ x64.T)).clamp_min_(0.0)
            if with_pos:
                p64 = p.double()
                for c in range(2):
                    df = p64[:, c][:, None] - p64[None, :, c]
                    ref += lam_g * df * df
            ref = torch.exp(-ref)
            res = {"n": n, "d": d, "positions": with_pos}
            for name, min_n in (("tc", 1), ("simt", 0)):
                ops.DENSE_TC_MIN_N = min_n
                K = ops.similarity_dense(x, pos_a=p, pos_b=p, lambda_h=1.0, lambda_g=lam_g)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                e0.record()
                for _ in range(5):
                    K = ops.similarity_dense(x, pos_a=p, pos_b=p, lambda_h=1.0, lambda_g=lam_g)
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / 5
                rel = ((K.double() - ref).abs() / ref)
                res[name] = {"ms": ms, "tflops_algorithmic": 2.0 * n * n * d / (ms * 1e-3) / 1e12,
                             "max_rel_err": float(rel.max()), "mean_rel_err": float(rel.mean())}
                del K, rel
            print(json.dumps(res), flush=True)
            del ref, x64


if __name__ == "__main__":
    main()
