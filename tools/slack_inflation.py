#!/usr/bin/env python
"""VERDICT r01 item 8(a), asked when the operands were split BF16: would a 1xBF16 (hi*hi only) MAIN pass pay?  Its
scores are off by up to SLACK_HI * scale (2^-8 + 2^-11, the constant of that time) instead of SLACK_3X * scale (2^-11),
scale = |x_i|^2 + max |x|^2, so every column whose
exact score lies within that slack of the row's k-th best has to be kept as a candidate and re-ranked exactly.
This counts those columns per row with float64 scores (torch on the GPU, sampled rows), for both slacks:
    python tools/slack_inflation.py [c2|c3]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import synth          # noqa: E402

SLACK_3X, SLACK_HI = 2.0 ** -11, 2.0 ** -8 + 2.0 ** -11


def main():
    for cfg in (sys.argv[1:] or ["c2", "c3"]):
        n, d, k, seed = {"c2": (50_000, 1024, 16, 1236), "c3": (100_000, 1536, 16, 1237)}[cfg]
        x = torch.from_numpy(synth.make_features(n, d, seed, "gaussian")).cuda().double()
        sq = (x * x).sum(1)
        rows = torch.randperm(n, generator=torch.Generator().manual_seed(1))[:4096].cuda()
        out = {"config": cfg, "n": n, "d": d, "k": k, "rows_sampled": int(rows.numel())}
        counts = {"3x": [], "hi": []}
        for a in range(0, rows.numel(), 1024):
            rr = rows[a:a + 1024]
            s = (sq[rr, None] + sq[None, :] - 2.0 * (x[rr] @ x.T)).clamp_min_(0.0)
            s[torch.arange(rr.numel(), device="cuda"), rr] = float("inf")
            kth = torch.topk(s, k, dim=1, largest=False).values[:, -1]
            scale = sq[rr] + sq.max()
            for name, sl in (("3x", SLACK_3X), ("hi", SLACK_HI)):
                counts[name].append((s <= (kth + sl * scale)[:, None]).sum(1))
        for name in counts:
            c = torch.cat(counts[name]).double()
            out[f"candidates_within_slack_{name}"] = {"mean": float(c.mean()), "median": float(c.median()),
                                                      "p99": float(torch.quantile(c, 0.99)), "max": float(c.max()),
                                                      "rows_over_24": float((c > 24).double().mean()),
                                                      "rows_over_64": float((c > 64).double().mean())}
        print(json.dumps(out), flush=True)
        del x


if __name__ == "__main__":
    main()
