"""Compute-side cost of the staged kNN (hg_knn_staged_*) on ONE GPU: the per-GPU workload of the
row-partitioned config C5 (rows [0, n/G) of an n x d slide, all columns local) as one hg_knn_feature call
and as G column panels with 1..3 list-slot lanes.  Modes alternate round by round.
    python tools/bench_staged.py [n] [d] [G] [k]"""
import json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import multigpu, ops, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 250_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 1536
G = int(sys.argv[3]) if len(sys.argv) > 3 else 8
k = int(sys.argv[4]) if len(sys.argv) > 4 else 16
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(5)
x = torch.randn((n, d), device=dev, generator=gen) * (1.0 / np.sqrt(d))
ranges = [multigpu.row_range_aligned(n, r, G) for r in range(G)]
rows = ranges[0]
widest = max(b - a for a, b in ranges)


def one_shot():
    return ops.knn_feature(x, k, "gaussian", rows=rows, return_cos=True)


def staged(lanes):
    def run():
        st = ops.StagedKnn(n, d, k, "gaussian", rows, widest, lanes=lanes, device=dev)
        for a, b in ranges:
            st.add_panel(x[a:b], a, b)
        return st.finish(x, return_cos=True)
    return run


def staged_g(g, lanes=1):
    width = -(-(-(-n // g)) // 256) * 256
    cuts = [(a, min(n, a + width)) for a in range(0, n, width)]
    own = next(i for i, (a, b) in enumerate(cuts) if a <= rows[0] and rows[1] <= b)
    order = [cuts[own]] + [c for i, c in enumerate(cuts) if i != own]

    def run():
        st = ops.StagedKnn(n, d, k, "gaussian", rows, min(width, n), lanes=lanes, device=dev)
        for a, b in order:
            st.add_panel(x[a:b], a, b)
        return st.finish(x, return_cos=True)
    return run


if os.environ.get("HG_BENCH_STAGED_SWEEP"):
    modes = {"one_shot": one_shot}
    for g in (1, 2, 4, 8):
        modes[f"staged_panels{g}"] = staged_g(g)
else:
    modes = {"one_shot": one_shot, "staged_lanes1": staged(1), "staged_lanes2": staged(2), "staged_lanes3": staged(3)}
times = {m: [] for m in modes}
ref = None
for rnd in range(4):
    for name, fn in modes.items():
        out = fn()
        if ref is None:
            ref = out
        same = all(torch.equal(a, b) for a, b in zip(out, ref))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(3):
            fn()
        e1.record()
        torch.cuda.synchronize()
        times[name].append(e0.elapsed_time(e1) / 3)
        if rnd == 0:
            print(json.dumps({"mode": name, "equals_one_shot": same}), flush=True)
for name, t in times.items():
    print(json.dumps({"mode": name, "n": n, "d": d, "panels": G, "rows": rows[1] - rows[0], "k": k,
                      "ms_median": float(np.median(t[1:])), "ms_rounds": t}), flush=True)
