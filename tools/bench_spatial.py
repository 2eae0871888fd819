"""Spatial kNN (A3) timing: brute force vs the uniform-grid path, cell-size sweep.
    python tools/bench_spatial.py [N] [K]"""
import json, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import ops, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 8
pos = torch.from_numpy(synth.make_positions(n, 1237)).cuda()


def timed(reps=20):
    for _ in range(3):
        ops.knn_spatial(pos, k)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        out = ops.knn_spatial(pos, k)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


os.environ["HG_B200_SP_GRID"] = "0"
ms, ref = timed(3)
print(json.dumps({"n": n, "k": k, "path": "brute", "ms": ms}))
os.environ["HG_B200_SP_GRID"] = "1"
for cell in (1, 1.5, 2, 3, 4, 6, 8):
    os.environ["HG_B200_SP_CELL"] = str(cell)
    ms, out = timed()
    same = bool(torch.equal(out[0], ref[0]) and torch.equal(out[1], ref[1]))
    print(json.dumps({"n": n, "k": k, "path": "grid", "cell_target": cell, "ms": ms, "equals_brute": same,
                      "algorithmic_GBps": n * (8 + 8 * k) / ms / 1e6}))
