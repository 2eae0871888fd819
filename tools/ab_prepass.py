"""A/B of the symmetric mode's threshold pre-pass: (column stride, 3xBF16 | hi-only product).  The
configurations are interleaved over several rounds (the power cap makes back-to-back runs drift) and
the median time of the fused tensor-core region is reported with the bucket occupancy."""
import ctypes, os, statistics, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import _lib, ops, synth
lib = _lib.load_library()
CONFIGS = [(16, 0), (8, 1), (16, 1), (8, 0)]


def run(n, d, k, with_pos, rounds, reps):
    feats, pos = synth.make_slide(n, d, 1, "gaussian")
    x = torch.from_numpy(feats).cuda()
    kw = dict(positions=torch.from_numpy(pos).cuda(), lambda_h=1.0, lambda_g=1.0 / 128) if with_pos else {}
    times = {c: [] for c in CONFIGS}
    ref = None
    for r in range(rounds):
        for stride, hi in CONFIGS:
            os.environ["HG_B200_TC_SYM"], os.environ["HG_B200_TC_PRE_HI"] = str(stride), str(hi)
            if r == 0:
                os.environ["HG_B200_TC_DEBUG"] = "1"
                idx, _ = ops.knn_feature(x, k, "gaussian", **kw)
                os.environ.pop("HG_B200_TC_DEBUG")
                ref = idx.clone() if ref is None else ref
                assert torch.equal(idx, ref)
            ops.knn_feature(x, k, "gaussian", **kw)
            torch.cuda.synchronize()
            lib.hg_profile_enable(1)
            for _ in range(reps):
                ops.knn_feature(x, k, "gaussian", **kw)
            ms, c = ctypes.c_float(0), ctypes.c_int(0)
            lib.hg_profile_collect(ctypes.byref(ms), ctypes.byref(c)); lib.hg_profile_enable(0)
            times[(stride, hi)].append(ms.value / reps)
    for (stride, hi), t in times.items():
        print(f"n={n} d={d} k={k} pos={with_pos} stride={stride} hi_only={hi}: median {statistics.median(t):.3f} ms  "
              f"min {min(t):.3f}  all {[round(v, 3) for v in t]}", flush=True)


run(50000, 1024, 16, False, rounds=5, reps=4)
run(100000, 1536, 16, True, rounds=3, reps=2)
run(20000, 1024, 8, False, rounds=5, reps=6)
