"""A/B of the symmetric (upper-triangle) tensor-core pass against the classic full pass:
same inputs, neighbour lists compared exactly, time of the fused stage from the in-library timer."""
import ctypes, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import _lib, ops, synth
lib = _lib.load_library()

def run(n, d, k, modes, metric="gaussian", with_pos=False, reps=5, pre_hi=0):
    os.environ["HG_B200_TC_PRE_HI"] = str(pre_hi)
    feats, pos = synth.make_slide(n, d, 1, "gaussian" if metric != "cosine" else "cosine")
    x = torch.from_numpy(feats).cuda()
    p = torch.from_numpy(pos).cuda() if with_pos else None
    ref = None
    for mode in modes:
        os.environ["HG_B200_TC_SYM"] = str(mode)
        kw = dict(positions=p, lambda_h=1.0, lambda_g=1.0 / 128) if with_pos else {}
        os.environ["HG_B200_TC_DEBUG"] = "1"
        idx, val = ops.knn_feature(x, k, metric, **kw)
        os.environ.pop("HG_B200_TC_DEBUG")
        ops.knn_feature(x, k, metric, **kw)
        torch.cuda.synchronize()
        if ref is None: ref = (idx.clone(), val.clone())
        same = int((idx != ref[0]).any(1).sum())
        vdiff = float((val - ref[1]).abs().max())
        lib.hg_profile_enable(1)
        for _ in range(reps): ops.knn_feature(x, k, metric, **kw)
        ms, c = ctypes.c_float(0), ctypes.c_int(0)
        lib.hg_profile_collect(ctypes.byref(ms), ctypes.byref(c)); lib.hg_profile_enable(0)
        t = ms.value / c.value
        print(f"n={n} d={d} k={k} {metric} pos={with_pos} pre_hi={pre_hi} sym={mode}: {t:.3f} ms  {2.0*n*n*d/t/1e9:.1f} TFLOP/s alg  "
              f"rows differing from first={same} max|dval|={vdiff:.2e}", flush=True)

if __name__ == "__main__":
    quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
    run(50000, 1024, 16, [0, 16, 8])
    run(50000, 1024, 16, [16, 8, 16, 8], pre_hi=1)
    if not quick:
        run(100000, 1536, 16, [0, 16], with_pos=True, reps=2)
        run(100000, 1536, 16, [16, 8], with_pos=True, reps=2, pre_hi=1)
        run(20000, 1024, 8, [0, 8, 16], metric="cosine")
        run(20000, 1024, 8, [8, 16], metric="cosine", pre_hi=1)
        run(9000, 512, 8, [0, 8])
        run(9000, 512, 8, [8], pre_hi=1)
