import ctypes, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import _lib, ops, synth
lib = _lib.load_library()
def run(n, d, k, cgs, reps=5):
    x = torch.from_numpy(synth.make_features(n, d, 1, "gaussian")).cuda()
    ref = None
    for cg in cgs:
        os.environ["HG_B200_TC_CG"] = str(cg)
        for _ in range(2): idx, val = ops.knn_feature(x, k, "gaussian")
        torch.cuda.synchronize()
        if ref is None: ref = idx.clone()
        same = bool((idx == ref).all())
        lib.hg_profile_enable(1)
        for _ in range(reps): ops.knn_feature(x, k, "gaussian")
        ms, c = ctypes.c_float(0), ctypes.c_int(0)
        lib.hg_profile_collect(ctypes.byref(ms), ctypes.byref(c)); lib.hg_profile_enable(0)
        t = ms.value / c.value
        print(f"n={n} d={d} cg={cg}: {t:.3f} ms  {2.0*n*n*d/t/1e9:.1f} TFLOP/s alg  same_as_first={same}", flush=True)
run(50000, 1024, 16, [1, 2, 1, 2, 1, 2], reps=6)
run(100000, 1536, 16, [1, 2, 1, 2], reps=2)
run(4096, 1024, 8, [1, 2, 1, 2], reps=20)
