import ctypes, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import _lib, ops, synth
lib = _lib.load_library()
def run(n, d, k, plans, reps=5):
    x = torch.from_numpy(synth.make_features(n, d, 1, "gaussian")).cuda()
    for plan in plans:
        if plan: os.environ["HG_B200_TC_PLAN"] = plan
        else: os.environ.pop("HG_B200_TC_PLAN", None)
        for _ in range(2): ops.knn_feature(x, k, "gaussian")
        torch.cuda.synchronize()
        lib.hg_profile_enable(1)
        for _ in range(reps): ops.knn_feature(x, k, "gaussian")
        ms, c = ctypes.c_float(0), ctypes.c_int(0)
        lib.hg_profile_collect(ctypes.byref(ms), ctypes.byref(c)); lib.hg_profile_enable(0)
        t = ms.value / c.value
        print(f"n={n} d={d} plan={plan or 'auto'}: {t:.3f} ms  {2.0*n*n*d/t/1e9:.1f} TFLOP/s alg", flush=True)
run(50000, 1024, 16, ["", "1,3", "2,6", "2,4", "4,4", "4,8", "1,1", ""], reps=6)
run(100000, 1536, 16, ["", "4,8", "4,4", "2,6", "1,2"], reps=2)
