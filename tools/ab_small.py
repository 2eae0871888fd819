"""Small-slide sweep of the fused kNN stage (BASELINE config C4 sizes): work-list plans
(HG_B200_TC_PLAN=cc,splits) and the symmetric mode, time of the tensor-core region per slide."""
import ctypes, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import _lib, ops, synth
lib = _lib.load_library()


def timed(x, k, reps=20):
    for _ in range(3): ops.knn_feature(x, k, "gaussian")
    torch.cuda.synchronize()
    lib.hg_profile_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): ops.knn_feature(x, k, "gaussian")
    e1.record(); torch.cuda.synchronize()
    ms, c = ctypes.c_float(0), ctypes.c_int(0)
    lib.hg_profile_collect(ctypes.byref(ms), ctypes.byref(c)); lib.hg_profile_enable(0)
    return ms.value / c.value, e0.elapsed_time(e1) / reps


quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
for n in (2048, 3072, 4096, 6144, 8192, 10240):
    x = torch.from_numpy(synth.make_features(n, 1024, 1, "gaussian")).cuda()
    for plan in ([""] if quick else ["", "1,1", "1,2", "1,4", "1,8", "2,2", "2,4", "2,8", "4,4", "4,8"]):
        os.environ["HG_B200_TC_SYM"] = "0"
        if plan: os.environ["HG_B200_TC_PLAN"] = plan
        else: os.environ.pop("HG_B200_TC_PLAN", None)
        t, tot = timed(x, 16)
        print(f"n={n} plan={plan or 'auto':5s} sym=0: tc {t*1e3:7.1f} us  knn stage {tot*1e3:7.1f} us  {2.0*n*n*1024/t/1e9:6.1f} TFLOP/s alg", flush=True)
    os.environ.pop("HG_B200_TC_PLAN", None)
    for sym, hi in (((8, 0),) if quick else ((8, 0), (8, 1), (4, 1))):
        os.environ["HG_B200_TC_SYM"] = str(sym); os.environ["HG_B200_TC_PRE_HI"] = str(hi)
        t, tot = timed(x, 16)
        print(f"n={n} plan=auto  sym={sym} pre_hi={hi}: tc {t*1e3:7.1f} us  knn stage {tot*1e3:7.1f} us  {2.0*n*n*1024/t/1e9:6.1f} TFLOP/s alg", flush=True)
    os.environ.pop("HG_B200_TC_PRE_HI", None)
