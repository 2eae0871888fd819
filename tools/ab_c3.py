"""C3-shaped sweeps of the fused kernel on one box: unit plans (HG_B200_TC_PLAN), the L2 budget of the A blocks
(HG_B200_TC_A_MB), pre-pass stride, and the D / positions / N factors in isolation.   python tools/ab_c3.py"""
import ctypes, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import _lib, ops, synth
lib = _lib.load_library()
KEYS = ("HG_B200_TC_PLAN", "HG_B200_TC_A_MB", "HG_B200_TC_SYM", "HG_B200_TC_PRE_HI", "HG_B200_TC_PAIR", "HG_B200_TC_HINTS")


def timed(x, k, kw, env, reps=3):
    for key in KEYS:
        os.environ.pop(key, None)
    os.environ.update(env)
    for _ in range(2):
        ops.knn_feature(x, k, "gaussian", **kw)
    torch.cuda.synchronize()
    lib.hg_profile_enable(1)
    for _ in range(reps):
        ops.knn_feature(x, k, "gaussian", **kw)
    ms, c = ctypes.c_float(0), ctypes.c_int(0)
    lib.hg_profile_collect(ctypes.byref(ms), ctypes.byref(c)); lib.hg_profile_enable(0)
    for key in KEYS:
        os.environ.pop(key, None)
    return ms.value / reps


def case(n, d, pos, envs, k=16):
    feats, p = synth.make_slide(n, d, 1, "gaussian")
    x = torch.from_numpy(feats).cuda()
    kw = dict(positions=torch.from_numpy(p).cuda(), lambda_h=1.0, lambda_g=1.0 / 128) if pos else {}
    for env in envs:
        t = timed(x, k, kw, env)
        print(f"n={n} d={d} pos={int(pos)} {env or 'default'}: {t:.3f} ms  {2.0 * n * n * d / t / 1e9:.1f} TFLOP/s alg", flush=True)
    del x


plans = [{}] + [{"HG_B200_TC_PLAN": p} for p in ("4,8", "4,4", "2,8", "2,6", "2,4", "1,8", "1,4")]
budgets = [{"HG_B200_TC_A_MB": str(m)} for m in (16, 24, 32, 64, 100)]
extra = [{"HG_B200_TC_SYM": "8"}, {"HG_B200_TC_SYM": "32"}, {"HG_B200_TC_PRE_HI": "1"}, {"HG_B200_TC_PAIR": "0"},
         {"HG_B200_TC_HINTS": "nn"}, {"HG_B200_TC_HINTS": "ll"}, {"HG_B200_TC_HINTS": "lf"}, {}]
case(100000, 1536, True, plans + budgets + extra)
for n, d, pos in ((100000, 1536, False), (100000, 1024, True), (100000, 1024, False), (50000, 1536, False), (50000, 1024, False)):
    case(n, d, pos, [{}])
