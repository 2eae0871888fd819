"""Same-box A/B of two builds of libhg_b200.so: runs `tools/prof_tc.py`-style timing of the fused kernel in
alternating subprocesses (HG_B200_LIB selects the library).   python tools/ab_lib.py OLD.so [rounds]"""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
old = os.path.abspath(sys.argv[1])
rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 3
CODE = r'''
import ctypes, json, os, sys, torch
sys.path.insert(0, os.path.join(%r, "multimodal-fusion_b200"))
from hg_b200 import _lib, ops, synth
lib = _lib.load_library()
res = {}
CASES = (("c2", 50000, 1024, 16, False, 6), ("c3", 100000, 1536, 16, True, 3), ("n4096", 4096, 1024, 16, False, 30), ("c5rows", 100000, 1536, 16, False, 3))
if os.environ.get("AB_CASES") == "c3":
    CASES = (("c3", 100000, 1536, 16, True, 3),)
if os.environ.get("AB_CASES") == "isolate":
    CASES = (("n100k_d1536_nopos", 100000, 1536, 16, False, 3), ("n100k_d1024_pos", 100000, 1024, 16, True, 3),
             ("n50k_d1536_nopos", 50000, 1536, 16, False, 4), ("n100k_d1024_nopos", 100000, 1024, 16, False, 3))
for name, n, d, k, pos, reps in CASES:
    feats, p = synth.make_slide(n, d, 1, "gaussian")
    x = torch.from_numpy(feats).cuda()
    kw = dict(positions=torch.from_numpy(p).cuda(), lambda_h=1.0, lambda_g=1.0 / 128) if pos else {}
    if name == "c5rows": kw["rows"] = (0, n // 8)          # classic full pass on a row block (row-partitioned mode)
    for _ in range(2): ops.knn_feature(x, k, "gaussian", **kw)
    torch.cuda.synchronize()
    lib.hg_profile_enable(1)
    for _ in range(reps): ops.knn_feature(x, k, "gaussian", **kw)
    ms, c = ctypes.c_float(0), ctypes.c_int(0)
    lib.hg_profile_collect(ctypes.byref(ms), ctypes.byref(c)); lib.hg_profile_enable(0)
    res[name] = ms.value / reps
    del x
print(json.dumps(res))
''' % ROOT
for r in range(rounds):
    for tag, path in (("prev", old), ("new", "")):
        env = dict(os.environ)
        if path: env["HG_B200_LIB"] = path
        else: env.pop("HG_B200_LIB", None)
        out = subprocess.run([sys.executable, "-c", CODE], env=env, capture_output=True, text=True)
        line = out.stdout.strip().splitlines()[-1] if out.stdout.strip() else out.stderr[-500:]
        print(tag, r, line, flush=True)
