"""A/B of environment switches of the fused kernel in one process, interleaved rounds, lists checked equal.
    python tools/ab_env.py [--cases c2,c3] "" "HG_B200_TC_MC=1" "HG_B200_TC_SYM_PRODS=3" "HG_B200_TC_MC=1;HG_B200_TC_PAIR=0"
Every argument is one setting (NAME=VALUE pairs separated by ";", "" = defaults); fused-kernel ms from the in-stream
profile events (hg_profile_*)."""
import ctypes, json, os, statistics, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import _lib, ops, synth
lib = _lib.load_library()
ALL = {"c2": (50000, 1024, 16, False, None, 5), "c3": (100000, 1536, 16, True, None, 3), "n4096": (4096, 1024, 16, False, None, 30),
       "n16k": (16384, 1024, 16, False, None, 10), "c5rows": (100000, 1536, 16, False, (0, 12500), 3),
       "c3d1024": (100000, 1024, 16, True, None, 3), "n10k": (10240, 1024, 16, False, None, 20), "n8192": (8192, 1024, 16, False, None, 20), "n24k": (24576, 1024, 16, False, None, 10),
       "n10kd256": (10240, 256, 16, False, None, 20), "c2d256": (50000, 256, 16, False, None, 8), "c2d512": (50000, 512, 16, False, None, 6)}
args = sys.argv[1:]
cases = ["c2", "c3"]
if args and args[0] == "--cases":
    cases = args[1].split(","); args = args[2:]
settings = args or [""]
names = set()
for s in settings:
    for kv in filter(None, s.split(";")): names.add(kv.split("=")[0])
for name in cases:
    n, d, k, pos, rows, reps = ALL[name]
    feats, p = synth.make_slide(n, d, 1, "gaussian")
    x = torch.from_numpy(feats).cuda()
    kw = dict(positions=torch.from_numpy(p).cuda(), lambda_h=1.0, lambda_g=1.0 / 128) if pos else {}
    if rows: kw["rows"] = rows
    t = {s: [] for s in settings}
    ref = None
    for r in range(3):
        for s in settings:
            for nm in names: os.environ.pop(nm, None)
            for kv in filter(None, s.split(";")):
                a, b = kv.split("="); os.environ[a] = b
            idx, _ = ops.knn_feature(x, k, "gaussian", **kw)
            if ref is None: ref = idx.clone()
            assert torch.equal(idx, ref), f"{name}: setting '{s}' gives different lists"
            torch.cuda.synchronize()
            lib.hg_profile_enable(1)
            for _ in range(reps): ops.knn_feature(x, k, "gaussian", **kw)
            ms, c = ctypes.c_float(0), ctypes.c_int(0)
            lib.hg_profile_collect(ctypes.byref(ms), ctypes.byref(c)); lib.hg_profile_enable(0)
            t[s].append(ms.value / reps)
    for nm in names: os.environ.pop(nm, None)
    print(json.dumps({"case": name, "n": n, "d": d, "fused_kernel_ms": {s or "default": round(statistics.median(v), 4) for s, v in t.items()},
                      "all": {s or "default": [round(q, 3) for q in v] for s, v in t.items()}}), flush=True)
    del x
