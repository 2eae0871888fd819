"""A/B of the multicast-pair variant of the fused kernel (HG_B200_TC_MC) in one process, interleaved rounds."""
import ctypes, os, statistics, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import _lib, ops, synth
lib = _lib.load_library()
CASES = (("c2", 50000, 1024, 16, False, None, 5), ("c3", 100000, 1536, 16, True, None, 2), ("n4096", 4096, 1024, 16, False, None, 30),
         ("n8192", 8192, 1024, 16, False, None, 20), ("c5rows", 100000, 1536, 16, False, (0, 12500), 3))
for name, n, d, k, pos, rows, reps in CASES:
    feats, p = synth.make_slide(n, d, 1, "gaussian")
    x = torch.from_numpy(feats).cuda()
    kw = dict(positions=torch.from_numpy(p).cuda(), lambda_h=1.0, lambda_g=1.0 / 128) if pos else {}
    if rows: kw["rows"] = rows
    t = {0: [], 1: []}
    ref = None
    for r in range(3):
        for mc in (0, 1):
            os.environ["HG_B200_TC_MC"] = str(mc)
            idx, _ = ops.knn_feature(x, k, "gaussian", **kw)
            if ref is None: ref = idx.clone()
            assert torch.equal(idx, ref), f"{name}: mc={mc} differs"
            torch.cuda.synchronize()
            lib.hg_profile_enable(1)
            for _ in range(reps): ops.knn_feature(x, k, "gaussian", **kw)
            ms, c = ctypes.c_float(0), ctypes.c_int(0)
            lib.hg_profile_collect(ctypes.byref(ms), ctypes.byref(c)); lib.hg_profile_enable(0)
            t[mc].append(ms.value / reps)
    print(f"{name}: single {statistics.median(t[0]):.4f} ms  {[round(v, 4) for v in t[0]]}   multicast pairs {statistics.median(t[1]):.4f} ms  {[round(v, 4) for v in t[1]]}", flush=True)
    del x
