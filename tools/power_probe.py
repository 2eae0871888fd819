"""SM clock, power draw and throttle reasons (NVML, 10 ms samples) while one configuration of the feature kNN runs back
to back for a few seconds; fused-kernel ms from the in-stream profile events.
    python tools/power_probe.py [c2|c3] [seconds]"""
import ctypes, json, os, statistics, sys, threading, time, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
import pynvml
from hg_b200 import _lib, ops, synth
lib = _lib.load_library()
cfg = sys.argv[1] if len(sys.argv) > 1 else "c3"
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 4.0
n, d, k, pos = {"c2": (50000, 1024, 16, False), "c3": (100000, 1536, 16, True)}[cfg]
feats, p = synth.make_slide(n, d, 1, "gaussian")
x = torch.from_numpy(feats).cuda()
kw = dict(positions=torch.from_numpy(p).cuda(), lambda_h=1.0, lambda_g=1.0 / 128) if pos else {}
pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
samples, stop = [], False
def sampler():
    while not stop:
        samples.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(h) / 1e3,
                        pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)))
        time.sleep(0.01)
for _ in range(2): ops.knn_feature(x, k, "gaussian", **kw)
torch.cuda.synchronize()
t = threading.Thread(target=sampler); t.start()
lib.hg_profile_enable(1)
t0, reps = time.perf_counter(), 0
while time.perf_counter() - t0 < secs:
    ops.knn_feature(x, k, "gaussian", **kw); reps += 1
    if reps % 4 == 0: torch.cuda.synchronize()
torch.cuda.synchronize()
ms, c = ctypes.c_float(0), ctypes.c_int(0)
lib.hg_profile_collect(ctypes.byref(ms), ctypes.byref(c)); lib.hg_profile_enable(0)
stop = True; t.join()
half = samples[len(samples) // 2:]
print(json.dumps({"config": cfg, "env": {k2: v for k2, v in os.environ.items() if k2.startswith("HG_B200_TC")}, "reps": reps,
                  "fused_kernel_ms": ms.value / reps, "sm_mhz_median_second_half": statistics.median(s[0] for s in half),
                  "sm_mhz_min": min(s[0] for s in half), "power_w_median": statistics.median(s[1] for s in half),
                  "power_w_max": max(s[1] for s in half), "power_limit_w": pynvml.nvmlDeviceGetEnforcedPowerLimit(h) / 1e3,
                  "throttle_reason_bits_seen": sorted({hex(s[2]) for s in half})}))
