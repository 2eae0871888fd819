import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import synth
from hg_b200.pipeline import SlidePipeline
dev = torch.device("cuda", 0)
x_host = torch.from_numpy(synth.make_features(50000, 1024, 1236, "gaussian")).pin_memory()
pipe = SlidePipeline(dev, k=16, kernel="gaussian", lambda_h=1.0, return_csr=True)
def stat():
    s = torch.cuda.memory_stats(dev)
    return s["num_device_alloc"], s["reserved_bytes.all.current"] >> 20
for rep in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter(); ts = []
    for i in range(8):
        a = time.perf_counter()
        pipe.submit(x_host)
        b = time.perf_counter()
        if len(pipe._done) >= 2:
            next(pipe.results())
        c = time.perf_counter()
        ts.append((round((b - a) * 1e3, 2), round((c - b) * 1e3, 2), stat()))
    for _ in pipe.results(flush=True):
        pass
    torch.cuda.synchronize()
    print("staged" if os.environ.get("HG_B200_STAGED") else "fused", "rep", rep, "total ms/step", (time.perf_counter() - t0) * 1e3 / 8)
    print(ts)
