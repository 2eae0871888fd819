"""A1 / A2 dense kernels: FP32 SIMT tiles against the tensor-core dense-output mode.   python tools/bench_dense.py"""
import json, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import ops, synth
for n, d in [(4096, 1024), (8192, 1024), (16384, 1024), (32768, 1024)]:
    feats, pos = synth.make_slide(n, d, 50, "gaussian")
    x, p = torch.from_numpy(feats).cuda(), torch.from_numpy(pos).cuda()
    row = {"n": n, "d": d}
    for name, min_n in (("simt_ms", 0), ("tc_ms", 1)):
        ops.DENSE_TC_MIN_N = min_n
        for _ in range(2):
            K = ops.similarity_dense(x, pos_a=p, pos_b=p, lambda_h=1.0, lambda_g=1.0 / 4096)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 3
        e0.record()
        for _ in range(reps):
            K = ops.similarity_dense(x, pos_a=p, pos_b=p, lambda_h=1.0, lambda_g=1.0 / 4096)
        e1.record(); torch.cuda.synchronize()
        row[name] = e0.elapsed_time(e1) / reps
        del K
    row["speedup"] = row["simt_ms"] / row["tc_ms"]
    row["tc_tflops_algorithmic"] = 2.0 * n * n * d / row["tc_ms"] / 1e9
    row["tc_output_GBps"] = 4.0 * n * n / row["tc_ms"] / 1e6
    print(json.dumps(row), flush=True)
