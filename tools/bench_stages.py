"""Bandwidth of the HBM-bound stages (A3 spatial kNN, A5 edge assembly, CSR, operand split) at a size
that fills the machine -- at slide sizes (<= 100k patches) these kernels are a few microseconds long and
launch/latency-bound, so their achieved GB/s says nothing about the kernels themselves.
    python tools/bench_stages.py [N]            # default N = 4,000,000 nodes, k = 16
Algorithmic bytes follow SURVEY.md 8(d) / DESIGN.md section 3; peak = MEASURED_PEAKS.json hbm_gbs."""
import json, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import ops, synth, _lib

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
k = 16
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
dev = torch.device("cuda", 0)
lib = _lib.load_library()


def timed(fn, reps=5):
    for _ in range(2):
        out = fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


def emit(stage, ms, nbytes, **kw):
    gbs = nbytes / ms / 1e6
    print(json.dumps(dict(stage=stage, n=n, ms=round(ms, 4), algorithmic_bytes=int(nbytes), achieved_GBps=round(gbs, 1),
                          frac_of_measured_hbm_peak=round(gbs / peak, 3), **kw)), flush=True)


g = torch.Generator(device=dev).manual_seed(1)
# neighbour lists shaped like kNN output: every row points into a window around itself (no self, no duplicates)
offs = torch.arange(1, k + 1, device=dev, dtype=torch.int64) * 37
rows = torch.arange(n, device=dev, dtype=torch.int64).unsqueeze(1)
sign = torch.where(torch.rand(n, k, device=dev, generator=g) < 0.5, -1, 1)
idx = ((rows + sign * offs) % n).to(torch.int32).contiguous()
ms, ei = timed(lambda: ops.build_edges([idx], n))
e = ei.shape[1]
bits = (n * n).bit_length()
passes = (bits + 7) // 8
# emit 4 B read + 8 B written per slot; per radix pass keys read twice + written once; unique: read twice + write; decode 8 read + 16 written
alg = n * k * (12 + passes * 24 + 16) + e * 8 + e * 24
emit("A5 edge assembly (emit + radix sort + unique + decode)", ms, alg, edges=e, radix_passes=passes, slots=n * k)
ms, csr = timed(lambda: ops.csr_from_coo(ei, n))
p2 = (n.bit_length() + 7) // 8
alg = e * (16 + 12) + e * p2 * (24 + 8) + e * (12 + 16) + (n + 1) * 16 + 2 * e * 16
emit("CSR incidence (keys + stable sort by larger endpoint + bounds + fill)", ms, alg, edges=e, radix_passes=p2)
pos = torch.from_numpy(synth.make_positions(n, 5)).to(dev)
ms, _ = timed(lambda: ops.knn_spatial(pos, 8))
emit("A3 spatial kNN, cell grid (k = 8)", ms, n * (8 + 8 * 8), note="compulsory bytes only: positions in, lists out")
d = 1024
nn = min(n, 1_000_000)
x = torch.randn(nn, d, device=dev)
nbytes = lib.hg_operand_bytes(nn, d)
opsbuf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
norms = torch.empty(nn, device=dev)
st = torch.cuda.current_stream().cuda_stream
ms, _ = timed(lambda: _lib.check(lib.hg_split_operands(x.data_ptr(), nn, d, opsbuf.data_ptr(), norms.data_ptr(), st)))
emit("operand split FP32 -> BF16 hi/lo planes + norms", ms, nn * d * 8, rows=nn, d=d)
f = 256
xf = torch.randn(n, f, device=dev)
ms, _ = timed(lambda: ops.hypergraph_propagate(xf, csr[0], csr[1]))
# gathered rows are credited once per pass (the neighbour windows are L2-resident), as SURVEY.md 8(d) does for A6
emit("(f)3 propagation, F = 256 (two gather passes)", ms, 2 * (2 * n * f * 4) + 2 * 2 * e * 8, edges=e,
     l2_gather_bytes=2 * e * f * 4)
