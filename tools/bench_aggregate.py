"""Row (f)4 timing: super-patch aggregation in one device call (hg_aggregate_super_patches) next to the
torch formulation it replaced (index_add_ means + one gather / reduction per cluster with a host
read-back each, the reference's loop shape at preprocess_hypergraph.py:157-184).
    python tools/bench_aggregate.py"""
import json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
import hg_b200
from hg_b200 import ops, synth


def torch_loop(feats, pos, labels, s, lh, lg, K):
    counts = torch.bincount(labels, minlength=s)
    inv = (1.0 / counts.float()).unsqueeze(1)
    sf = torch.zeros(s, feats.shape[1], device=feats.device).index_add_(0, labels, feats) * inv
    sp = torch.zeros(s, pos.shape[1], device=feats.device).index_add_(0, labels, pos) * inv
    order = torch.argsort(labels, stable=True)
    bounds = torch.cumsum(counts, 0).tolist()
    intra, start = [], 0
    for c in range(s):
        idx = order[start:bounds[c]]
        start = bounds[c]
        m = idx.numel()
        if m > 1:
            blk = K[idx][:, idx] if K is not None else ops.similarity_dense(
                feats[idx], pos_a=pos[idx], pos_b=pos[idx], lambda_h=lh, lambda_g=lg, device=feats.device)
            intra.append(float((blk.sum() - blk.diagonal().sum()) / (m * (m - 1))))
    return sf, sp, float(np.mean(intra))


def wall(fn, reps):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(reps):
        out = fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t) / reps * 1e3, out


for n, d, s, dense in [(4096, 1024, 100, True), (16384, 1024, 100, True), (50_000, 1024, 100, False),
                       (200_000, 1024, 400, False), (50_000, 1024, 5000, False)]:
    feats_np, pos_np = synth.make_slide(n, d, 5, "gaussian")
    feats, pos = torch.from_numpy(feats_np).cuda(), torch.from_numpy(pos_np).cuda()
    labels = torch.from_numpy(np.random.default_rng(1).permutation(np.arange(n) % s)).cuda()
    lh, lg = 1.0 / d, 1.0 / 4096.0
    K = hg_b200.compute_combined_similarity(feats, pos, lh, lg) if dense else None

    def ours():
        agg = ops.aggregate_super_patches(feats, pos, labels, s, lh, lg, dense_k=K)
        return agg["features"], agg["positions"], agg["summary"].tolist()[0]
    ms_new, a = wall(ours, 10)
    ms_old, b = wall(lambda: torch_loop(feats, pos, labels, s, lh, lg, K), 2)
    pairs = float((torch.bincount(labels).double() ** 2).sum())
    print(json.dumps({"n": n, "d": d, "clusters": s, "dense_k": dense, "ms_device_call": ms_new, "ms_torch_loop": ms_old,
                      "speedup": ms_old / ms_new, "avg_intra": a[2], "avg_intra_torch": b[2],
                      "max_feat_diff": float((a[0] - b[0]).abs().max()),
                      "same_cluster_pairs": pairs,
                      "tile_gflops": None if dense else 2 * pairs * d / ms_new / 1e6}), flush=True)
