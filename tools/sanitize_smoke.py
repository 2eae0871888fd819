"""Small-size pass over every kernel family for compute-sanitizer (memcheck / racecheck):
    compute-sanitizer --tool memcheck python tools/sanitize_smoke.py
Sizes are tiny on purpose (the tool slows kernels 10-100x); results are still checked."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import hg_b200
import oracle_np as oracle
from hg_b200 import ops, synth
from hg_b200.pipeline import build_batch

torch.cuda.set_device(0)
n, d, k = 700, 96, 8
feats, pos = synth.make_slide(n, d, 3, "gaussian")
x, p = torch.from_numpy(feats).cuda(), torch.from_numpy(pos).cuda()
kw = dict(k=k, kernel="gaussian", lambda_h=1.0, lambda_g=1.0 / 128, spatial_k=4, return_csr=True)
os.environ["HG_B200_SP_GRID"] = "1"                       # grid path at a size the tool can afford
out = hg_b200.knn_hyperedges(x, p, **kw)
os.environ["HG_B200_SP_GRID"] = "0"
ref = hg_b200.knn_hyperedges(x, p, **kw)
for key in out:
    assert torch.equal(out[key], ref[key]), key
os.environ.pop("HG_B200_SP_GRID")
ei = oracle.edges_from_knn(np.concatenate([out["knn_idx"].cpu().numpy(), out["spatial_idx"].cpu().numpy()], 1))
assert (out["edge_index"].cpu().numpy() == ei).all()
rowptr, col, eid = oracle.csr_from_coo(ei, n)
assert (out["csr_rowptr"].cpu().numpy() == rowptr).all() and (out["csr_col"].cpu().numpy() == col).all()
os.environ["HG_B200_TC_SYM"] = "4"                        # symmetric tensor-core pass, forced
sym = ops.knn_feature(x, k, "gaussian")
os.environ["HG_B200_TC_SYM"] = "0"
full = ops.knn_feature(x, k, "gaussian")
os.environ.pop("HG_B200_TC_SYM")
assert torch.equal(sym[0], full[0])
slides = [torch.from_numpy(synth.make_features(s, 64, 9 + s, "gaussian")).cuda() for s in (300, 520, 257, 400)]
outs = build_batch(slides, lanes=2, k=5, kernel="euclidean", return_csr=True)
for s, o in zip(slides, outs):
    r = hg_b200.knn_hyperedges(s, k=5, kernel="euclidean", return_csr=True)
    for key in ("edge_index", "csr_rowptr", "csr_col", "csr_eid"):
        assert torch.equal(o[key], r[key]), key
xf = torch.randn(n, 20, device="cuda")
prop = ops.hypergraph_propagate(xf, out["csr_rowptr"], out["csr_col"])
np.testing.assert_allclose(prop.cpu().numpy(), oracle.hypergraph_conv_propagate(xf.cpu().numpy(), ei, n)[0], rtol=1e-5, atol=1e-6)
labels = torch.from_numpy(np.random.default_rng(0).integers(0, 12, n).astype(np.int32))
cl = hg_b200.knn_hyperedges(x, k=4, kernel="cosine", labels=labels, num_hyperedges=6, return_csr=True)
K = hg_b200.compute_combined_similarity(x[:300], p[:300], 1.0, 1.0 / 128)
ops.similarity_stats(K)
hg_b200.build_weighted_hypergraph(x[:200], p[:200], 1.0, 1.0 / 128, 1.0)
from hg_b200.kmeans import kmeans
for kn, kd, kc in [(700, 96, 7), (300, 33, 40)]:
    lab, cen, inertia = kmeans(x[:kn, :kd].contiguous(), kc, n_init=2, max_iter=20)
    assert lab.min() >= 0 and lab.max() < kc and np.isfinite(inertia)
torch.cuda.synchronize()
print("sanitize_smoke ok")
