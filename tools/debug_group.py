import os, sys
os.environ["CUDA_LAUNCH_BLOCKING"] = "1"
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
import hg_b200
from hg_b200 import synth
from hg_b200.pipeline import build_batch
sizes = [int(a) for a in sys.argv[1:]] or [300, 5000, 2049]
slides = [torch.from_numpy(synth.make_features(s, 128, 100 + i, "gaussian")).cuda() for i, s in enumerate(sizes)]
kw = dict(k=8, kernel="gaussian", lambda_h=1.0, return_csr=True)
refs = [hg_b200.knn_hyperedges(x, **kw) for x in slides]
torch.cuda.synchronize()
print("per-slide builds ok", flush=True)
try:
    outs = build_batch(slides, lanes=2, grouped=False, **kw)
    torch.cuda.synchronize()
    print("ungrouped batch ok", flush=True)
    outs = build_batch(slides, lanes=2, grouped=True, **kw)
    torch.cuda.synchronize()
    for x, o in zip(slides, outs):
        ref = hg_b200.knn_hyperedges(x, **kw)
        print(x.shape[0], "knn equal", bool((o["knn_idx"] == ref["knn_idx"]).all()), "edges", o["edge_index"].shape, ref["edge_index"].shape,
              bool(o["edge_index"].shape == ref["edge_index"].shape and (o["edge_index"] == ref["edge_index"]).all()))
except Exception as e:
    print("FAILED:", e)
