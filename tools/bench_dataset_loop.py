"""process_dataset over a folder of synthetic .npz slide containers: the reference's sequential loop
(load -> build -> save, HG_B200_IO_OVERLAP=0) against the read-ahead / writer-thread loop.
    python tools/bench_dataset_loop.py [N_wsi] [slides]"""
import json, os, sys, tempfile, time, contextlib, io
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
import hg_b200
from hg_b200 import preprocess_hypergraph as pp, store, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
slides = int(sys.argv[2]) if len(sys.argv) > 2 else 12
d, n_tma = 1024, 256
pp.KMEANS_BACKEND = os.environ.get("HG_B200_KMEANS", "cuda")
tmp = tempfile.mkdtemp(prefix="hg_b200_loop_")
names = []
for i in range(slides):
    feats, pos = synth.make_slide(n, d, 300 + i, "gaussian")
    name = f"slide_{i}.npz"
    with store.open_store(os.path.join(tmp, name), "w") as f:
        g = f.create_group("wsi")
        g["features"] = feats
        g["positions"] = pos
        f.create_group("tma")["features"] = synth.make_features(n_tma, d, 400 + i, "gaussian")
    names.append(name)
csv = os.path.join(tmp, "cases.csv")
with open(csv, "w") as f:
    f.write("case_id,h5_file_path\n" + "\n".join(f"c{i},{nm}" for i, nm in enumerate(names)) + "\n")
for overlap in (False, True, False, True):
    pp.IO_OVERLAP = overlap
    torch.cuda.synchronize()
    t = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
        res = hg_b200.process_dataset(csv, tmp, num_wsi_super_patches=100, num_groups=10, hypergraph_k=5, num_hyperedges=10,
                                      lambda_h=1.0 / d, lambda_g=1.0 / 4096)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t
    assert len(res) == slides
    print(json.dumps({"n_wsi": n, "slides": slides, "kmeans": pp.KMEANS_BACKEND, "io_overlap": overlap,
                      "total_s": dt, "s_per_slide": dt / slides}), flush=True)
