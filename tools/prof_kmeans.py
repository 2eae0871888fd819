"""One restart of the GPU KMeans at a given size, for a launch list:
    ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file out.csv python tools/prof_kmeans.py 50000 1024 100"""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
from hg_b200 import synth
from hg_b200.kmeans import kmeans
n, d, c = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (50000, 1024, 100)
x = torch.from_numpy(synth.make_features(n, d, 80, "gaussian")).cuda()
labels, centers, inertia = kmeans(x, c, n_init=1)
torch.cuda.synchronize()
print("inertia", inertia)
