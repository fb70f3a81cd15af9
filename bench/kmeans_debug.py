import os, sys, torch
sys.path.insert(0, "/root/repo")
from dryad_b200.models import kmeans as KM
from dryad_b200.parallel.kmeans import KMeans
from dryad_b200.parallel.symm import SymmetricArena
pts, centres = KM.generate(8_000_000, 1024, 5, "cuda")
km = KMeans(SymmetricArena(16 << 20), 1024, 8_000_000)
km.set_centroids(centres + 0.05)
for _ in range(2): km.iterate(pts)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): km.iterate(pts)
e1.record(); torch.cuda.synchronize()
print("DBG", os.environ.get("DRYAD_KM_DEBUG"), "ms/iter", e0.elapsed_time(e1) / 3)
