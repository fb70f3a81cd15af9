"""ncu target 4: k-means iterations only (CTA-pair kernel)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.models import kmeans as KM  # noqa: E402
from dryad_b200.parallel.kmeans import KMeans  # noqa: E402
from dryad_b200.parallel.symm import SymmetricArena  # noqa: E402

dev = torch.device("cuda", 0)
pts, centres = KM.generate(2_000_000, 1024, 5, dev)
km = KMeans(SymmetricArena(16 << 20), 1024, 2_000_000)
km.set_centroids(centres + 0.05)
for _ in range(3):
    km.iterate(pts)
torch.cuda.synchronize()
print("ok")
