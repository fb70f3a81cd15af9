"""ncu target 3: one GroupBy step (bucketised table), one fused dimension-join step and two k-means
iterations with the CTA-pair kernel (DRYAD_KM_PAIR=1)."""
import os
import sys

import torch

os.environ.setdefault("DRYAD_KM_PAIR", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.models import groupby as GB, joins as J, kmeans as KM  # noqa: E402
from dryad_b200.parallel.kmeans import KMeans  # noqa: E402
from dryad_b200.parallel.symm import SymmetricArena  # noqa: E402

dev = torch.device("cuda", 0)
sync = torch.cuda.synchronize
g = GB.bench(0, 1, dev, 40_000_000, 1 << 20, steps=1, warmup=1, max_over_ranks=lambda x: x, barrier=sync)
r = J.bench(0, 1, dev, 20_000_000, 1 << 20, steps=1, warmup=1, max_over_ranks=lambda x: x, barrier=sync)
pts, centres = KM.generate(2_000_000, 1024, 5, dev)
km = KMeans(SymmetricArena(16 << 20), 1024, 2_000_000)
km.set_centroids(centres + 0.05)
for _ in range(2):
    km.iterate(pts)
sync()
print("ok", r["valid"], g["valid"])
