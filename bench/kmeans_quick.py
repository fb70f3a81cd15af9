"""k-means iteration times (exact / TF32-only / fused certificate) on one GPU; prints one JSON line."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.models import kmeans as KM  # noqa: E402
from dryad_b200.parallel.kmeans import KMeans  # noqa: E402
from dryad_b200.parallel.symm import SymmetricArena  # noqa: E402

n, k = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000, 1024
pts, centres = KM.generate(n, k, 5, "cuda")
res = {"n": n, "k": k}
for name, exact in (("exact", True), ("tf32_only", False)):
    km = KMeans(SymmetricArena(8 << 20), k, n, exact=exact)
    km.set_centroids(centres + 0.05)
    for _ in range(3):
        km.iterate(pts)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        km.iterate(pts)
    e1.record()
    torch.cuda.synchronize()
    res[name + "_ms"] = round(e0.elapsed_time(e1) / 10, 3)
    if exact:
        res["queued"], res["changed"] = km.ucount.tolist()
print(json.dumps(res))
