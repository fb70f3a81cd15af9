"""Correctness + timing of the CTA-pair k-means kernel against the one-CTA kernel (same inputs).
Run each mode in its own process (the mode is read from DRYAD_KM_PAIR at call time)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.models import kmeans as KM  # noqa: E402
from dryad_b200.parallel.kmeans import KMeans  # noqa: E402
from dryad_b200.parallel.symm import SymmetricArena  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000
pts, centres = KM.generate(n, 1024, 5, "cuda")
out = {}
for mode, ew in (("0", "16"), ("1", "8")):
    os.environ["DRYAD_KM_PAIR"] = mode
    os.environ["DRYAD_KM_EPI_WARPS"] = ew
    km = KMeans(SymmetricArena(16 << 20), 1024, n)
    km.set_centroids(centres + 0.05)
    km.iterate(pts)
    torch.cuda.synchronize()
    out[mode + ew] = (km.assign[:n].clone(), km.centroids.clone(), km.counts.clone() if hasattr(km, "counts") else None)
    for _ in range(2):
        km.iterate(pts)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        km.iterate(pts)
    e1.record()
    torch.cuda.synchronize()
    print("pair ew=" + ew if mode == "1" else "single", "ms/iter", round(e0.elapsed_time(e1) / 5, 3), flush=True)
a0, c0, _ = out["016"]
for key in ("18",):
    a1, c1, _ = out[key]
    diff = int((a0 != a1).sum().item())
    print(key, "assignments differing:", diff, "of", n, " max centroid delta:", float((c0 - c1).abs().max().item()))
    assert diff <= n // 10000, "pair kernel disagrees with the single-CTA kernel"
