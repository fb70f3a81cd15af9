import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.models import kmeans as KM
from dryad_b200.parallel.kmeans import KMeans
from dryad_b200.parallel.symm import SymmetricArena
n = 8_000_000
pts, centres = KM.generate(n, 1024, 5, "cuda")
for pair in ("0", "1"):
    for dbg in ("0", "1"):
        os.environ["DRYAD_KM_PAIR"] = pair
        os.environ["DRYAD_KM_DEBUG"] = dbg
        km = KMeans(SymmetricArena(16 << 20), 1024, n)
        km.set_centroids(centres + 0.05)
        for _ in range(2): km.iterate(pts)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): km.iterate(pts)
        e1.record(); torch.cuda.synchronize()
        print("pair", pair, "skip_reds", dbg, "ms/iter", round(e0.elapsed_time(e1) / 5, 3), flush=True)
