"""Iterative k-means workload (BASELINE.json config 5: points x 128 dims, k = 1024)."""
from __future__ import annotations

import torch


def generate(n: int, k: int, seed: int, device, rank: int = 0, spread: float = 0.35):
    """Synthetic clustered points: k Gaussian blobs in 128-d (shared centres, per-rank points)."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    centres = torch.randn((k, 128), dtype=torch.float32, device=device, generator=g)
    g.manual_seed(seed * 31 + rank + 1)
    which = torch.randint(0, k, (n,), device=device, generator=g)
    pts = torch.empty((n, 128), dtype=torch.float32, device=device)
    chunk = 1 << 22
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        pts[s:e] = centres[which[s:e]] + spread * torch.randn((e - s, 128), dtype=torch.float32, device=device, generator=g)
    return pts, centres


def reference_iteration(points: torch.Tensor, centroids: torch.Tensor, chunk: int = 1 << 18):
    """Plain PyTorch (cuBLAS) iteration: the baseline and, in fp64 on small inputs, the oracle."""
    k = centroids.shape[0]
    cn = (centroids * centroids).sum(1)
    sums = torch.zeros_like(centroids)
    counts = torch.zeros(k, dtype=torch.int64, device=points.device)
    assign = torch.empty(points.shape[0], dtype=torch.int64, device=points.device)
    for s in range(0, points.shape[0], chunk):
        x = points[s:s + chunk]
        d = cn[None, :] - 2.0 * (x @ centroids.T)
        a = d.argmin(1)
        assign[s:s + chunk] = a
        sums.index_add_(0, a, x)
        counts += torch.bincount(a, minlength=k)
    new_c = torch.where(counts[:, None] > 0, sums / counts.clamp(min=1)[:, None].to(sums.dtype), centroids)
    return assign, sums, counts, new_c


class KMeansStep:
    """One Lloyd iteration as a DEVICE function for `apply_per_partition`: called with this
    partition's points (column "x": [n, 128] fp32) and the current centroids (column "c": [k, 128]
    fp32, one replica per partition), returns the new centroids (identical bits on every partition).
    The kernels are the tcgen05 assign + certificate + fp32 re-check and the cross-GPU all-reduce of
    parallel/kmeans.py.  With `do_while` this is BASELINE.json config 5 expressed as a query
    (Apply loop; LinqToDryad/DryadLinqQueryable.cs:1280-1306):

        step = KMeansStep(ctx)
        cents = cents0.do_while(lambda cur: points.apply_per_partition(step, cur), cond)
    """

    device = True

    def __init__(self, context, exact: bool = True):
        self.context, self.exact = context, exact
        self._km = None
        self.iterations = 0

    def _engine(self, k: int, n: int, dev):
        from ..parallel.kmeans import KMeans
        from ..parallel.symm import SymmetricArena
        km = self._km
        if km is None or km.k != k or km.assign.numel() < n:
            ctx = self.context
            arena = ctx._spmd_arena() if ctx.world > 1 else SymmetricArena(4 << 20, 0, 1, None, dev)
            km = self._km = KMeans(arena, k, n, exact=self.exact)
        return km

    def __call__(self, points, cents):
        x = points.cols["x"]
        c = cents.cols["c"]
        km = self._engine(c.shape[0], x.shape[0], x.device)
        km.set_centroids(c.to(torch.float32))
        km.iterate(x.contiguous())
        self.iterations += 1
        return {"c": km.centroids.clone()}
