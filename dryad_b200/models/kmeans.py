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
