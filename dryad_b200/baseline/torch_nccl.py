"""PyTorch + NCCL baselines for BASELINE.json configs 2-5: `all_to_all_single`
shuffle, `torch.sort`, `index_add_` / `bincount` aggregation, `all_reduce`, `torch.matmul`.

Run under torchrun:  python -m dryad_b200.baseline.torch_nccl --gpus N [--rows-per-gpu R]
Prints one JSON line on rank 0 with device-timed (CUDA events, max over ranks) numbers.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist


def _a2a_rows(x: torch.Tensor, dest: torch.Tensor, world: int):
    """Shuffle rows of x (any trailing shape) to rank dest[i] with NCCL all_to_all_single."""
    order = torch.argsort(dest, stable=True)
    xs = x[order]
    counts = torch.bincount(dest, minlength=world)
    if world == 1:
        return xs
    recv_counts = torch.empty_like(counts)
    dist.all_to_all_single(recv_counts, counts)
    sc, rc = counts.tolist(), recv_counts.tolist()
    out = torch.empty((sum(rc),) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_to_all_single(out, xs, rc, sc)
    return out


def terasort_step(rows: torch.Tensor, world: int, rank: int):
    """Sample -> all_gather -> separators -> bucketize -> all_to_all -> torch.sort -> index."""
    n = rows.shape[0]
    k = rows[:, :8].to(torch.int64)
    hi = torch.zeros(n, dtype=torch.int64, device=rows.device)
    for b in range(8):
        hi = (hi << 8) | k[:, b]
    key = hi ^ (-0x8000000000000000)              # unsigned order as signed int64
    if world > 1:
        m = max(1024, n // 1000)
        idx = torch.randint(0, n, (m,), device=rows.device)
        samples = [torch.empty(m, dtype=torch.int64, device=rows.device) for _ in range(world)]
        dist.all_gather(samples, key[idx].contiguous())
        s = torch.sort(torch.cat(samples)).values
        seps = s[(torch.arange(1, world, device=rows.device) * (s.numel() // world))]
        dest = torch.bucketize(key, seps, right=False)
        rows = _a2a_rows(rows, dest, world)
        k = rows[:, :8].to(torch.int64)
        hi = torch.zeros(rows.shape[0], dtype=torch.int64, device=rows.device)
        for b in range(8):
            hi = (hi << 8) | k[:, b]
        key = hi ^ (-0x8000000000000000)
    lo = (rows[:, 8].to(torch.int64) << 8) | rows[:, 9].to(torch.int64)
    o1 = torch.argsort(lo, stable=True)             # LSD: low 2 bytes first, then the 8-byte prefix
    o2 = torch.argsort(key[o1], stable=True)
    return rows[o1[o2]]


def groupby_step(keys: torch.Tensor, vals: torch.Tensor, nkeys: int, world: int):
    """Local partial aggregate (index_add_/bincount on dense key ids) + all_reduce of the partials."""
    s = torch.zeros(nkeys, dtype=torch.int64, device=keys.device).index_add_(0, keys, vals)
    c = torch.bincount(keys, minlength=nkeys)
    if world > 1:
        dist.all_reduce(s)
        dist.all_reduce(c)
    return s, c


def groupby_hashed_step(keys: torch.Tensor, vals: torch.Tensor, world: int):
    """The same GroupBy for general (sparse 64-bit) keys, the way a PyTorch + NCCL user writes it:
    local partial aggregate (unique + index_add_/bincount), all_to_all of the partials by key hash,
    final aggregate (unique + index_add_).  Returns (keys, sums, counts) of the groups this rank owns."""
    uk, inv = torch.unique(keys, return_inverse=True)
    s = torch.zeros(uk.numel(), dtype=torch.int64, device=keys.device).index_add_(0, inv, vals)
    c = torch.bincount(inv, minlength=uk.numel())
    if world == 1:
        return uk, s, c
    dest = ((uk * -7046029254386353131) >> 40) % world       # multiplicative hash of the key bits
    part = _a2a_rows(torch.stack([uk, s, c], dim=1), dest, world)
    fk, finv = torch.unique(part[:, 0], return_inverse=True)
    fs = torch.zeros(fk.numel(), dtype=torch.int64, device=keys.device).index_add_(0, finv, part[:, 1])
    fc = torch.zeros(fk.numel(), dtype=torch.int64, device=keys.device).index_add_(0, finv, part[:, 2])
    return fk, fs, fc


def gemm_tf32_only(points: torch.Tensor, centroids: torch.Tensor, chunk: int = 1 << 20):
    """cuBLAS TF32 GEMM of the k-means distance shape and NOTHING else (no argmin, no sums): the
    library's time for the tensor-core part of one iteration, in row chunks so the product fits."""
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    ct = centroids.t().contiguous()
    out = torch.empty((min(chunk, points.shape[0]), centroids.shape[0]), dtype=torch.float32, device=points.device)
    for i in range(0, points.shape[0], chunk):
        blk = points[i:i + chunk]
        torch.matmul(blk, ct, out=out[:blk.shape[0]])
    torch.backends.cuda.matmul.allow_tf32 = prev
    return out


def join_step(fk, fv, dk, da, n_dim, world, threshold=500):
    """Broadcast join: all_gather the dimension table, filter, gather-by-key, multiply, sum."""
    if world > 1:
        parts_k = [torch.empty_like(dk) for _ in range(world)]
        parts_a = [torch.empty_like(da) for _ in range(world)]
        dist.all_gather(parts_k, dk)
        dist.all_gather(parts_a, da)
        dk, da = torch.cat(parts_k), torch.cat(parts_a)
    table = torch.empty(n_dim, dtype=torch.int64, device=fk.device)
    table[dk] = da
    sel = fv >= threshold
    return (fv[sel] * table[fk[sel]]).sum()


def kmeans_step(points, centroids, world):
    from ..models.kmeans import reference_iteration
    _, sums, counts, _ = reference_iteration(points, centroids)
    if world > 1:
        dist.all_reduce(sums)
        dist.all_reduce(counts)
    return torch.where(counts[:, None] > 0, sums / counts.clamp(min=1)[:, None].to(sums.dtype), centroids)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--rows-per-gpu", type=int, default=40_000_000)
    ap.add_argument("--groupby-rows-per-gpu", type=int, default=125_000_000)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
    from dryad_b200.models import groupby as GB, joins as J, kmeans as KM, terasort as TS

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn):
        for _ in range(a.warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            fn()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1) / a.steps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    res = {"impl": "pytorch+nccl baseline", "n_gpus": world}
    n = a.rows_per_gpu
    rows = TS.generate_rows(n, 7, dev, rank)
    ms = timed(lambda: terasort_step(rows, world, rank))
    res["terasort_ms"] = ms
    res["terasort_GBs"] = world * n * 100 / (ms * 1e-3) / 1e9
    del rows
    torch.cuda.empty_cache()
    g = a.groupby_rows_per_gpu
    keys, vals = GB.generate(g, 1 << 20, 11, dev, rank)
    ms = timed(lambda: groupby_step(keys, vals, 1 << 20, world))
    res["groupby_ms"] = ms
    res["groupby_rows_per_s"] = world * g / (ms * 1e-3)
    del keys, vals
    torch.cuda.empty_cache()
    fk, fv, dk, da, _ = J.generate(50_000_000, 1 << 20, 13, dev, rank, world)
    ms = timed(lambda: join_step(fk, fv, dk, da, 1 << 20, world))
    res["broadcast_join_ms"] = ms
    res["broadcast_join_rows_per_s"] = world * 50_000_000 / (ms * 1e-3)
    del fk, fv
    torch.cuda.empty_cache()
    torch.backends.cuda.matmul.allow_tf32 = True
    pts, centres = KM.generate(8_000_000, 1024, 5, dev, rank)
    c0 = centres + 0.05
    ms = timed(lambda: kmeans_step(pts, c0, world))
    res["kmeans_iter_ms"] = ms
    if rank == 0:
        print(json.dumps(res))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
