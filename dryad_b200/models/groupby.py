"""GroupBy-Reduce workload: decomposable Sum + Count per key over synthetic (key, value) rows
(BASELINE.json config 3: 1 B rows over 8 GPUs = 125 M rows per GPU)."""
from __future__ import annotations

import torch


def generate(n: int, nkeys: int, seed: int, device, rank: int = 0):
    g = torch.Generator(device=device)
    g.manual_seed(seed * 7919 + rank)
    keys = torch.randint(0, nkeys, (n,), dtype=torch.int64, device=device, generator=g)
    vals = torch.randint(0, 1000, (n,), dtype=torch.int64, device=device, generator=g)
    return keys, vals


def bench(rank, world, dev, n, nkeys, steps, warmup, max_over_ranks, barrier):
    from ..ops import core as K
    from ..parallel.dagg import DistributedHashAggregate
    from ..parallel.symm import SymmetricArena
    keys, vals = generate(n, nkeys, 11, dev, rank)
    # owner tables hold ~nkeys/world groups (1.25x for hash imbalance), the local partial table all nkeys
    fcap = K.agg_capacity(int(1.25 * nkeys) // world + 1024)
    arena = SymmetricArena(fcap * 8 * 3 + (fcap // 3 + 1024) * world * 24 + (16 << 20), rank, world, None, dev)
    agg = DistributedHashAggregate(arena, ["sum_i64", "count"], K.agg_capacity(nkeys), fcap)
    for _ in range(max(warmup, 1)):
        gk, gl = agg.run(keys, [vals, None])
    barrier()
    l0 = K.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        gk, gl = agg.run(keys, [vals, None])
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1)) / steps
    # correctness: total count == rows, total sum matches
    tot = torch.stack([gl[1].sum(), gl[0].sum(), vals.sum(), torch.tensor(gk.numel(), device=dev)])
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(tot)
    ok = int(tot[0]) == world * n and int(tot[1]) == int(tot[2])
    return {"metric": "groupby_sum_count_rows_per_s", "value": world * n / (ms * 1e-3), "unit": "rows/s",
            "ms_per_step": ms, "rows_per_gpu": n, "distinct_keys": nkeys, "groups_out": int(tot[3]),
            "valid": bool(ok), "gpu_launches": K.launch_count() - l0}
