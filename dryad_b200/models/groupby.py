"""GroupBy-Reduce workload: decomposable Sum + Count per key over synthetic (key, value) rows
(BASELINE.json config 3: 1 B rows over 8 GPUs = 125 M rows per GPU)."""
from __future__ import annotations

import torch


def generate(n: int, nkeys: int, seed: int, device, rank: int = 0, sparse: bool = False):
    """(key, value) columns with `nkeys` distinct keys.  sparse=False: dense ids 0..nkeys-1 (dictionary
    codes).  sparse=True: the ids pushed through the 64-bit mixer (a bijection), i.e. general 64-bit
    keys spread over the whole range, which no dense array can index."""
    g = torch.Generator(device=device)
    g.manual_seed(seed * 7919 + rank)
    keys = torch.randint(0, nkeys, (n,), dtype=torch.int64, device=device, generator=g)
    vals = torch.randint(0, 1000, (n,), dtype=torch.int64, device=device, generator=g)
    if sparse:
        from ..ops import core as K
        keys = K.hash_u64(keys)
    return keys, vals


def validate_sample(rank, world, dev, keys, vals, gk, gl, nsample: int = 2048) -> bool:
    """Per-key check on a sample: for `nsample` key values drawn from the data, the (sum, count) this
    framework returned (each key lives on exactly one rank) equals a plain torch reduction over ALL
    ranks' rows of that key.  Catches a wrong key -> owner mapping that totals alone would not."""
    import torch.distributed as dist
    sample = keys[:: max(1, keys.numel() // nsample)][:nsample].clone()
    if world > 1:
        dist.broadcast(sample, 0)
    sample = torch.unique(sample)
    ns = sample.numel()
    sel = torch.isin(keys, sample)
    pos = torch.searchsorted(sample, keys[sel])
    want = torch.zeros((2, ns), dtype=torch.int64, device=dev)
    want[0].index_add_(0, pos, vals[sel])
    want[1].index_add_(0, pos, torch.ones_like(pos))
    got = torch.zeros((3, ns), dtype=torch.int64, device=dev)
    gsel = torch.isin(gk, sample)
    gpos = torch.searchsorted(sample, gk[gsel])
    got[0].index_add_(0, gpos, gl[0][gsel])
    got[1].index_add_(0, gpos, gl[1][gsel])
    got[2].index_add_(0, gpos, torch.ones_like(gpos))          # every sampled key on exactly ONE rank
    if world > 1:
        dist.all_reduce(want)
        dist.all_reduce(got)
    return bool(torch.equal(want, got[:2]) and bool((got[2] == 1).all()))


def bench(rank, world, dev, n, nkeys, steps, warmup, max_over_ranks, barrier, sparse: bool = False):
    from ..ops import core as K
    from ..parallel.dagg import DistributedHashAggregate
    from ..parallel.symm import SymmetricArena
    keys, vals = generate(n, nkeys, 11, dev, rank, sparse=sparse)
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
    ok = int(tot[0]) == world * n and int(tot[1]) == int(tot[2]) and 0 < int(tot[3]) <= nkeys
    ok = ok and validate_sample(rank, world, dev, keys, vals, gk, gl)
    return {"metric": "groupby_sum_count_rows_per_s", "value": world * n / (ms * 1e-3), "unit": "rows/s",
            "ms_per_step": ms, "rows_per_gpu": n, "distinct_keys": nkeys, "groups_out": int(tot[3]),
            "keys": "sparse 64-bit (mix64 of ids): hashed table" if sparse else "dense ids: direct-addressed table",
            "valid": bool(ok), "validation": "totals + per-key sum/count on a 2048-key sample against torch",
            "gpu_launches": K.launch_count() - l0}


def bench_api(rank, world, dev, n, nkeys, steps, warmup, max_over_ranks, barrier):
    """The same GroupBy through the PUBLIC API, end to end: every step copies this rank's (key, value)
    columns from pinned host memory, runs `from_columns(...).group_by(k).select(sum, count)` on the
    GPU backend (SPMD over `world` ranks: partial aggregate -> fused hash shuffle -> final
    aggregate) and reads the number of groups and the total count back."""
    import dryad_b200 as D
    keys, vals = generate(n, nkeys, 11, dev, rank)
    hk = torch.empty(n, dtype=torch.int64).pin_memory(); hk.copy_(keys)
    hv = torch.empty(n, dtype=torch.int64).pin_memory(); hv.copy_(vals)
    del keys, vals
    ctx = D.DryadLinqContext(world, platform="B200" if world > 1 else "LOCAL", device=str(dev))
    ctx.group_capacity_hint = max(ctx.group_capacity_hint, 2 * nkeys)

    def step():
        k = hk.to(dev, non_blocking=True)
        v = hv.to(dev, non_blocking=True)
        cols = ctx.from_columns({"k": k, "v": v}).group_by(lambda r: r.k).select(
            lambda g: {"k": g.key, "s": g.sum(lambda r: r.v), "c": g.count()}).to_columns()
        if not cols:
            return 0, 0
        return int(cols["k"].numel()), int(cols["c"].sum().item())          # D2H: 16 bytes

    for _ in range(max(1, min(warmup, 2))):
        g, c = step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        g, c = step()
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1)) / steps
    tot = torch.tensor([g, c], dtype=torch.int64, device=dev)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(tot)
    res = {"value": world * n / (ms * 1e-3), "unit": "rows/s", "ms_per_step": ms, "h2d_bytes_per_step": 16 * n,
           "d2h_bytes_per_step": 16, "groups_out": int(tot[0]), "valid": int(tot[1]) == world * n,
           "host_fallbacks": ctx._backend().host_fallbacks}
    ctx.dispose()
    return res
