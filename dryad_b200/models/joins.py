"""Join workload (BASELINE.json config 4): large fact table x small dimension table with a
Where/Select pipeline — `fact.where(v >= t).join(dim, k).select(v * attr)` reduced to a checksum."""
from __future__ import annotations

import torch


def generate(n_fact: int, n_dim: int, seed: int, device, rank: int = 0, world: int = 1):
    g = torch.Generator(device=device)
    g.manual_seed(seed * 104729 + rank)
    fk = torch.randint(0, n_dim, (n_fact,), dtype=torch.int64, device=device, generator=g)
    fv = torch.randint(0, 1000, (n_fact,), dtype=torch.int64, device=device, generator=g)
    per = (n_dim + world - 1) // world
    lo, hi = rank * per, min(n_dim, (rank + 1) * per)
    dk = torch.arange(lo, hi, dtype=torch.int64, device=device)
    da = (dk * 2654435761) % 97
    return fk, fv, dk, da, lo


def bench(rank, world, dev, n_fact, n_dim, steps, warmup, max_over_ranks, barrier, threshold=500):
    import os
    from ..expr import col, gather, probe_slot, where
    from ..expr.compile import run_on_gpu
    from ..ops import core as K
    from ..parallel.djoin import BroadcastJoin
    from ..parallel.symm import SymmetricArena
    fk, fv, dk, da, base = generate(n_fact, n_dim, 13, dev, rank, world)
    arena = SymmetricArena(K._next_pow2(2 * n_dim + 64) * 12 + n_dim * 12 + (16 << 20), rank, world, None, dev)
    bj = BroadcastJoin(arena, n_dim, 1)

    fused = os.environ.get("DRYAD_JOIN_FUSED", "1") != "0"

    def step():
        bj.build(dk, [da], base)                                      # fused broadcast + build
        if fused:
            # probe + payload gather + Where + Select + Sum in ONE generated kernel: the fact side is
            # read once and nothing per-row is written (expr.probe / expr.gather, expr/codegen.py)
            # the payload column is permuted into slot order once per build (1 Mi elements), so the
            # probe kernel touches two L2 sectors per fact row (key bucket, payload) instead of three
            attr_slot = bj.slot_payload(bj.cols[0])
            hit = probe_slot("dim", col("k"))
            keep = (col("v") >= threshold) & (hit >= 0)
            out, cnt = run_on_gpu({"r": where(keep, col("v") * gather("attr", hit), 0)}, {"k": fk, "v": fv}, n_fact,
                                  predicate=keep, tables={"dim": bj}, arrays={"attr": attr_slot},
                                  reduce={"r": "sum"}, count_flags=True)
            return out["r"], cnt
        di, matched = bj.probe_unique(fk)                             # local probe: the fact side never moves
        a = K.gather_rows(bj.cols[0], di)
        # Where(v >= t) and Select(v * attr) fused into one register program; unmatched rows give 0
        out, flags = run_on_gpu({"r": where((col("v") >= threshold) & (col("m") != 0), col("v") * col("a"), 0)},
                                {"v": fv, "a": a, "m": matched}, n_fact, predicate=(col("v") >= threshold) & (col("m") != 0))
        return K.reduce(out["r"], "sum"), K.reduce(flags, "sum")

    for _ in range(max(warmup, 1)):
        s, m = step()
    barrier()
    l0 = K.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        s, m = step()
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1)) / steps
    sel = fv >= threshold
    ref = int((fv[sel] * ((fk[sel] * 2654435761) % 97)).sum().item())
    m = int(m.item())
    ok = int(s.item()) == ref and m == int(sel.sum().item())
    return {"metric": "broadcast_join_fact_rows_per_s", "value": world * n_fact / (ms * 1e-3), "unit": "rows/s",
            "ms_per_step": ms, "fact_rows_per_gpu": n_fact, "dim_rows": n_dim, "matches_local": m, "valid": bool(ok), "fused_probe_pipeline": fused,
            "gpu_launches": K.launch_count() - l0}


def bench_shuffle_join(rank, world, dev, n_fact, n_dim_per_gpu, steps, warmup, max_over_ranks, barrier):
    """Hash join where NEITHER side fits a broadcast: fact (key, value) and a dimension table with
    `n_dim_per_gpu` unique keys per GPU are both hash-partitioned to the key's owner with the fused
    shuffle, joined there, and Select(v * attr) is reduced to a checksum."""
    from ..ops import core as K
    from ..parallel.djoin import ShuffleHashJoin
    from ..parallel.symm import SymmetricArena
    n_dim = n_dim_per_gpu * world
    g = torch.Generator(device=dev)
    g.manual_seed(977 + rank)
    fk = K.hash_u64(torch.randint(0, n_dim, (n_fact,), dtype=torch.int64, device=dev, generator=g))   # sparse 64-bit keys
    fv = torch.randint(0, 1000, (n_fact,), dtype=torch.int64, device=dev, generator=g)
    ids = torch.arange(rank * n_dim_per_gpu, (rank + 1) * n_dim_per_gpu, dtype=torch.int64, device=dev)
    dk = K.hash_u64(ids)
    da = (ids * 2654435761) % 97
    capl, capr = int(n_fact * 1.15) + 65536, int(n_dim_per_gpu * 1.15) + 65536
    arena = SymmetricArena((capl + capr) * 16 + world * 8192 * 16 + (64 << 20), rank, world, None, dev)
    shj = ShuffleHashJoin(arena, 2, 2, capl, capr, n_fact, n_dim_per_gpu)

    def step():
        lc, rc, rows, matched = shj.run_unique([fk, fv], [dk, da])
        attr = K.gather_rows(rc[1], rows.clamp_(min=0))
        return (lc[1] * attr * matched).sum(), matched.sum(dtype=torch.int64)

    for _ in range(max(warmup, 1)):
        s_, m_ = step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        s_, m_ = step()
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1)) / steps
    # reference: attr is a function of the id, which the mixer (a bijection) hides from the join
    idsf = torch.randint(0, n_dim, (n_fact,), dtype=torch.int64, device=dev, generator=torch.Generator(device=dev).manual_seed(977 + rank))
    want = (fv * ((idsf * 2654435761) % 97)).sum()
    tot = torch.stack([s_, m_, want])
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(tot)
    ok = int(tot[0]) == int(tot[2]) and int(tot[1]) == world * n_fact
    return {"metric": "shuffle_hash_join_fact_rows_per_s", "value": world * n_fact / (ms * 1e-3), "unit": "rows/s",
            "ms_per_step": ms, "fact_rows_per_gpu": n_fact, "dim_rows_per_gpu": n_dim_per_gpu, "valid": bool(ok)}
