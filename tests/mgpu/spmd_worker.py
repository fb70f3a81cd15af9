"""Multi-rank worker (launched by torchrun / mp.spawn).  `--mode cpu` exercises the SPMD executor
on the host backend over gloo; `--mode gpu` checks the fused multi-GPU kernels over NVLink."""
import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import dryad_b200 as D  # noqa: E402


def cpu_checks(rank, world):
    ctx = D.DryadLinqContext(platform="B200", device="cpu")
    assert ctx.world == world and ctx.rank == rank
    data = list(range(200))
    q = ctx.from_enumerable(data, world)
    got = q.where(lambda x: x % 3 == 0).select(lambda x: x + 1).partitions()
    mine = [x + 1 for x in data[rank * 100:(rank + 1) * 100] if x % 3 == 0] if world == 2 else None
    if mine is not None:
        assert got[0] == mine, (rank, got[0][:5], mine[:5])
    words = ["a b c", "a b", "a", "c c c d"] * 5
    wc = (ctx.from_enumerable(words, world).select_many(lambda l: l.split(" ")).group_by(lambda w: w)
          .select(lambda g: (g.key, g.count())))
    local = wc.partitions()[0]
    allp = [None] * world
    dist.all_gather_object(allp, local)
    merged = sorted(sum(allp, []))
    assert merged == [("a", 15), ("b", 10), ("c", 20), ("d", 5)], merged
    # ordering: range partition + sort, globally ordered by rank
    vals = [(i * 7919) % 1000 for i in range(400)]
    srt = ctx.from_enumerable(vals, world).order_by(lambda x: x).partitions()[0]
    allp = [None] * world
    dist.all_gather_object(allp, srt)
    assert sum(allp, []) == sorted(vals)
    # sliding window: only the w-1 boundary records move between neighbours
    seq = list(range(100, 160))
    per = len(seq) // world
    win = ctx.from_enumerable(seq, world).sliding_window(lambda w_: sum(w_), 4).partitions()[0]
    allp = [None] * world
    dist.all_gather_object(allp, win)
    assert sum(allp, []) == [sum(seq[j:j + 4]) for j in range(len(seq) - 3)], allp
    # Zip pairs records by GLOBAL position: sides with different partition boundaries are re-aligned
    left = list(range(60))
    right = [x * x for x in range(100)]
    zq = ctx.from_enumerable(left, world).where(lambda x: x % 7 != 0).zip(ctx.from_enumerable(right, world), lambda a, b: a * 100000 + b)
    allp = [None] * world
    dist.all_gather_object(allp, zq.partitions()[0])
    kept = [x for x in left if x % 7 != 0]
    assert sum(allp, []) == [a * 100000 + b for a, b in zip(kept, right)], allp
    # Reverse permutes whole partitions between ranks (partition i <- partition world-1-i), then reverses locally
    rv = ctx.from_enumerable(left, world).select(lambda x: x * 3).reverse().partitions()[0]
    allp = [None] * world
    dist.all_gather_object(allp, rv)
    assert sum(allp, []) == [x * 3 for x in reversed(left)], allp
    # a single-partition stage over a cross-product edge (range partition into ONE range): every rank takes part
    one = ctx.from_enumerable(left, world).range_partition(lambda x: x, 1).select(lambda x: x + 1).partitions()
    allp = [None] * world
    dist.all_gather_object(allp, one[0] if one else [])
    assert sorted(sum(allp, [])) == [x + 1 for x in left]
    # streamed rank-to-rank channel (pointwise FIFO edge): blocks flow while the producer still produces
    from dryad_b200.channel.pipe import PipeReader, PipeWriter
    if world >= 2 and rank < 2:
        if rank == 0:
            w = PipeWriter(1, "cpu", depth=4)
            for j in range(10):
                w.write(torch.arange(1000 * j, 1000 * j + 777, dtype=torch.int64))
            w.write(torch.linspace(0, 1, 33, dtype=torch.float64))
            w.close()
            assert w.blocks == 11
        else:
            got = list(PipeReader(0, "cpu"))
            assert len(got) == 11 and got[-1].dtype == torch.float64 and got[-1].numel() == 33
            assert all(torch.equal(got[j], torch.arange(1000 * j, 1000 * j + 777, dtype=torch.int64)) for j in range(10))
    dist.barrier()
    # scalar aggregate lands on rank 0
    total = ctx.from_enumerable(vals, world).sum_as_query().partitions()[0]
    if rank == 0:
        assert total == [sum(vals)]


def fault_checks(rank, world, device="cpu"):
    """Failure handling of SPMD jobs under the control plane (runtime/spmd_control.py): a transient
    failure on ONE rank is retried by ALL ranks with a new version, a persistent failure aborts every
    rank after the failure budget, and an aborted job commits nothing."""
    import tempfile
    import time as _time
    from dryad_b200 import DryadLinqException, ErrorCode
    from dryad_b200.plan.qnodes import NodeType
    ctx = D.DryadLinqContext(platform="B200", device=device)
    ctx.spmd_control_timeout = 60.0
    data = list(range(300))
    seen = []

    def transient(node, part, version):
        seen.append((node.uid, node.type, part, version))
        head = node.members[0].type if node.type == NodeType.Super else node.type
        if head != NodeType.InputTable and part == world - 1 and version < 2:
            raise RuntimeError("injected transient fault")
    ctx.fault_injector = transient
    got = ctx.from_enumerable(data, world).select(lambda x: x * 2).group_by(lambda x: x % 5).select(
        lambda g: (g.key, g.count())).partitions()[0]
    allp = [None] * world
    dist.all_gather_object(allp, got)
    assert sorted(sum(allp, [])) == [(k, 60) for k in range(5)], allp
    versions = sorted({v for (_, _, p, v) in seen})
    assert versions[-1] >= 2, versions              # every rank re-ran the stage with versions 1 and 2
    cp = ctx.__dict__["_control_plane"]
    if rank == 0:
        assert cp.commands_delivered >= 5            # launch descriptors really travelled through the mailbox

    # persistent failure on one rank: all ranks abort (no hang), within the failure budget
    ctx.fault_injector = None

    def bad(x):
        if rank == world - 1:
            raise ValueError("user code is broken on one rank")
        return x
    t0 = _time.time()
    try:
        ctx.from_enumerable(data, world).select(bad).group_by(lambda x: x % 3).select(lambda g: (g.key, g.count())).partitions()
        raise AssertionError("the job should have failed")
    except DryadLinqException as ex:
        assert ex.error_code == ErrorCode.DidNotCompleteSuccessfully, ex
        assert "user code is broken" in str(ex) or "failed" in str(ex), str(ex)
    assert _time.time() - t0 < 60
    dist.barrier()

    # an aborted job leaves no part files; the same job without the fault commits all of them
    out = os.path.join(tempfile.gettempdir(), f"dryad_spmd_out_{world}.pt")
    base = os.path.splitext(out)[0]
    if rank == 0:
        for f_ in os.listdir(tempfile.gettempdir()):
            if f_.startswith(os.path.basename(base)):
                os.remove(os.path.join(tempfile.gettempdir(), f_))
    dist.barrier()
    try:
        ctx.from_enumerable(data, world).select(bad).to_store("partfile://" + out, delete_if_exists=True).submit_and_wait()
        raise AssertionError("the job should have failed")
    except DryadLinqException:
        pass
    dist.barrier()
    left = [f_ for f_ in os.listdir(tempfile.gettempdir()) if f_.startswith(os.path.basename(base))]
    assert left == [], left
    ctx.from_enumerable(data, world).select(lambda x: x + 1).to_store("partfile://" + out, delete_if_exists=True).submit_and_wait()
    dist.barrier()
    if rank == 0:
        back = D.DryadLinqContext(1).from_store("partfile://" + out).to_list()
        assert sorted(back) == [x + 1 for x in data]
    dist.barrier()


def gpu_checks(rank, world):
    from dryad_b200.models import terasort
    from dryad_b200.ops import core as K
    from dryad_b200.parallel.dagg import DistributedHashAggregate
    from dryad_b200.parallel.dsort import RowSorter
    from dryad_b200.parallel.shuffle import ColSpec, FusedShuffle
    from dryad_b200.parallel.symm import SymmetricArena
    from dryad_b200.utils.hashing import hash_bucket_np
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    arena = SymmetricArena(2 << 30, rank, world, None, dev)
    arena.barrier()
    torch.cuda.synchronize()

    # 1. fused hash shuffle of (key, value) columns: every key lands on its owner, nothing is lost
    n = 1_000_003 + rank * 1000
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    keys = torch.randint(-2**62, 2**62, (n,), dtype=torch.int64, device=dev, generator=g)
    vals = torch.arange(n, dtype=torch.int64, device=dev) + rank * 10_000_000
    sh = FusedShuffle(arena, [ColSpec(8, torch.int64), ColSpec(8, torch.int64)], int(1_010_000 * 1.3) + 4096, max_rows_in=1_010_000)   # same size on every rank
    sh.poison()                                         # debug protocol checker: poisoned buffers + checksums
    sh.run([keys, vals], mode="hash", key_col=0, key_kind=0)
    sh.verify([keys, vals])
    m = sh.received_rows()[0]
    rk, rv = sh.recv[0][0][:m], sh.recv[1][0][:m]
    import numpy as np
    owner = hash_bucket_np(rk.cpu().numpy().view(np.uint64), world)
    assert (owner == rank).all()
    tot = torch.tensor([m, int(rv.sum().item()), n, int(vals.sum().item())], dtype=torch.int64, device=dev)
    dist.all_reduce(tot)
    assert int(tot[0]) == int(tot[2]) and int(tot[1]) == int(tot[3]), tot
    # stability: rows from one source keep their source order
    src0 = rv[(rv >= 0) & (rv < 10_000_000)]
    assert bool((src0[1:] > src0[:-1]).all())

    # 2. distributed sort of 100-byte rows
    nr = 500_000
    rows = terasort.generate_rows(nr, seed=5, device=dev, rank=rank)
    sorter = RowSorter(arena, nr, 100, key_len=10)
    out, mm = sorter.sort(rows)
    res = K.check_sorted_rows(out, 0, 10).cpu()
    cs_in = K.check_sorted_rows(rows, 0, 10).cpu()[1]
    assert int(res[0]) == 0
    first = out[0, :10].cpu().numpy().tobytes() if mm else b""
    last = out[mm - 1, :10].cpu().numpy().tobytes() if mm else b""
    bounds = [None] * world
    dist.all_gather_object(bounds, (first, last, mm, int(res[1]), int(cs_in)))
    assert sum(b[2] for b in bounds) == nr * world
    assert (sum(b[3] for b in bounds) - sum(b[4] for b in bounds)) % (1 << 64) == 0
    ne = [b for b in bounds if b[2]]
    assert all(a[1] <= b[0] for a, b in zip(ne, ne[1:])), "partitions are not globally ordered"

    # 2a. the same with the rows travelling in two key sub-range passes (filtered scatter passes into
    #     per-sub-range receive buffers), and in descending order (one fused shuffle, no side stream)
    for kw in ({"passes": 2}, {"descending": True}):
        mark = arena.mark()
        s2 = RowSorter(arena, nr, 100, key_len=10, **kw)
        out2, m2 = s2.sort(rows)
        desc = bool(kw.get("descending"))
        r2 = K.check_sorted_rows(out2, 0, 10, descending=desc).cpu()
        tt = torch.tensor([int(r2[0]), m2, int(r2[1]) & ((1 << 56) - 1), int(cs_in) & ((1 << 56) - 1)], dtype=torch.int64, device=dev)
        dist.all_reduce(tt)
        assert int(tt[0]) == 0 and int(tt[1]) == nr * world, (kw, tt)
        assert (int(tt[2]) - int(tt[3])) % (1 << 56) == 0, kw
        # (rank boundaries may differ by a few rows from the one-pass sort: the separators are picked
        # at 1/(world*passes) quantiles instead of 1/world)
        lo2 = out2[0, :10].cpu().numpy().tobytes() if m2 else b""
        hi2 = out2[m2 - 1, :10].cpu().numpy().tobytes() if m2 else b""
        b2 = [None] * world
        dist.all_gather_object(b2, (lo2, hi2, m2))
        ne2 = [b for b in b2 if b[2]]
        if desc:
            assert all(a[1] >= b[0] for a, b in zip(ne2, ne2[1:])), "descending partitions are not globally ordered"
        else:
            assert all(a[1] <= b[0] for a, b in zip(ne2, ne2[1:])), "two-pass partitions are not globally ordered"
        torch.cuda.synchronize()
        dist.barrier()
        arena.release_to(mark)

    # 3. fused GroupBy reduce-scatter
    nk = 50_000
    gk = torch.randint(0, nk, (2_000_000,), dtype=torch.int64, device=dev, generator=g)
    gv = torch.randint(0, 100, (2_000_000,), dtype=torch.int64, device=dev, generator=g)
    full_k = [torch.zeros(nk, dtype=torch.int64, device=dev) for _ in range(2)]
    full_k[0].index_add_(0, gk, gv)
    full_k[1].index_add_(0, gk, torch.ones_like(gv))
    dist.all_reduce(full_k[0]); dist.all_reduce(full_k[1])
    for merge in ("push", "atomics"):          # partials pushed through the fused shuffle / remote atomics
        agg = DistributedHashAggregate(arena, ["sum_i64", "count"], 2 * nk, 4 * nk, merge=merge)
        for _ in range(2):
            ok_, ol = agg.run(gk, [gv, None])
        own = hash_bucket_np(ok_.cpu().numpy().view(np.uint64), world)
        assert (own == rank).all(), merge
        assert torch.equal(ol[0], full_k[0][ok_]) and torch.equal(ol[1], full_k[1][ok_]), merge
        cnt_g = torch.tensor([ok_.numel()], dtype=torch.int64, device=dev)
        dist.all_reduce(cnt_g)
        assert int(cnt_g.item()) == int((full_k[1] > 0).sum().item()), merge

    # 3a. NVLS: in-switch all-reduce and multicast copy on a symmetric buffer (when the box has it)
    from dryad_b200.parallel.nvls import NvlsBuffer
    nv = NvlsBuffer.create(8 << 20, dev, rank, world)
    if nv is not None:
        xf, xo = nv.alloc((1 << 18,), torch.float32)
        xu, uo = nv.alloc((4099,), torch.int64)
        bd, bo = nv.alloc((world, 4096), torch.int64)
        gg = torch.Generator(device=dev).manual_seed(50 + rank)
        mine_f = torch.randint(-1000, 1000, (1 << 18,), device=dev, generator=gg).float()
        mine_u = torch.randint(0, 1 << 40, (4099,), dtype=torch.int64, device=dev, generator=gg)
        xf.copy_(mine_f); xu.copy_(mine_u); bd.zero_()
        want_f, want_u = mine_f.clone(), mine_u.clone()
        dist.all_reduce(want_f); dist.all_reduce(want_u)
        arena.barrier(slot=2)
        nv.allreduce_f32(xo, 1 << 18)
        nv.allreduce_u64(uo, 4099)
        nv.broadcast(bo + rank * 4096 * 8, torch.full((4096,), rank + 1, dtype=torch.int64, device=dev))
        arena.barrier(slot=3)
        torch.cuda.synchronize()
        assert torch.equal(xf, want_f) and torch.equal(xu, want_u)          # integer-valued floats: exact
        assert torch.equal(bd, (torch.arange(world, device=dev) + 1)[:, None].expand(world, 4096))
        nvls_on = True
    else:
        nvls_on = False

    # 3b. broadcast join: the small side is replicated by the build kernel, the fact side stays put
    from dryad_b200.models import joins
    jr = joins.bench(rank, world, dev, 1_000_000, 50_000, steps=1, warmup=1,
                     max_over_ranks=lambda x: x, barrier=lambda: (dist.barrier(), torch.cuda.synchronize()))
    assert jr["valid"], jr
    # 3c. k-means: tcgen05 assign on every rank + peer-load all-reduce; all ranks end with equal centroids
    from dryad_b200.models import kmeans as KM
    from dryad_b200.parallel.kmeans import KMeans
    pts, centres = KM.generate(200_000, 256, seed=21, device=dev, rank=rank)
    km = KMeans(arena, 256, 200_000)
    km.set_centroids(centres + 0.05)
    for _ in range(2):
        km.iterate(pts)
    torch.cuda.synchronize()
    allc = [torch.zeros_like(km.centroids) for _ in range(world)]
    dist.all_gather(allc, km.centroids)
    assert all(torch.equal(allc[0], c) for c in allc)
    assert int(km.total_counts.sum().item()) == 200_000 * world
    # same with the peer-load all-reduce (no NVLS): partials may only be cleared after every rank
    # has read them, so centroids stay bit-identical across ranks over several iterations
    km2 = KMeans(arena, 256, 200_000, use_nvls=False)
    km2.set_centroids(centres + 0.05)
    for _ in range(6):
        km2.iterate(pts)
    torch.cuda.synchronize()
    allc = [torch.zeros_like(km2.centroids) for _ in range(world)]
    dist.all_gather(allc, km2.centroids)
    assert all(torch.equal(allc[0], c) for c in allc), "P2P k-means all-reduce diverged across ranks"
    assert int(km2.total_counts.sum().item()) == 200_000 * world

    # 3d. streamed channel between GPUs (NCCL p2p on a side stream, overlapped with the producer)
    from dryad_b200.channel.pipe import PipeReader, PipeWriter
    if rank < 2:
        if rank == 0:
            w = PipeWriter(1, dev, depth=4)
            for j in range(8):
                blk = torch.full((1 << 20,), j, dtype=torch.int64, device=dev) * 3 + 1    # produced on the compute stream
                w.write(blk)
            w.close()
        else:
            got = list(PipeReader(0, dev))
            assert len(got) == 8 and all(int(b[0]) == 3 * j + 1 and int(b.sum()) == (3 * j + 1) * (1 << 20) for j, b in enumerate(got))
    dist.barrier()

    # 3e. HBM ring FIFO with flag-based flow control: 40 blocks through a 4-slot ring in both
    #     directions between ranks 0 and 1, no host involvement per block beyond the launches
    from dryad_b200.parallel.ring import HbmRing
    fwd = HbmRing(arena, 0, 1, slots=4, slot_bytes=1 << 20)
    bwd = HbmRing(arena, 1, 0, slots=4, slot_bytes=1 << 20)
    if rank < 2:
        out_ring, in_ring = (fwd, bwd) if rank == 0 else (bwd, fwd)
        nblk, words = 40, (1 << 20) // 8
        got = torch.empty((nblk, words), dtype=torch.int64, device=dev)
        blocks = [torch.full((words - 3 * j,), 1000 * rank + j, dtype=torch.int64, device=dev) for j in range(nblk)]
        for j in range(nblk):
            out_ring.push(blocks[j])
            in_ring.pop(got[j])
        torch.cuda.synchronize()
        out_ring.check(); in_ring.check()
        peer = 1 - rank
        for j in range(nblk):
            assert bool((got[j, : words - 3 * j] == 1000 * peer + j).all()), (rank, j)
        assert int(in_ring.out_len.item()) == (words - 3 * (nblk - 1)) * 8
    dist.barrier()

    # 4. the same sort through the public API (SPMD executor -> RowSorter)
    ctx = D.DryadLinqContext(platform="B200", device=str(dev))
    host = rows.cpu().pin_memory()
    res_q = ctx.from_rows(host, key_off=0, key_len=10).order_by(D.RowKey(0, 10)).submit_and_wait()
    viol, cs, cnt = terasort.validate(res_q)
    t = torch.tensor([viol, cnt], dtype=torch.int64, device=dev)
    dist.all_reduce(t)
    assert int(t[0]) == 0 and int(t[1]) == nr * world
    torch.cuda.synchronize()

    # 5. GroupBy and Join through the public API on columnar data: the HashPartition stages are the
    #    fused partition + all-to-all kernel (no object exchange), results equal the oracle
    nrec = 200_000
    gq = torch.Generator().manual_seed(7 + rank)
    kk = torch.randint(0, 5000, (nrec,), dtype=torch.int64, generator=gq)
    vv = torch.randint(0, 100, (nrec,), dtype=torch.int64, generator=gq)
    tbl = ctx.from_columns({"k": kk, "v": vv})
    res = tbl.group_by(lambda r: r.k).select(lambda g: {"k": g.key, "s": g.sum(lambda r: r.v), "c": g.count()}).partitions()[0]
    be = ctx._backend()
    assert be.host_fallbacks == 0, be.host_fallbacks
    gathered = [None] * world
    dist.all_gather_object(gathered, (kk, vv))
    fk = torch.cat([g_[0] for g_ in gathered]); fv = torch.cat([g_[1] for g_ in gathered])
    want_s = torch.zeros(5000, dtype=torch.int64).index_add_(0, fk, fv)
    want_c = torch.bincount(fk, minlength=5000)
    from dryad_b200.utils.hashing import hash_bucket
    mine = {r["k"]: (r["s"], r["c"]) for r in res}
    assert all(hash_bucket(k_, world) == rank for k_ in mine)
    assert all(mine[k_] == (int(want_s[k_]), int(want_c[k_])) for k_ in mine)
    cnt_groups = torch.tensor([len(mine)], dtype=torch.int64, device=dev)
    dist.all_reduce(cnt_groups)
    assert int(cnt_groups.item()) == int((want_c > 0).sum().item())
    # ... and it ran as ONE fused local-aggregate + NVLink reduce-scatter, with no pickled exchange
    assert getattr(be, "fused_aggregates", 0) >= 1, "GroupBy did not take the DistributedHashAggregate path"
    assert ctx.last_object_collectives == 0, ctx.last_object_collectives
    torch.cuda.synchronize()

    # an explicit HashPartition stage is the fused partition + all-to-all kernel
    hp = tbl.hash_partition(lambda r: r.k, world).to_columns()
    assert getattr(be, "fused_exchanges", 0) >= 1 and ctx.last_object_collectives == 0
    assert all(hash_bucket(int(k_), world) == rank for k_ in hp["k"].cpu().tolist()[:2000])
    tot_hp = torch.tensor([hp["k"].numel()], dtype=torch.int64, device=dev)
    dist.all_reduce(tot_hp)
    assert int(tot_hp.item()) == nrec * world

    # sliding window on columns: the w-1 boundary rows move as one P2P neighbour write into the arena
    seq = torch.arange(1000 * rank, 1000 * rank + 1000, dtype=torch.int64)
    win = ctx.from_columns({"value": seq}).sliding_window(lambda w_: w_[0] + w_[1] + w_[2], 3).to_columns()["value"].cpu()
    lo = 1000 * rank - (2 if rank > 0 else 0)
    full = torch.arange(lo, 1000 * rank + 1000, dtype=torch.int64)
    assert torch.equal(win, full[:-2] + full[1:-1] + full[2:]), (rank, win[:4])
    assert ctx.__dict__.get("last_p2p_halos", 0) >= 1 and ctx.last_object_collectives == 0 and be.host_fallbacks == 0

    # N:1 merges of columnar partials are streamed rank-to-rank channels: the p2p pipe, then the HBM ring
    for pipe in ("nccl", "ring"):
        ctx.spmd_pipe = pipe
        tot = tbl.select(lambda r: r.v).sum_as_query().to_columns()
        if rank == 0:
            assert int(tot["value"][0].item()) == int(fv.sum()), (pipe, tot)
        assert ctx.__dict__.get("last_streamed_edges", 0) >= 1 and ctx.last_object_collectives == 0, pipe
        top = tbl.select(lambda r: r.v).order_by(lambda x: x).take(5).to_columns()
        if rank == 0:
            assert top["value"].cpu().tolist() == sorted(fv.tolist())[:5], pipe
    ctx.spmd_pipe = "nccl"
    assert be.host_fallbacks == 0, be.fallback_log

    # 5a. dimension join through the API: the planner picks the broadcast join (small inner side), the
    #     SPMD executor runs it as BroadcastJoin + generated probe; no object collectives
    nd = 3000                                            # dimension rows in total, sliced over the ranks
    per = (nd + world - 1) // world
    dlo, dhi = rank * per, min(nd, (rank + 1) * per)
    dim = ctx.from_columns({"id": torch.arange(dlo, dhi, dtype=torch.int64), "w": (torch.arange(dlo, dhi, dtype=torch.int64) * 7) % 13})
    fk = torch.randint(0, nd + 500, (nrec,), dtype=torch.int64, generator=gq)        # some keys have no match
    fact = ctx.from_columns({"k": fk, "v": vv})
    jq = fact.join(dim, lambda r: r.k, lambda d: d.id, lambda r, d: {"k": r.k, "x": r.v * d.w})
    jcols = jq.to_columns()
    assert getattr(be, "fused_broadcast_joins", 0) >= 1, "Join did not take the BroadcastJoin path"
    assert ctx.last_object_collectives == 0 and be.host_fallbacks == 0, (ctx.last_object_collectives, be.host_fallbacks)
    sel = fk < nd
    want_x = (vv[sel] * ((fk[sel] * 7) % 13))
    got_k, got_x = jcols["k"].cpu(), jcols["x"].cpu()
    assert got_k.numel() == int(sel.sum()) and torch.equal(got_k, fk[sel]) and torch.equal(got_x, want_x)
    torch.cuda.synchronize()

    # 5b. skew: half of all rows carry ONE key.  OrderBy puts them into one partition (its receive
    #     buffer is re-created from the published counts), GroupBy folds them into one owner slot.
    skew_rows = terasort.generate_rows(nr, seed=9, device=dev, rank=rank)
    skew_rows[::2, :10] = 0x41
    sk_sorter = RowSorter(arena, nr, 100, key_len=10)
    sk_out, sk_m = sk_sorter.sort(skew_rows)
    rs = K.check_sorted_rows(sk_out, 0, 10).cpu()
    tt = torch.tensor([int(rs[0]), sk_m], dtype=torch.int64, device=dev)
    dist.all_reduce(tt)
    assert int(tt[0]) == 0 and int(tt[1]) == nr * world, tt
    assert sk_sorter.regrown >= 1, "the skewed sort was expected to re-create its receive side"
    presorted, pm = sk_sorter.sort(sk_out[:sk_m].clone() if sk_m <= nr else sk_out[:nr].clone())   # already-sorted input
    rs = K.check_sorted_rows(presorted, 0, 10).cpu()
    assert int(rs[0]) == 0
    sk_keys = torch.where(torch.arange(nrec) % 2 == 0, torch.zeros(nrec, dtype=torch.int64), kk + 1)
    sk_res = ctx.from_columns({"k": sk_keys, "v": vv}).group_by(lambda r: r.k).select(
        lambda g: {"k": g.key, "c": g.count()}).to_columns()
    tot_c = torch.tensor([int(sk_res["c"].sum().item()) if sk_res else 0], dtype=torch.int64, device=dev)
    dist.all_reduce(tot_c)
    assert int(tot_c.item()) == nrec * world
    torch.cuda.synchronize()


def wordcount_checks(rank, world, dev):
    """WordCount through the API on `world` GPUs: every rank reads its own byte range of the text
    file, tokenizes on the GPU, partial-aggregates on token hashes, the partials (hash, length,
    padded bytes, count) cross NVLink in the fused hash shuffle, owners aggregate and verify."""
    import tempfile
    from dryad_b200.models import wordcount as WC
    path = os.path.join(tempfile.gettempdir(), "dryad_wc_corpus.txt")
    if rank == 0:
        with open(path, "wb") as f:
            f.write(WC.generate_text(20_000, words_per_line=10, seed=11, vocab=3000))
    dist.barrier()
    ctx = D.DryadLinqContext(platform="B200", device=str(dev))
    q = (ctx.from_store("file://" + path).select_many(lambda l: l.split(" ")).group_by(lambda w: w)
         .select(lambda g: D.Pair(g.key, g.count())))
    mine = dict(q.partitions()[0])
    be = ctx._backend()
    assert be.host_fallbacks == 0, be.host_fallbacks
    assert ctx.last_object_collectives == 0
    everyone = [None] * world
    dist.all_gather_object(everyone, mine)
    merged = {}
    for d_ in everyone:
        assert not (set(d_) & set(merged)), "a word is owned by two ranks"
        merged.update(d_)
    if rank == 0:
        from collections import Counter
        with open(path, "rb") as f:
            want = Counter(w.decode() for line in f.read().split(b"\n")[:-1] for w in line.split(b" "))
        assert merged == dict(want), (len(merged), len(want))
    dist.barrier()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="cpu")
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if a.mode == "cpu":
        dist.init_process_group("gloo")
        cpu_checks(rank, world)
        fault_checks(rank, world)
    else:
        dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
        torch.cuda.set_device(dev)
        dist.init_process_group("nccl", device_id=dev)
        gpu_checks(rank, world)
        wordcount_checks(rank, world, dev)
        fault_checks(rank, world, str(dev))
    dist.barrier()
    if rank == 0:
        print(f"SPMD {a.mode} checks passed on {world} ranks")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
