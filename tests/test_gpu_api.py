"""The public query API on the GPU backend must agree with the LocalDebug oracle (the reference's
cluster-vs-LocalDebug pattern, DryadLinqTests/Utils.cs:301-360), and must actually use the CUDA
kernels (no silent host fallback) for traced stages."""
import pytest
import torch

import dryad_b200 as D

pytestmark = pytest.mark.gpu


def _both(build, data, parts=(1, 3)):
    ld = D.DryadLinqContext(1, local_debug=True)
    want = sorted(build(ld.from_enumerable(data, 1)).to_list(), key=repr)
    for p in parts:
        ctx = D.DryadLinqContext(p, device="cuda")
        got = sorted(build(ctx.from_enumerable(data, p)).to_list(), key=repr)
        assert got == want, f"partitions={p}"
    return ctx


def test_select_where_on_gpu_uses_kernels():
    from dryad_b200.ops import core as K
    data = list(range(1, 5001))
    before = K.launch_count()
    ctx = _both(lambda q: q.where(lambda x: x % 3 == 0).select(lambda x: x * 2 + 1), data)
    assert K.launch_count() > before
    assert ctx._backend().host_fallbacks == 0


def test_records_group_by_sum_count_avg():
    data = [{"k": i % 17, "v": i} for i in range(4000)]
    _both(lambda q: q.group_by(lambda r: r.k).select(
        lambda g: {"k": g.key, "s": g.sum(lambda r: r.v), "c": g.count(), "a": g.average(lambda r: r.v),
                   "mx": g.max(lambda r: r.v)}), data)


def test_order_by_hash_partition_distinct_join_aggregates():
    data = [(i * 7919) % 1000 for i in range(3000)]
    ld = D.DryadLinqContext(1, local_debug=True)
    for p in (1, 4):
        ctx = D.DryadLinqContext(p, device="cuda")
        q = ctx.from_enumerable(data, p)
        assert q.order_by(lambda x: x).to_list() == sorted(data)
        assert q.order_by_descending(lambda x: x).to_list() == sorted(data, reverse=True)
        assert sorted(q.distinct().to_list()) == sorted(set(data))
        assert q.count() == len(data) and q.sum() == sum(data) and q.min() == min(data) and q.max() == max(data)
        parts = q.hash_partition(lambda x: x, 4).partitions()
        assert sorted(sum(parts, [])) == sorted(data)
        from dryad_b200.utils.hashing import hash_bucket
        for j, part in enumerate(parts):
            assert all(hash_bucket(x, 4) == j for x in part)
        rp = q.range_partition(lambda x: x, 4).partitions()
        nonempty = [x for x in rp if x]
        assert all(max(a) <= min(b) for a, b in zip(nonempty, nonempty[1:]))
    left = [{"id": i, "a": i * 2} for i in range(500)]
    right = [{"id": i % 100, "b": i} for i in range(300)]
    want = sorted(ld.from_enumerable(left).join(ld.from_enumerable(right), lambda l: l.id, lambda r: r.id,
                                                lambda l, r: {"id": l.id, "ab": l.a + r.b}).to_list(), key=repr)
    ctx = D.DryadLinqContext(3, device="cuda")
    got = sorted(ctx.from_enumerable(left, 3).join(ctx.from_enumerable(right, 2), lambda l: l.id, lambda r: r.id,
                                                   lambda l, r: {"id": l.id, "ab": l.a + r.b}).to_list(), key=repr)
    assert got == want


def test_opaque_lambdas_fall_back_to_host_and_still_agree():
    data = [f"w{i % 13}" for i in range(500)]
    _both(lambda q: q.group_by(lambda w: w).select(lambda g: (g.key, g.count())), data)


def test_rows_order_by_through_api():
    from dryad_b200.models import terasort
    rows = terasort.generate_rows(100_000, seed=3)
    host = rows.cpu().pin_memory()
    ctx = D.DryadLinqContext(1, device="cuda")
    res = ctx.from_rows(host, key_off=0, key_len=10).order_by(D.RowKey(0, 10)).submit_and_wait()
    viol, checksum, n = terasort.validate(res)
    assert viol == 0 and n == 100_000 and checksum == terasort.key_checksum(rows)
    out = res.expr.params["table"].partitions[0].cols["row"].cpu()
    assert terasort.validate_sorted(out)


def test_merge_join_and_take_after_order_on_gpu():
    ld = D.DryadLinqContext(1, local_debug=True)
    left = [{"k": (i * 37) % 50, "a": i} for i in range(400)]
    right = [{"k": i % 60, "b": -i} for i in range(300)]

    def build(c, n):
        l = c.from_enumerable(left, n).order_by(lambda r: r.k)
        return l.join(c.from_enumerable(right, n), lambda r: r.k, lambda r: r.k, lambda x, y: {"k": x.k, "s": x.a + y.b})
    canon = lambda rows: sorted(tuple(sorted(dict(r).items())) for r in rows)
    want = canon(build(ld, 1).to_list())
    ctx = D.DryadLinqContext(1, device="cuda")
    q = build(ctx, 1)
    assert "algorithm=merge" in q.explain()
    assert canon(q.to_list()) == want
    vals = [(i * 7919) % 1000 for i in range(2000)]
    c4 = D.DryadLinqContext(4, device="cuda")
    assert c4.from_enumerable(vals, 4).order_by(lambda x: x).take(10).to_list() == sorted(vals)[:10]


def test_join_against_unique_dimension_runs_as_one_fused_kernel():
    """Dimension-table join (unique build keys): probe + gather + result selector are one generated
    kernel; a build side with duplicate keys takes the general pair-emitting path.  Both agree with
    the oracle."""
    from dryad_b200.ops import core as K
    fact = [{"k": (i * 31) % 140, "v": i} for i in range(3000)]            # keys 100..139 have no match
    dim = [{"k": i, "w": i * 3 + 1, "f": i * 0.5} for i in range(100)]
    dup = dim + [{"k": 7, "w": -1, "f": 0.25}]
    sel = lambda a, b: {"k": a.k, "x": a.v * b.w, "g": b.f + 1.0}    # noqa: E731
    ld = D.DryadLinqContext(1, local_debug=True)
    for d in (dim, dup):
        want = sorted(ld.from_enumerable(fact).join(ld.from_enumerable(d), lambda a: a.k, lambda b: b.k, sel).to_list(), key=repr)
        ctx = D.DryadLinqContext(2, device="cuda")
        before = K.launch_count()
        got = sorted(ctx.from_enumerable(fact, 2).join(ctx.from_enumerable(d, 1), lambda a: a.k, lambda b: b.k, sel).to_list(), key=repr)
        assert got == want
        assert K.launch_count() > before and ctx._backend().host_fallbacks == 0


def test_fused_probe_reduce_matches_torch():
    from dryad_b200.expr import col, gather, probe, where
    from dryad_b200.expr.compile import run_on_gpu
    from dryad_b200.ops import core as K
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(5)
    n, nd = 1_000_003, 5000
    fk = torch.randint(0, nd + 700, (n,), dtype=torch.int64, device=dev, generator=g)
    fv = torch.randint(0, 1000, (n,), dtype=torch.int64, device=dev, generator=g)
    dk = torch.randperm(nd, device=dev, generator=g).to(torch.int64)
    da = (dk * 2654435761) % 97
    jt = K.JoinTable(dk.contiguous())
    assert jt.unique
    hit = probe("dim", col("k"))
    keep = (col("v") >= 500) & (hit >= 0)
    out, cnt = run_on_gpu({"r": where(keep, col("v") * gather("a", hit), 0)}, {"k": fk, "v": fv}, n, predicate=keep,
                          tables={"dim": jt}, arrays={"a": da}, reduce={"r": "sum"}, count_flags=True)
    m = (fv >= 500) & (fk < nd)
    assert int(cnt.item()) == int(m.sum().item())
    assert int(out["r"].item()) == int((fv[m] * ((fk[m] * 2654435761) % 97)).sum().item())
    assert not K.JoinTable(torch.cat([dk, dk[:3]]).contiguous()).unique


def test_non_decomposable_group_by_and_set_operations_on_gpu():
    """GroupBy that materialises its groups runs sort + run boundaries on the GPU (K7); Union /
    Intersect / Except run on the hash-set kernels (K10); all agree with the oracle."""
    data = [{"k": (i * 37) % 23, "v": i} for i in range(3000)]
    norm = lambda gs: sorted((g.key, list(g)) for g in gs)          # noqa: E731
    ld = D.DryadLinqContext(1, local_debug=True)
    want = norm(ld.from_enumerable(data).group_by(lambda r: r.k, lambda r: r.v).to_list())
    for p in (1, 3):
        ctx = D.DryadLinqContext(p, device="cuda")
        got = norm(ctx.from_enumerable(data, p).group_by(lambda r: r.k, lambda r: r.v).to_list())
        assert got == want          # (groups are host objects: stages after the GroupBy run on the host)
    a = [(i * 7) % 400 for i in range(2000)]
    b = [(i * 11) % 600 + 200 for i in range(1500)]
    for p in (1, 2):
        ctx = D.DryadLinqContext(p, device="cuda")
        qa, qb = ctx.from_enumerable(a, p), ctx.from_enumerable(b, p)
        assert sorted(qa.union(qb).to_list()) == sorted(set(a) | set(b))
        assert sorted(qa.intersect(qb).to_list()) == sorted(set(a) & set(b))
        assert sorted(qa.except_(qb).to_list()) == sorted(set(a) - set(b))
        assert ctx._backend().host_fallbacks == 0


def test_group_by_dense_integer_keys_uses_direct_table_and_agrees():
    """Large partitions with a dense integer key range take the direct-addressed aggregate table."""
    import numpy as np
    n = 300_000
    rng = np.random.default_rng(9)
    k = rng.integers(-50, 4000, n)
    v = rng.integers(-100, 100, n)
    ctx = D.DryadLinqContext(1, device="cuda")
    got = ctx.from_columns({"k": k, "v": v}).group_by(lambda r: r.k).select(
        lambda g: {"k": g.key, "s": g.sum(lambda r: r.v), "c": g.count(), "mx": g.max(lambda r: r.v)}).to_list()
    want_s = np.zeros(4050, dtype=np.int64); np.add.at(want_s, k + 50, v)
    want_c = np.bincount(k + 50, minlength=4050)
    want_m = np.full(4050, -1000); np.maximum.at(want_m, k + 50, v)
    assert len(got) == int((want_c > 0).sum())
    for r in got:
        j = r["k"] + 50
        assert (r["s"], r["c"], r["mx"]) == (want_s[j], want_c[j], want_m[j])
    assert ctx._backend().host_fallbacks == 0


def test_gpu_timeline_records_device_time_per_vertex(monkeypatch):
    """DRYAD_TRACE=1: one CUDA-event span (+ NVTX range) per vertex, resolved once at job end and
    written next to the event log."""
    import json
    import os
    monkeypatch.setenv("DRYAD_TRACE", "1")
    ctx = D.DryadLinqContext(2, device="cuda")
    data = list(range(20000))
    got = ctx.from_enumerable(data, 2).where(lambda x: x % 2 == 0).select(lambda x: x * 3).to_list()
    assert sorted(got) == [x * 3 for x in data if x % 2 == 0]
    ex = ctx.last_executor
    assert ex.gpu_timeline and all(s["gpu_ms"] >= 0 and "stage" in s for s in ex.gpu_timeline)
    path = os.path.join(ex.reporter.job_dir, "gpu_timeline.jsonl")
    lines = [json.loads(l) for l in open(path)]
    assert len(lines) == len(ex.gpu_timeline) >= 2


def test_columnar_ingress_and_egress_stay_on_the_device():
    import numpy as np
    n = 200_000
    rng = np.random.default_rng(2)
    k, v = rng.integers(0, 3000, n), rng.integers(0, 50, n)
    ctx = D.DryadLinqContext(2, device="cuda")
    cols = ctx.from_columns({"k": torch.from_numpy(k).cuda(), "v": torch.from_numpy(v).cuda()}, 2).group_by(lambda r: r.k).select(
        lambda g: {"k": g.key, "s": g.sum(lambda r: r.v), "c": g.count()}).to_columns()
    assert cols["k"].is_cuda and cols["k"].numel() == len(set(k.tolist()))
    order = torch.argsort(cols["k"])
    want_s = np.zeros(3000, dtype=np.int64); np.add.at(want_s, k, v)
    occ = np.nonzero(np.bincount(k, minlength=3000))[0]
    assert cols["k"][order].cpu().tolist() == occ.tolist()
    assert cols["s"][order].cpu().tolist() == want_s[occ].tolist()
    assert int(cols["c"].sum()) == n


def test_key_minus_one_is_an_ordinary_key_on_the_gpu_backend():
    """int64 -1 has the bit pattern the hash tables reserve for EMPTY: queries whose keys contain
    it must still agree with the oracle (host path for that vertex, or the direct table)."""
    data = [{"k": (i % 7) - 1, "v": i} for i in range(2000)]               # keys -1 .. 5
    ld = D.DryadLinqContext(1, local_debug=True)
    q = lambda c, p: c.from_enumerable(data, p).group_by(lambda r: r.k).select(      # noqa: E731
        lambda g: {"k": g.key, "s": g.sum(lambda r: r.v), "c": g.count()})
    want = sorted(q(ld, 1).to_list(), key=repr)
    for p in (1, 3):
        assert sorted(q(D.DryadLinqContext(p, device="cuda"), p).to_list(), key=repr) == want
    vals = [(i * 13) % 11 - 1 for i in range(500)]
    other = [-1, 3, 99]
    ctx = D.DryadLinqContext(2, device="cuda")
    a, b = ctx.from_enumerable(vals, 2), ctx.from_enumerable(other, 1)
    assert sorted(a.distinct().to_list()) == sorted(set(vals))
    assert sorted(a.intersect(b).to_list()) == sorted(set(vals) & set(other))
    assert sorted(a.union(b).to_list()) == sorted(set(vals) | set(other))
    left = [{"id": i % 5 - 1, "a": i} for i in range(50)]
    right = [{"id": i - 1, "b": i * 10} for i in range(5)]
    wantj = sorted(ld.from_enumerable(left).join(ld.from_enumerable(right), lambda l: l.id, lambda r: r.id,
                                                 lambda l, r: (l.id, l.a + r.b)).to_list())
    gotj = sorted(ctx.from_enumerable(left, 2).join(ctx.from_enumerable(right, 1), lambda l: l.id, lambda r: r.id,
                                                    lambda l, r: (l.id, l.a + r.b)).to_list())
    assert gotj == wantj


def test_wordcount_from_text_store_runs_on_the_gpu_without_host_fallback(tmp_path):
    """samples/WordCount.cs.pp:72-75 through the public API on the GPU backend: text file -> string
    column (views into the uploaded bytes) -> tokenizer kernel -> hash aggregate on 64-bit token
    hashes with byte-for-byte verification -> (word, count) pairs; equal to the LocalDebug oracle."""
    from dryad_b200.models import wordcount as WC
    text = WC.generate_text(3000, words_per_line=9, seed=3, vocab=700)
    text += b"\n  double  spaces and a trailing space \n\nlast line without newline"
    path = tmp_path / "corpus.txt"
    path.write_bytes(text)

    def query(ctx):
        return (ctx.from_store("file://" + str(path)).select_many(lambda l: l.split(" ")).group_by(lambda w: w)
                .select(lambda g: D.Pair(g.key, g.count())))

    want = dict(query(D.DryadLinqContext(1, local_debug=True)).to_list())
    ctx = D.DryadLinqContext(1, device="cuda")
    got = dict(query(ctx).to_list())
    assert got == want
    assert "" in got and got[""] >= 3                      # empty tokens: String.Split(' ') semantics
    assert ctx._backend().host_fallbacks == 0
    # a second decomposable lane and a numeric result next to the string key
    q2 = (ctx.from_store("file://" + str(path)).select_many(lambda l: l.split(" ")).group_by(lambda w: w)
          .select(lambda g: {"word": g.key, "n": g.count(), "twice": g.count() * 2}))
    recs = q2.to_list()
    assert {r["word"]: r["n"] for r in recs} == want and all(r["twice"] == 2 * r["n"] for r in recs)
    assert ctx._backend().host_fallbacks == 0


def test_string_columns_partition_and_concat_round_trip():
    """StrColumn plumbing: pack into fixed-width slots, rebuild, gather, concat, hashes equal the host's."""
    from dryad_b200.runtime.columnar import ColumnBatch, StrColumn
    from dryad_b200.utils.hashing import hash_bytes_py
    words = [b"alpha", b"", b"be", b"gamma-delta-epsilon", b"z"] * 50
    blob = b"\n".join(words) + b"\n"
    buf = torch.frombuffer(bytearray(blob), dtype=torch.uint8).cuda()
    sc = StrColumn(buf, None, None, None, whole_lines=True)
    assert sc.to_list() == [w.decode() for w in words]
    hs = sc.key_hashes().cpu().tolist()
    assert [h & ((1 << 64) - 1) for h in hs] == [hash_bytes_py(w) for w in words]
    packed = sc.pack(20)
    back = StrColumn.from_packed(packed, sc.lens.to(torch.int64), sc.hashes)
    assert back.to_list() == sc.to_list()
    idx = torch.tensor([3, 0, 249], dtype=torch.int32, device="cuda")
    b = ColumnBatch({"w": sc, "i": torch.arange(250, device="cuda")})
    t = b.take(idx)
    assert t.cols["w"].to_list() == ["gamma-delta-epsilon", "alpha", "z"] and t.cols["i"].tolist() == [3, 0, 249]
    c = ColumnBatch.concat([ColumnBatch({"w": back}), ColumnBatch({"w": sc})])
    assert c.n == 500 and c.cols["w"].to_list() == sc.to_list() * 2


def test_kmeans_as_a_do_while_query_matches_the_direct_engine():
    """BASELINE config 5 through the API: `do_while` submits one job per iteration whose body is
    `points.apply_per_partition(KMeansStep, centroids)` — a device function on the GPU backend."""
    from dryad_b200.models import kmeans as KM
    from dryad_b200.parallel.kmeans import KMeans
    from dryad_b200.parallel.symm import SymmetricArena
    n, k, iters = 50_000, 256, 3
    pts, centres = KM.generate(n, k, seed=4, device="cuda")
    init = centres + 0.05
    ctx = D.DryadLinqContext(1, device="cuda")
    step = KM.KMeansStep(ctx)
    points = ctx.from_columns({"x": pts})
    cents0 = ctx.from_columns({"c": init})

    class Budget:
        device = True

        def __init__(self, limit):
            self.left = limit

        def __call__(self, new_c, old_c):
            self.left -= 1
            return {"value": torch.tensor([self.left > 0], device=new_c.cols["c"].device)}
    budget = Budget(iters)
    final = cents0.do_while(lambda cur: points.apply_per_partition(step, cur),
                            lambda cur, nxt: nxt.apply_per_partition(budget, cur))
    got = final.to_columns()["c"]
    assert step.iterations == iters and ctx._backend().host_fallbacks == 0
    km = KMeans(SymmetricArena(8 << 20), k, n)
    km.set_centroids(init)
    for _ in range(iters):
        km.iterate(pts)
    torch.cuda.synchronize()
    assert torch.allclose(got, km.centroids, rtol=1e-5, atol=1e-5)


def _same_order(build, data, parts=(1, 3)):
    """Exact (ordered) agreement with the LocalDebug oracle; returns the last GPU context."""
    ld = D.DryadLinqContext(1, local_debug=True)
    want = build(ld.from_enumerable(data, 1)).to_list()
    for p in parts:
        ctx = D.DryadLinqContext(p, device="cuda")
        got = build(ctx.from_enumerable(data, p)).to_list()
        assert got == want, f"partitions={p}: {got[:8]} vs {want[:8]}"
        assert ctx._backend().host_fallbacks == 0, f"partitions={p}: {ctx._backend().host_fallbacks} host fallbacks"
    return ctx


def test_order_dependent_operators_run_on_the_gpu():
    """Global index (count -> scan -> offsets, K16), Take/Skip(While), Reverse and Zip on columns."""
    data = [(i * 7919) % 5003 for i in range(4001)]
    _same_order(lambda q: q.select_indexed(lambda x, i: x * 2 + i), data)
    _same_order(lambda q: q.where_indexed(lambda x, i: (x + i) % 3 == 0), data)
    _same_order(lambda q: q.take(37), data)
    _same_order(lambda q: q.skip(3950), data)
    _same_order(lambda q: q.take_while(lambda x: x != 2500), data)
    _same_order(lambda q: q.skip_while(lambda x: x != 2500), data)
    _same_order(lambda q: q.take_while_indexed(lambda x, i: i < 1234), data)
    _same_order(lambda q: q.reverse(), data)
    _same_order(lambda q: q.zip(q.select(lambda x: x * 3), lambda a, b: a + b), data)
    recs = [{"k": i % 11, "v": i} for i in range(3000)]
    _same_order(lambda q: q.zip(q.select(lambda r: r.v + 1), lambda r, w: {"k": r.k, "s": r.v + w}), recs)


def test_scalar_aggregates_run_on_the_gpu_and_sum_is_checked_exactly():
    data = [(i * 104729) % 100003 - 50000 for i in range(6000)]
    for build in (lambda q: q.sum_as_query(), lambda q: q.average_as_query(), lambda q: q.min_as_query(),
                  lambda q: q.max_as_query(), lambda q: q.count_as_query(lambda x: x % 7 == 0),
                  lambda q: q.any_as_query(lambda x: x == data[4321]), lambda q: q.any_as_query(lambda x: x > 10**9),
                  lambda q: q.all_as_query(lambda x: x > -60000), lambda q: q.all_as_query(lambda x: x > 0),
                  lambda q: q.first_as_query(), lambda q: q.last_as_query(),
                  lambda q: q.first_as_query(lambda x: x % 1000 == 0), lambda q: q.last_as_query(lambda x: x % 1000 == 0),
                  lambda q: q.contains_as_query(data[17]), lambda q: q.contains_as_query(123456789),
                  lambda q: q.select(lambda x: x * 0.5).average_as_query()):
        _same_order(build, data)
    # checked arithmetic is exact at the int64 boundary (no float shadow)
    top = (1 << 63) - 1
    ctx = D.DryadLinqContext(2, device="cuda")
    assert ctx.from_enumerable([top - 5, 2, 3], 2).sum_as_query().to_list() == [top]
    with pytest.raises(D.DryadLinqException):
        ctx.from_enumerable([top - 5, 2, 4], 2).sum_as_query().to_list()
    with pytest.raises(D.DryadLinqException):
        ctx.from_enumerable([-(1 << 63), -1], 1).sum_as_query().to_list()
    assert ctx.from_enumerable([1 << 62, 1 << 62, -(1 << 62), -(1 << 62), 7], 2).sum_as_query().to_list() == [7]


def test_float_zero_keys_and_mixed_key_kinds_agree_with_the_oracle():
    """-0.0 and 0.0 are one key; a float outer key joined with an int inner key compares values."""
    vals = [0.0, -0.0, 1.5, -0.0, 0.0, 1.5, 2.0]
    _both(lambda q: q.group_by(lambda x: x).select(lambda g: (g.key + 0.0, g.count())), vals, parts=(1, 2))
    left = [float(i % 5) for i in range(40)]
    ld = D.DryadLinqContext(1, local_debug=True)
    want = sorted(ld.from_enumerable(left, 1).join(ld.from_enumerable([1, 3], 1), lambda a: a, lambda b: b, lambda a, b: (a, b)).to_list())
    ctx = D.DryadLinqContext(2, device="cuda")
    got = sorted(ctx.from_enumerable(left, 2).join(ctx.from_enumerable([1, 3], 1), lambda a: a, lambda b: b, lambda a, b: (a, b)).to_list())
    assert got == want and len(got) == 16


def test_group_join_and_sliding_window_run_on_the_gpu():
    outer = [{"id": i, "w": i % 5} for i in range(300)]
    inner = [{"fk": (i * 7) % 330, "v": i} for i in range(4000)]        # ids 300..329 have no outer row; some outers have no match

    def gj(c, parts):
        o, n = c.from_enumerable(outer, parts), c.from_enumerable(inner, parts)
        return o.group_join(n, lambda r: r.id, lambda s: s.fk,
                            lambda r, g: {"id": r.id, "n": g.count(), "s": g.sum(lambda s: s.v) + r.w})
    want = sorted(gj(D.DryadLinqContext(1, local_debug=True), 1).to_list(), key=repr)
    for parts in (1, 3):
        ctx = D.DryadLinqContext(parts, device="cuda")
        got = sorted(gj(ctx, parts).to_list(), key=repr)
        assert got == want and ctx._backend().host_fallbacks == 0, parts
    data = [(i * 31) % 97 for i in range(2000)]
    _same_order(lambda q: q.sliding_window(lambda w: sum(w), 4), data)
    _same_order(lambda q: q.sliding_window(lambda w: w[-1] - w[0], 7), data)
    _same_order(lambda q: q.sliding_window(lambda w: (w[0], w[1] * 2), 2), data)
