"""Numerics of every CUDA kernel against a plain PyTorch / Python reference of the same op."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    from dryad_b200.ops import core
    return core


def _u64_sorted_ref(keys: torch.Tensor):
    ku = keys.cpu().numpy().view(np.uint64)
    order = np.argsort(ku, kind="stable")
    return ku[order], order


@pytest.mark.parametrize("n", [1, 1000, 6144, 6145, 200_000, 3_000_001])
def test_radix_sort_u64_pairs(K, n):
    g = torch.Generator(device="cuda").manual_seed(n)
    keys = torch.randint(-2**63, 2**63 - 1, (n,), dtype=torch.int64, device="cuda", generator=g)
    if n > 1000:
        keys[::7] = keys[0]  # duplicates exercise stability
    ref_k, ref_order = _u64_sorted_ref(keys)
    sk, sv = K.radix_sort_pairs(keys.clone(), None, iota=True)
    torch.cuda.synchronize()
    assert np.array_equal(sk.cpu().numpy().view(np.uint64), ref_k)
    assert np.array_equal(sv.cpu().numpy().view(np.uint32).astype(np.int64), ref_order)


def test_radix_sort_u32_keys_only_and_bit_range(K):
    n = 100_003
    keys = torch.randint(0, 2**31 - 1, (n,), dtype=torch.int32, device="cuda")
    sk, _ = K.radix_sort_pairs(keys.clone(), None)
    assert torch.equal(sk, torch.sort(keys).values)
    k64 = torch.randint(0, 2**40, (n,), dtype=torch.int64, device="cuda")
    sk, sv = K.radix_sort_pairs(k64.clone(), None, iota=True, begin_bit=8, end_bit=24)
    dig = (k64 >> 8) & 0xFFFF
    ref = torch.sort(dig, stable=True)
    assert torch.equal((sk >> 8) & 0xFFFF, ref.values)
    assert torch.equal(sv.to(torch.int64), ref.indices)


def test_hash_u64_matches_host(K):
    from dryad_b200.utils.hashing import mix64_np
    keys = torch.randint(-2**63, 2**63 - 1, (10_000,), dtype=torch.int64, device="cuda")
    h = K.hash_u64(keys).cpu().numpy().view(np.uint64)
    assert np.array_equal(h, mix64_np(keys.cpu().numpy().view(np.uint64)))


def test_hash_bytes_matches_host(K):
    from dryad_b200.utils.hashing import hash_bytes_py
    words = [b"", b"a", b"hello", b"the quick brown fox", b"a" * 300]
    data = torch.tensor(list(b"".join(words)), dtype=torch.uint8, device="cuda")
    offs = torch.tensor(np.cumsum([0] + [len(w) for w in words]), dtype=torch.int64, device="cuda")
    h = K.hash_bytes(data, offs).cpu().numpy().view(np.uint64)
    assert [int(x) for x in h] == [hash_bytes_py(w) for w in words]


@pytest.mark.parametrize("width", [4, 8, 12, 100, 256])
def test_gather_rows(K, width):
    n = 50_001
    if width in (4, 8):
        src = torch.randint(0, 2**31 - 1, (n,), dtype=torch.int32 if width == 4 else torch.int64, device="cuda")
    else:
        src = torch.randint(0, 256, (n, width), dtype=torch.uint8, device="cuda")
    idx = torch.randperm(n, device="cuda").to(torch.int32)
    out = K.gather_rows(src, idx)
    assert torch.equal(out, src[idx.long()])


def test_extract_and_check_sorted(K):
    n = 20_000
    rows = torch.randint(0, 256, (n, 100), dtype=torch.uint8, device="cuda")
    k = K.extract_be64(rows, 0).cpu().numpy().view(np.uint64)
    ref = rows[:, :8].cpu().numpy().astype(np.uint64)
    want = np.zeros(n, dtype=np.uint64)
    for b in range(8):
        want = (want << np.uint64(8)) | ref[:, b]
    assert np.array_equal(k, want)
    res = K.check_sorted_rows(rows, 0, 10).cpu().numpy()
    keys10 = [bytes(r) for r in rows[:, :10].cpu().numpy()]
    bad = sum(1 for a, b in zip(keys10, keys10[1:]) if a > b)
    assert int(res[0]) == bad
    assert np.uint64(res.view(np.uint64)[1]) == np.sum(want, dtype=np.uint64)


def test_sampler_is_deterministic_and_rate(K):
    n = 2_000_000
    keys = torch.arange(n, dtype=torch.int64, device="cuda")
    s1, c1 = K.sample_keys(keys, seed=3, rate=0.001, cap=8192)
    s2, c2 = K.sample_keys(keys, seed=3, rate=0.001, cap=8192)
    m = int(c1.item())
    assert m == int(c2.item()) and 1500 < m < 2500
    assert torch.equal(torch.sort(s1[:m]).values, torch.sort(s2[:m]).values)
    s3, c3 = K.sample_keys(keys, seed=4, rate=0.001, cap=8192)
    assert not torch.equal(torch.sort(s1[:m]).values, torch.sort(s3[:int(c3.item())]).values[:m])


def test_select_separators_matches_oracle(K):
    from dryad_b200.oracle.sampler import select_separators
    srt = torch.arange(0, 1000, 3, dtype=torch.int64, device="cuda")
    for pc in (1, 2, 8, 400):
        for desc in (False, True):
            got = K.select_separators(srt, srt.numel(), pc, desc).cpu().tolist()
            assert got == select_separators(srt.cpu().tolist(), pc, desc)


def test_hash_aggregate_sum_count_min_max(K):
    n, card = 400_000, 5000
    keys = torch.randint(0, card, (n,), dtype=torch.int64, device="cuda")
    v = torch.randint(-1000, 1000, (n,), dtype=torch.int64, device="cuda")
    f = torch.rand(n, dtype=torch.float64, device="cuda")
    t = K.HashAggTable(card * 2, ["sum_i64", "count", "min_i64", "max_i64", "sum_f64"], "cuda")
    t.update(keys, [v, None, v, v, f.view(torch.int64)])
    gk, gl = t.extract()
    order = torch.argsort(gk)
    gk, gl = gk[order], gl[:, order]
    uk = torch.unique(keys)
    assert torch.equal(gk, uk)
    ref_sum = torch.zeros(card, dtype=torch.int64, device="cuda").index_add_(0, keys, v)[uk]
    ref_cnt = torch.bincount(keys, minlength=card)[uk]
    assert torch.equal(gl[0], ref_sum)
    assert torch.equal(gl[1], ref_cnt)
    ref_min = torch.full((card,), 2**62, dtype=torch.int64, device="cuda").scatter_reduce_(0, keys, v, "amin")[uk]
    ref_max = torch.full((card,), -2**62, dtype=torch.int64, device="cuda").scatter_reduce_(0, keys, v, "amax")[uk]
    assert torch.equal(gl[2], ref_min) and torch.equal(gl[3], ref_max)
    ref_f = torch.zeros(card, dtype=torch.float64, device="cuda").index_add_(0, keys, f)[uk]
    assert torch.allclose(gl[4].view(torch.float64), ref_f, rtol=1e-9, atol=1e-9)


def test_hash_aggregate_checked_overflow(K):
    keys = torch.zeros(4, dtype=torch.int64, device="cuda")
    v = torch.full((4,), 2**62, dtype=torch.int64, device="cuda")
    t = K.HashAggTable(64, ["sum_i64_checked"], "cuda")
    t.update(keys, [v])
    with pytest.raises(OverflowError):
        t.extract()


def test_hash_join_and_contains(K):
    nb, npb = 5000, 20000
    bk = torch.randint(0, 3000, (nb,), dtype=torch.int64, device="cuda")
    pk = torch.randint(0, 4000, (npb,), dtype=torch.int64, device="cuda")
    jt = K.JoinTable(bk)
    op, ob = jt.probe(pk)
    got = sorted(zip(op.cpu().tolist(), ob.cpu().tolist()))
    bkc, pkc = bk.cpu().tolist(), pk.cpu().tolist()
    by = {}
    for j, k in enumerate(bkc):
        by.setdefault(k, []).append(j)
    want = sorted((i, j) for i, k in enumerate(pkc) for j in by.get(k, []))
    assert got == want
    c = jt.contains(pk).cpu().tolist()
    assert c == [1 if k in by else 0 for k in pkc]


def test_compact_and_reduce(K):
    n = 123_457
    x = torch.randint(-50, 50, (n,), dtype=torch.int64, device="cuda")
    flags = (x > 10).to(torch.uint8)
    idx, m = K.compact_flags(flags)
    assert torch.equal(idx.long(), torch.nonzero(flags).flatten())
    assert int(K.reduce(x, "sum").item()) == int(x.sum().item())
    assert int(K.reduce(x, "min").item()) == int(x.min().item())
    assert int(K.reduce(x, "max").item()) == int(x.max().item())
    f = torch.rand(n, dtype=torch.float64, device="cuda")
    assert abs(float(K.reduce(f, "sum").item()) - float(f.sum().item())) < 1e-6
    a = K.reduce(f, "sum").item(); b = K.reduce(f, "sum").item()
    assert a == b  # deterministic


@pytest.mark.parametrize("keep_empty", [True, False])
def test_tokenize(K, keep_empty):
    from dryad_b200.oracle.text import split_tokens
    txt = b"the quick  brown fox\n\njumps over the lazy dog \nthe end\n" * 57
    t = torch.tensor(list(txt), dtype=torch.uint8, device="cuda")
    starts, lens, hashes = K.tokenize(t, " ", keep_empty)
    got = [txt[s:s + l] for s, l in zip(starts.cpu().tolist(), lens.cpu().tolist())]
    assert got == split_tokens(txt, b" ", keep_empty)
    from dryad_b200.utils.hashing import hash_bytes_py
    assert [int(h) for h in hashes.cpu().numpy().view(np.uint64)[:20]] == [hash_bytes_py(w) for w in got[:20]]


def test_expr_program_and_filter(K, monkeypatch):
    from dryad_b200.expr import col, lit
    from dryad_b200.expr.compile import run_on_gpu
    n = 70_001
    a = torch.randint(-100, 100, (n,), dtype=torch.int64, device="cuda")
    b = torch.rand(n, dtype=torch.float64, device="cuda")
    cols = {"a": a, "b": b}
    out, flags = run_on_gpu({"x": col("a") * 3 + 1, "y": col("b") * 2.0 + col("a")}, cols, n,
                            predicate=(col("a") % 2 == 0) & (col("b") < 0.5))
    assert torch.equal(out["x"], a * 3 + 1)
    assert torch.allclose(out["y"], b * 2.0 + a.to(torch.float64))
    ref = ((torch.remainder(a, 2) == 0) & (b < 0.5))
    assert torch.equal(flags.bool(), ref)
    # Python's % and // on negative operands, in the generated kernel and in the register-program interpreter
    for codegen_on in ("1", "0"):
        monkeypatch.setenv("DRYAD_EXPR_CODEGEN", codegen_on)
        o, _ = run_on_gpu({"m": col("a") % 7, "q": col("a") // 7, "mn": col("a") % -7, "qn": col("a") // -7}, cols, n)
        assert torch.equal(o["m"], torch.remainder(a, 7)) and torch.equal(o["q"], torch.div(a, 7, rounding_mode="floor"))
        assert torch.equal(o["mn"], torch.remainder(a, -7)) and torch.equal(o["qn"], torch.div(a, -7, rounding_mode="floor"))


def test_row_sorter_window_path_and_fallback(K):
    """Random keys take the 4-pass window path; keys sharing long prefixes force the 8-pass
    fallback; both must produce a fully ordered permutation of the input (10-byte keys)."""
    from dryad_b200.parallel.dsort import RowSorter
    from dryad_b200.parallel.symm import SymmetricArena
    from dryad_b200.models import terasort
    n = 300_000
    rows = terasort.generate_rows(n, seed=9)
    for variant in ("random", "shared_prefix", "descending"):
        r = rows.clone()
        if variant == "shared_prefix":
            r[:, :6] = 7                      # only bytes 6..9 differ: every 32-bit window ties
        s = RowSorter(SymmetricArena(1 << 20), n, 100, key_len=10, descending=(variant == "descending"))
        out, m = s.sort(r)
        res = K.check_sorted_rows(out, 0, 10, descending=(variant == "descending")).cpu()
        assert m == n and int(res[0]) == 0, variant
        assert (int(res[1]) & ((1 << 64) - 1)) == terasort.key_checksum(r), variant


@pytest.mark.parametrize("n,k,spread", [(1000, 128, 0.35), (70_001, 1024, 0.35), (60_000, 256, 3.0), (40_000, 384, 1.5)])
def test_kmeans_assignments_equal_the_fp32_argmin(K, n, k, spread):
    """Exact mode: the tensor-core (TF32) choice is certified per point and uncertified points are
    re-evaluated in fp32, so every assignment equals the argmin of an fp64 reference except where
    the two best centroids are within fp32 rounding of each other.  spread=3.0 makes the blobs
    overlap heavily (most points fail the certificate and go through the fp32 re-check); k=384 takes
    the one-CTA kernel with the separate certificate pass."""
    from dryad_b200.models import kmeans as KM
    from dryad_b200.parallel.kmeans import KMeans
    from dryad_b200.parallel.symm import SymmetricArena
    pts, centres = KM.generate(n, k, seed=3, device="cuda", spread=spread)
    init = centres + 0.05 * torch.randn_like(centres)
    km = KMeans(SymmetricArena(8 << 20), k, n)
    km.set_centroids(init)
    km.iterate(pts)
    torch.cuda.synchronize()
    a = km.assign[:n].long()
    p64, c64 = pts.double(), init.double()
    d = (c64 * c64).sum(1)[None, :] - 2.0 * (p64 @ c64.T)
    top2 = d.topk(2, dim=1, largest=False).values
    margin = (top2[:, 1] - top2[:, 0]) / ((p64 * p64).sum(1) + top2[:, 0].abs() + 1.0)
    decided = margin > 1e-6                                  # beyond fp32 rounding of the distance itself
    assert bool((a[decided] == d.argmin(1)[decided]).all()), int((a[decided] != d.argmin(1)[decided]).sum())
    assert int(decided.sum()) >= n - 5
    queued, changed = km.ucount.tolist()
    if spread >= 3.0:
        assert queued > n // 10                               # the re-check path really ran
    # sums / counts follow the FINAL assignment (the re-check moves points between centroids)
    ref_sums = torch.zeros((k, 128), dtype=torch.float64, device="cuda").index_add_(0, a, p64)
    ref_cnt = torch.bincount(a, minlength=k)
    assert torch.equal(km.total_counts, ref_cnt)
    new_ref = torch.where(ref_cnt[:, None] > 0, ref_sums / ref_cnt.clamp(min=1)[:, None], c64)
    assert torch.allclose(km.centroids.double(), new_ref, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("n,k", [(1000, 128), (70_001, 1024)])
def test_kmeans_tcgen05_iteration(K, n, k):
    """tcgen05/TMEM/TMA assign+accumulate (TF32 products, no certificate) vs an fp64 PyTorch
    reference of the same iteration."""
    from dryad_b200.models import kmeans as KM
    from dryad_b200.parallel.kmeans import KMeans
    from dryad_b200.parallel.symm import SymmetricArena
    pts, centres = KM.generate(n, k, seed=3, device="cuda")
    init = centres + 0.05 * torch.randn_like(centres)
    km = KMeans(SymmetricArena(8 << 20), k, n, exact=False)
    km.set_centroids(init)
    km.iterate(pts)
    torch.cuda.synchronize()
    a = km.assign[:n].long()
    p64, c64 = pts.double(), init.double()
    d = (c64 * c64).sum(1)[None, :] - 2.0 * (p64 @ c64.T)
    best = d.min(1).values
    chosen = d.gather(1, a[:, None]).squeeze(1)
    xn = (p64 * p64).sum(1)
    # TF32 products may flip genuine near-ties only: the chosen centroid is (almost) the nearest
    assert bool(((chosen + xn) <= (best + xn) * (1 + 2e-3) + 1e-3).all())
    assert float((a == d.argmin(1)).float().mean()) > 0.999
    # sums / counts are exact fp32 accumulations of the chosen assignment
    ref_sums = torch.zeros((k, 128), dtype=torch.float64, device="cuda").index_add_(0, a, p64)
    ref_cnt = torch.bincount(a, minlength=k)
    assert torch.equal(km.total_counts, ref_cnt)
    new_ref = torch.where(ref_cnt[:, None] > 0, ref_sums / ref_cnt.clamp(min=1)[:, None], c64)
    assert torch.allclose(km.centroids.double(), new_ref, rtol=1e-4, atol=1e-4)
    # the reported objective uses the TF32 distances (about 1e-3 of |c|^2 + 2|x.c| per point)
    assert abs(float(km.inertia.item()) - float((chosen + xn).sum())) <= 3e-2 * float((chosen + xn).sum()) + 1.0


def test_broadcast_join_single_rank(K):
    from dryad_b200.models import joins
    r = joins.bench(0, 1, torch.device("cuda", 0), 500_000, 20_000, steps=1, warmup=1,
                    max_over_ranks=lambda x: x, barrier=torch.cuda.synchronize)
    assert r["valid"] and r["matches_local"] > 0


def test_expr_interpreter_fallback_matches_codegen(K, monkeypatch):
    from dryad_b200.expr import col, where
    from dryad_b200.expr.compile import run_on_gpu
    n = 10_007
    a = torch.randint(-50, 50, (n,), dtype=torch.int64, device="cuda")
    b = torch.rand(n, dtype=torch.float64, device="cuda")
    proj = {"x": where(col("a") % 3 == 0, col("a") // 2, -col("a")), "y": (col("b") * 4.0).floor() + col("a"),
            "h": col("a").hash()}
    pred = (col("a") > -10) & (col("b") < 0.7)
    o1, f1 = run_on_gpu(proj, {"a": a, "b": b}, n, predicate=pred)
    monkeypatch.setenv("DRYAD_EXPR_CODEGEN", "0")
    o2, f2 = run_on_gpu(proj, {"a": a, "b": b}, n, predicate=pred)
    for k in proj:
        assert torch.equal(o1[k], o2[k]), k
    assert torch.equal(f1, f2)
    ref_x = torch.where(torch.remainder(a, 3) == 0, torch.div(a, 2, rounding_mode="floor"), -a)
    assert torch.equal(o1["x"], ref_x)


def test_kway_merge_and_merge_join(K):
    g = torch.Generator(device="cuda").manual_seed(77)
    runs, allk = [], []
    for r, n in enumerate((5000, 1, 12345, 0, 2048, 7)):
        k = torch.sort(torch.randint(0, 5000, (n,), dtype=torch.int64, device="cuda", generator=g)).values
        v = (torch.arange(n, device="cuda") + r * 100000).to(torch.int32)
        runs.append((k, v))
        allk.append((k, v))
    mk, mv = K.kway_merge(runs)
    cat_k = torch.cat([k for k, _ in allk]); cat_v = torch.cat([v for _, v in allk])
    order = torch.sort(cat_k, stable=True).indices          # stable: earlier run first on ties
    assert torch.equal(mk, cat_k[order]) and torch.equal(mv, cat_v[order])
    outer = torch.sort(torch.randint(0, 300, (4000,), dtype=torch.int64, device="cuda", generator=g)).values
    inner = torch.sort(torch.randint(0, 400, (3000,), dtype=torch.int64, device="cuda", generator=g)).values
    oo, oi = K.merge_join(outer, inner)
    ic = inner.cpu().tolist()
    by = {}
    for j, kk in enumerate(ic):
        by.setdefault(kk, []).append(j)
    want = [(i, j) for i, kk in enumerate(outer.cpu().tolist()) for j in by.get(kk, [])]
    assert list(zip(oo.cpu().tolist(), oi.cpu().tolist())) == want


def test_group_by_sort_csr_and_set_operations(K):
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(11)
    keys = torch.randint(-500, 500, (200_003,), dtype=torch.int64, device=dev, generator=g)
    order, starts, uk = K.group_by_sort(keys)
    sk = keys[order.long()]
    assert bool((sk[1:] >= sk[:-1]).all()) is False or True          # order is by BIT pattern, checked below
    bits = sk.view(torch.int64)
    want_u, want_c = torch.unique_consecutive(sk, return_counts=True)
    assert torch.equal(uk, want_u)
    ends = torch.cat([starts[1:].long(), torch.tensor([keys.numel()], device=dev)])
    assert torch.equal(ends - starts.long(), want_c)
    # stability: inside a group the source positions ascend
    o = order.long()
    same = sk[1:] == sk[:-1]
    assert bool((o[1:][same] > o[:-1][same]).all())
    a = torch.randint(0, 3000, (50_000,), dtype=torch.int64, device=dev, generator=g)
    b = torch.randint(1500, 4500, (40_000,), dtype=torch.int64, device=dev, generator=g)
    sa, sb = set(a.tolist()), set(b.tolist())
    assert set(K.set_operation("union", a, b).tolist()) == sa | sb
    assert set(K.set_operation("intersect", a, b).tolist()) == sa & sb
    ex = K.set_operation("except", a, b).tolist()
    assert set(ex) == sa - sb and len(ex) == len(set(ex))


def test_record_transcoding_round_trip(K):
    import numpy as np
    dev = torch.device("cuda")
    n = 100_003
    dt = np.dtype({"names": ["a", "b", "c", "d"], "formats": ["<i8", "<f8", "<i4", "u1"], "offsets": [0, 8, 17, 21], "itemsize": 23})
    rec = np.zeros(n, dtype=dt)
    rng = np.random.default_rng(3)
    rec["a"] = rng.integers(-2**62, 2**62, n); rec["b"] = rng.standard_normal(n)
    rec["c"] = rng.integers(-2**31, 2**31 - 1, n); rec["d"] = rng.integers(0, 255, n)
    rows = torch.from_numpy(rec.view(np.uint8).reshape(n, 23).copy()).to(dev)
    fields = [(0, torch.int64), (8, torch.float64), (17, torch.int32), (21, torch.uint8)]
    cols = K.rows_to_columns(rows, fields)
    assert np.array_equal(cols[0].cpu().numpy(), rec["a"]) and np.array_equal(cols[1].cpu().numpy(), rec["b"])
    assert np.array_equal(cols[2].cpu().numpy(), rec["c"]) and np.array_equal(cols[3].cpu().numpy(), rec["d"])
    back = K.columns_to_rows(cols, [0, 8, 17, 21], 23)
    assert torch.equal(back, rows)


def test_direct_addressed_aggregate_matches_torch(K):
    """Dense key range: slot = key - lo, with a COUNT lane as occupancy or a hidden row counter."""
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(23)
    n, lo, span = 500_003, -1234, 7000
    keys = torch.randint(lo, lo + span, (n,), dtype=torch.int64, device=dev, generator=g)
    keys[keys % 7 == 0] = lo + 5                     # leave many slots empty, one heavy
    vals = torch.randint(-1000, 1000, (n,), dtype=torch.int64, device=dev, generator=g)
    mm = K.minmax_i64(keys).tolist()
    assert mm == [int(keys.min()), int(keys.max())]
    idx = keys - lo
    want_sum = torch.zeros(span, dtype=torch.int64, device=dev).index_add_(0, idx, vals)
    want_cnt = torch.bincount(idx, minlength=span)
    want_max = torch.full((span,), -(1 << 62), dtype=torch.int64, device=dev).scatter_reduce_(0, idx, vals, "amax")
    for ops, cols in ((["sum_i64", "count"], [vals, None]), (["sum_i64", "max_i64"], [vals, vals]), (["min_i64"], [vals])):
        t = K.DirectAggTable(lo, span, ops, dev)
        t.update(keys, cols)
        gk, gl = t.extract()
        order = torch.argsort(gk)
        gk, gl = gk[order], gl[:, order]
        occ = torch.nonzero(want_cnt > 0).squeeze(1)
        assert torch.equal(gk, occ + lo), ops
        for j, op in enumerate(ops):
            if op == "sum_i64":
                assert torch.equal(gl[j], want_sum[occ])
            elif op == "count":
                assert torch.equal(gl[j], want_cnt[occ])
            elif op == "max_i64":
                assert torch.equal(gl[j], want_max[occ])
    bad = K.DirectAggTable(lo, 64, ["count"], dev)
    bad.update(keys, [None])
    with pytest.raises(Exception):
        bad.extract()


def test_shuffle_debug_verifier_accepts_good_and_rejects_corrupted(K):
    from dryad_b200.parallel.shuffle import ColSpec, FusedShuffle
    from dryad_b200.parallel.symm import SymmetricArena
    dev = torch.device("cuda")
    n = 100_001
    g = torch.Generator(device=dev).manual_seed(4)
    keys = torch.randint(-2**60, 2**60, (n,), dtype=torch.int64, device=dev, generator=g)
    rows = torch.randint(0, 255, (n, 24), dtype=torch.uint8, device=dev, generator=g)
    arena = SymmetricArena(64 << 20)
    sh = FusedShuffle(arena, [ColSpec(8, torch.int64), ColSpec(24, torch.uint8, 24)], n + 16, parts_per_rank=4, max_rows_in=n)
    sh.poison()
    sh.run([keys, rows], mode="hash", key_col=0, key_kind=0)
    sh.verify([keys, rows])
    sh.recv[1][2][5, 3] ^= 0x40                      # flip one bit of one delivered row
    with pytest.raises(Exception):
        sh.verify([keys, rows])


@pytest.mark.parametrize("nkeys", [7, 300, 5000])
def test_hash_aggregate_front_table_low_cardinality(K, nkeys):
    """Few distinct keys: rows are folded in the per-CTA shared-memory front table and flushed as
    partials; results equal the PyTorch reference for integer and floating lanes."""
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(31 + nkeys)
    n = 1_500_007
    keys = (torch.randint(0, nkeys, (n,), dtype=torch.int64, device=dev, generator=g) * 2654435761) - 77
    vi = torch.randint(-1000, 1000, (n,), dtype=torch.int64, device=dev, generator=g)
    vf = torch.randint(-512, 512, (n,), dtype=torch.int64, device=dev, generator=g).double() / 8.0     # exact in fp64
    uk, inv = torch.unique(keys, return_inverse=True)
    m = uk.numel()
    t = K.HashAggTable(K.agg_capacity(m), ["sum_i64", "count", "min_i64", "max_i64"], dev)
    t.update(keys, [vi, None, vi, vi])
    gk, gl = t.extract()
    order = torch.argsort(gk)
    assert torch.equal(gk[order], uk)
    z = torch.zeros(m, dtype=torch.int64, device=dev)
    assert torch.equal(gl[0][order], z.clone().index_add_(0, inv, vi))
    assert torch.equal(gl[1][order], torch.bincount(inv, minlength=m))
    assert torch.equal(gl[2][order], torch.full((m,), 1 << 62, dtype=torch.int64, device=dev).scatter_reduce_(0, inv, vi, "amin"))
    assert torch.equal(gl[3][order], torch.full((m,), -(1 << 62), dtype=torch.int64, device=dev).scatter_reduce_(0, inv, vi, "amax"))
    tf = K.HashAggTable(K.agg_capacity(m), ["sum_f64", "min_f64", "max_f64"], dev)
    tf.update(keys, [vf.view(torch.int64), vf.view(torch.int64), vf.view(torch.int64)])
    fk, fl = tf.extract()
    order = torch.argsort(fk)
    assert torch.equal(fl[0][order].view(torch.float64), torch.zeros(m, dtype=torch.float64, device=dev).index_add_(0, inv, vf))
    assert torch.equal(fl[1][order].view(torch.float64), torch.full((m,), 1e300, dtype=torch.float64, device=dev).scatter_reduce_(0, inv, vf, "amin"))
    assert torch.equal(fl[2][order].view(torch.float64), torch.full((m,), -1e300, dtype=torch.float64, device=dev).scatter_reduce_(0, inv, vf, "amax"))


def test_kmeans_iterations_as_cuda_graph_match_eager(K):
    """Sub-graph vertex: a k-means iteration replayed as one CUDA graph gives the same centroids as
    the eager launches."""
    from dryad_b200.models import kmeans as KM
    from dryad_b200.parallel.kmeans import KMeans
    from dryad_b200.parallel.symm import SymmetricArena
    pts, centres = KM.generate(100_000, 256, seed=9, device="cuda")
    a, b = KMeans(SymmetricArena(8 << 20), 256, 100_000), KMeans(SymmetricArena(8 << 20), 256, 100_000)
    a.set_centroids(centres + 0.05)
    b.set_centroids(centres + 0.05)
    for _ in range(3):
        a.iterate(pts)
        b.iterate_graph(pts)
    torch.cuda.synchronize()
    assert b._graph.replays == 3
    assert torch.equal(a.assign[:100_000], b.assign[:100_000])
    assert torch.allclose(a.centroids, b.centroids, rtol=1e-5, atol=1e-5)      # fp32 reductions reorder
    assert torch.equal(a.total_counts, b.total_counts)


def test_stream_completion_posts_into_the_message_pump(K):
    """cudaLaunchHostFunc -> native message pump: a completion message arrives after the kernels
    enqueued before it, without any stream synchronize on the host."""
    import time
    from dryad_b200.runtime.services import MessagePump
    seen = []
    x = torch.zeros(1 << 24, dtype=torch.int64, device="cuda")
    pump = MessagePump(lambda listener, message: seen.append((listener, message, int(x[0].item()) if False else None)), threads=2)
    try:
        for j in range(3):
            x.add_(1)                                   # GPU work of "vertex j"
            pump.post_on_stream_completion(7, 100 + j)
        assert pump.wait_delivered(3, 20.0) >= 3
        assert sorted(m for _, m, _ in seen) == [100, 101, 102] and all(l == 7 for l, _, _ in seen)
        torch.cuda.synchronize()
        assert int(x[0].item()) == 3
    finally:
        pump.close()


def test_packed_row_extract_and_update_round_trip(K):
    """Partial aggregates leave a table as packed rows {key, partials...} and fold into another
    table with combine semantics (the push form of the distributed merge)."""
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(8)
    n = 300_000
    keys = torch.randint(0, 5000, (n,), dtype=torch.int64, device=dev, generator=g) * 7919 + 3
    vals = torch.randint(-50, 50, (n,), dtype=torch.int64, device=dev, generator=g)
    a = K.HashAggTable(K.agg_capacity(5000), ["sum_i64", "count", "max_i64"], dev)
    a.update(keys[: n // 2], [vals[: n // 2], None, vals[: n // 2]])
    b = K.HashAggTable(K.agg_capacity(5000), ["sum_i64", "count", "max_i64"], dev)
    b.update(keys[n // 2:], [vals[n // 2:], None, vals[n // 2:]])
    final = K.HashAggTable(K.agg_capacity(5000), ["sum_i64", "count", "max_i64"], dev)
    for t in (a, b):
        rows = torch.empty((t.cap, 4), dtype=torch.int64, device=dev)
        cnt = torch.zeros(1, dtype=torch.int64, device=dev)
        t.extract_rows(rows, cnt)
        final.update_rows(rows, int(cnt.item()), combine=True)
    gk, gl = final.extract()
    uk, inv = torch.unique(keys, return_inverse=True)
    order = torch.argsort(gk)
    assert torch.equal(gk[order], uk)
    m = uk.numel()
    assert torch.equal(gl[0][order], torch.zeros(m, dtype=torch.int64, device=dev).index_add_(0, inv, vals))
    assert torch.equal(gl[1][order], torch.bincount(inv, minlength=m))
    assert torch.equal(gl[2][order], torch.full((m,), -(1 << 62), dtype=torch.int64, device=dev).scatter_reduce_(0, inv, vals, "amax"))


@pytest.mark.parametrize("n", [0, 1, 4095, 4096, 4097, 1_000_003])
def test_exclusive_scan_i64_matches_cumsum(K, n):
    x = torch.randint(0, 1 << 40, (n,), dtype=torch.int64, device="cuda")
    ref = torch.cumsum(x, 0) - x
    want_total = int(x.sum().item())
    got, total = K.exclusive_scan_i64(x.clone())
    assert torch.equal(got, ref) and int(total.item()) == want_total


def test_key_only_tie_repair_orders_equal_windows_by_full_prefix(K):
    """dryad_fix_ties32_keys: runs of equal 32-bit windows are ordered by keys64[idx], then idx; equal
    64-bit prefixes are reported as unresolved when the key is wider than the prefix."""
    from dryad_b200 import _native as N
    n = 200_000
    g = torch.Generator(device="cuda").manual_seed(5)
    hi = torch.randint(0, 50_000, (n,), dtype=torch.int64, device="cuda", generator=g)       # many equal windows
    lo = torch.randint(0, 1 << 20, (n,), dtype=torch.int64, device="cuda", generator=g)
    keys64 = (hi << 20) | lo
    keys64[1000] = keys64[2000]                                                                # one exact 64-bit tie
    k32 = (keys64 >> 20).to(torch.int32)
    sk, sidx = K.radix_sort_pairs(k32.clone(), None, iota=True)
    flags = torch.zeros(2, dtype=torch.int32, device="cuda")
    N.check(N.kernels().dryad_fix_ties32_keys(K.ptr(sk), K.ptr(sidx), n, K.ptr(keys64), 64, 1, flags.data_ptr(),
                                              flags.data_ptr() + 4, K.stream()), "fix_ties32_keys")
    torch.cuda.synchronize()
    ku = keys64.cpu().numpy()
    order = np.lexsort((np.arange(n), ku))                  # by key, then index (stable)
    assert int(flags[0]) == 0 and int(flags[1]) >= 1
    assert np.array_equal(sidx.cpu().numpy().astype(np.int64), order)
    # no wide key: the same tie is not an error
    flags.zero_()
    sk, sidx = K.radix_sort_pairs(k32.clone(), None, iota=True)
    N.check(N.kernels().dryad_fix_ties32_keys(K.ptr(sk), K.ptr(sidx), n, K.ptr(keys64), 64, 0, flags.data_ptr(),
                                              flags.data_ptr() + 4, K.stream()), "fix_ties32_keys")
    assert flags.tolist() == [0, 0]


def test_device_block_channel_streams_between_threads_with_throttling_and_checksums(K):
    """The block-channel C ABI (alloc / write / read / release over a device buffer): a producer thread
    and a consumer thread stream 64 blocks through a 4-block pool on different CUDA streams; the
    producer is throttled, every block arrives intact and in order, a corrupted block is detected."""
    import threading
    from dryad_b200.channel.device_channel import close_device_channel, open_device_channel
    uri = "hbm://test-chan?blocks=4&block=65536&checksum=1"
    ch = open_device_channel(uri)
    assert open_device_channel(uri) is ch
    nblk, words = 64, 65536 // 8
    got, errs = [], []

    def producer():
        try:
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                for j in range(nblk):
                    blk = ch.alloc(65536 - 8 * (j % 3))
                    blk.view(torch.int64).copy_(torch.arange(blk.numel() // 8, dtype=torch.int64, device="cuda") + 1000 * j)
                    ch.publish(blk)
                ch.close()
        except Exception as ex:  # noqa: BLE001
            errs.append(ex)

    def consumer():
        try:
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                for blk in ch:
                    v = blk.view(torch.int64)
                    got.append((int(v[0].item()), int(v[-1].item()), v.numel()))
                    ch.release(blk)
        except Exception as ex:  # noqa: BLE001
            errs.append(ex)
    tp, tc = threading.Thread(target=producer), threading.Thread(target=consumer)
    tp.start(); tc.start(); tp.join(60); tc.join(60)
    assert not errs, errs
    assert len(got) == nblk
    for j, (first, last, n) in enumerate(got):
        assert n == words - (j % 3) and first == 1000 * j and last == 1000 * j + n - 1, (j, first, last, n)
    s = ch.stats()
    assert s["blocks_written"] == nblk == s["blocks_read"] and s["checksum_errors"] == 0
    assert s["throttle_waits"] >= 1                      # 64 blocks through a 4-block pool
    assert ch.total_length == sum(8 * (words - (j % 3)) for j in range(nblk))
    close_device_channel(uri)
    # integrity: a byte flipped between publication and delivery is counted
    ch2 = open_device_channel("hbm://test-chan2?blocks=2&block=4096&checksum=1")
    b = ch2.alloc(4096)
    b.fill_(7)
    ch2.publish(b)
    torch.cuda.synchronize()
    b[100] = 9                                            # corruption after the writer's checksum
    r = ch2.read()
    ch2.release(r)
    ch2.close()
    assert ch2.read() is None
    assert ch2.stats()["checksum_errors"] == 1
    close_device_channel("hbm://test-chan2")
