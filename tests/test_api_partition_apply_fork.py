"""RangePartition API coverage, Apply / Fork, misc regressions (mirrors
DryadLinqTests/RangePartitionAPICoverageTests.cs, ApplyAndForkTests.cs, MiscBugFixTests.cs)."""
import pytest

import dryad_b200 as D
from dryad_b200 import DryadLinqException, ErrorCode
from dryad_b200.utils.hashing import hash_bucket
from util import check_both

VALS = [(i * 7919) % 3000 for i in range(3000)]    # 3 x 1000 so the 0.1 % sampler sees data


def _range_partitioned(parts, descending=False):
    ne = [p for p in parts if p]
    for a, b in zip(ne, ne[1:]):
        if descending:
            assert min(a) >= max(b)
        else:
            assert max(a) <= min(b)


@pytest.mark.parametrize("desc", [False, True])
def test_range_partition_sampled(desc):
    ctx = D.DryadLinqContext(3)
    parts = ctx.from_enumerable(VALS, 3).range_partition(lambda x: x, 4, descending=desc).partitions()
    assert len(parts) == 4 and sorted(sum(parts, [])) == sorted(VALS)
    _range_partitioned(parts, desc)


def test_range_partition_explicit_keys_and_inference():
    ctx = D.DryadLinqContext(3)
    parts = ctx.from_enumerable(VALS, 3).range_partition(lambda x: x, range_keys=[1000, 2000]).partitions()
    assert len(parts) == 3
    assert all(x <= 1000 for x in parts[0]) and all(1000 < x <= 2000 for x in parts[1]) and all(x > 2000 for x in parts[2])
    d = ctx.from_enumerable(VALS, 3).range_partition(lambda x: x, range_keys=[2000, 1000]).partitions()
    assert all(x >= 2000 for x in d[0]) and all(x < 1000 for x in d[2])     # direction inferred from the keys
    with pytest.raises(DryadLinqException) as ei:
        ctx.from_enumerable(VALS, 3).range_partition(lambda x: x, range_keys=[5, 9, 2]).partitions()
    assert ei.value.error_code == ErrorCode.PartitionKeysAreNotConsistentlyOrdered
    with pytest.raises(DryadLinqException) as ei:
        ctx.from_enumerable(VALS, 3).range_partition(lambda x: x, 2, range_keys=[1, 2, 3]).partitions()
    assert ei.value.error_code == ErrorCode.RangePartitionInputOutputMismatch
    with pytest.raises(DryadLinqException):
        ctx.from_enumerable(VALS, 3).range_partition(lambda x: x, range_keys=[1000, 2000], descending=True).partitions()


def test_range_partition_custom_comparer_and_single_partition():
    ctx = D.DryadLinqContext(2)
    cmp = lambda a, b: (a % 100 > b % 100) - (a % 100 < b % 100)
    parts = ctx.from_enumerable(VALS, 2).range_partition(lambda x: x, range_keys=[30, 60], comparer=cmp).partitions()
    assert all(x % 100 <= 30 for x in parts[0]) and all(x % 100 > 60 for x in parts[2])
    one = ctx.from_enumerable(VALS, 2).range_partition(lambda x: x, 1).partitions()
    assert len(one) == 1 and sorted(one[0]) == sorted(VALS)


def test_hash_partition_output_count_and_rule():
    ctx = D.DryadLinqContext(3)
    parts = ctx.from_enumerable(VALS, 3).hash_partition(lambda x: x, 5).partitions()
    assert len(parts) == 5
    for j, p in enumerate(parts):
        assert all(hash_bucket(x, 5) == j for x in p)
    assert len(ctx.from_enumerable(VALS, 3).hash_partition(lambda x: x).partitions()) == 3
    with pytest.raises(DryadLinqException):
        ctx.from_enumerable(VALS, 3).hash_partition(lambda x: x, 0)


def test_order_by_skips_shuffle_when_already_range_partitioned():
    ctx = D.DryadLinqContext(3)
    key = lambda x: x
    q = ctx.from_enumerable(VALS, 3).range_partition(key, 3).order_by(key)
    assert q.explain().count("RangePartition(") == 1
    assert q.to_list() == sorted(VALS)
    q2 = ctx.from_enumerable(VALS, 3).order_by(key).order_by(key)          # second sort is a no-op
    assert q2.explain().count("OrderBy(key=") == 1


def test_assume_partition_and_order():
    ctx = D.DryadLinqContext(2)
    data = [[1, 2, 3], [7, 8, 9]]
    q = ctx.from_partitions(data).assume_range_partition(lambda x: x, range_keys=[5]).assume_order_by(lambda x: x)
    q = q.order_by(lambda x: x)
    plan = q.explain()
    assert " RangePartition(" not in plan and "Sampler" not in plan and "OrderBy(key=" not in plan and q.to_list() == [1, 2, 3, 7, 8, 9]
    with pytest.raises(DryadLinqException) as ei:
        ctx.from_partitions(data).assume_range_partition(lambda x: x, range_keys=[1, 2]).to_list()
    assert ei.value.error_code == ErrorCode.BadSeparatorCount


def test_join_reuses_existing_partitioning():
    ctx = D.DryadLinqContext(3)
    a = ctx.from_enumerable([(i, i * i) for i in range(100)], 3).hash_partition(lambda r: r[0], 3)
    b = ctx.from_enumerable([(i % 10, i) for i in range(50)], 2)
    ctx.broadcast_join_rows = 0
    q = a.join(b, lambda r: r[0], lambda r: r[0], lambda x, y: (x[1], y[1]))
    assert q.explain().count("HashPartition(") == 2          # the explicit one + only the inner side
    check_both(lambda c, s: s([(i, i * i) for i in range(100)]).join(s([(i % 10, i) for i in range(50)]),
                                                                    lambda r: r[0], lambda r: r[0], lambda x, y: (x[1], y[1])))


def test_merge_join_on_sorted_inputs_including_descending():
    for desc in (False, True):
        def build(c, s, _d=desc):
            left = s([(i % 20, i) for i in range(60)]).order_by(lambda r: r[0], descending=_d)
            right = s([(i % 15, -i) for i in range(45)])
            return left.join(right, lambda r: r[0], lambda r: r[0], lambda x, y: (x[0], x[1], y[1]))
        check_both(build)
    ctx = D.DryadLinqContext(2)
    ctx.broadcast_join_rows = 0
    keyf = lambda r: r[0]
    left = ctx.from_enumerable([(i % 9, i) for i in range(30)], 2).range_partition(keyf, range_keys=[4]).order_by(keyf)
    q = left.join(ctx.from_enumerable([(i % 9, -i) for i in range(20)], 2), keyf, keyf, lambda x, y: (x[1], y[1]))
    plan = q.explain()
    assert "algorithm=merge" in plan and plan.count(" RangePartition(") == 2     # inner side follows the outer's keys
    assert len(q.to_list()) == sum(1 for i in range(30) for j in range(20) if i % 9 == j % 9)


def test_apply_variants():
    data = list(range(40))
    check_both(lambda c, s: s(data).apply(lambda xs: [sum(xs), len(xs)]), sort=False)
    check_both(lambda c, s: s(data).apply(lambda xs, ys: [sum(xs) + sum(ys)], s([1, 2, 3])), sort=False)
    ctx = D.DryadLinqContext(4)
    per = ctx.from_enumerable(data, 4).apply_per_partition(lambda xs: [len(xs)]).to_list()
    assert per == [10, 10, 10, 10]
    left_h = ctx.from_enumerable(data, 4).apply_per_partition(lambda xs, ys: [(len(xs), sorted(ys))],
                                                              ctx.from_enumerable([5, 6], 2), is_first_only=True).to_list()
    assert left_h == [(10, [5, 6])] * 4
    full = ctx.from_enumerable(data, 4).apply_per_partition(lambda xs, ys: [sum(xs) - sum(ys)], ctx.from_enumerable(data, 4)).to_list()
    assert full == [0, 0, 0, 0]
    with pytest.raises(DryadLinqException) as ei:
        ctx.from_enumerable(data, 4).apply_per_partition(lambda xs, ys: xs, ctx.from_enumerable(data, 3)).to_list()
    assert ei.value.error_code == ErrorCode.HomomorphicApplyNeedsSamePartitionCount
    idx = ctx.from_enumerable(data, 4).apply_with_partition_index(lambda xs, i: [i]).to_list()
    assert idx == [0, 1, 2, 3]
    with pytest.raises(DryadLinqException) as ei:
        ctx.from_enumerable(data, 2).apply(lambda xs: 1 / 0).to_list()
    assert "ZeroDivisionError" in str(ei.value)


def test_fork_and_multi_output():
    ctx = D.DryadLinqContext(3)
    f = ctx.from_enumerable(range(30), 3).fork(lambda x: (x if x % 2 == 0 else None, x * 10 if x % 3 == 0 else None))
    evens, tens = ctx.submit_and_wait(f.first, f.second)
    assert sorted(evens.to_list()) == list(range(0, 30, 2)) and sorted(tens.to_list()) == [x * 10 for x in range(0, 30, 3)]
    f2 = ctx.from_enumerable(range(30), 3).fork(lambda x: (x, None))
    with pytest.raises(DryadLinqException) as ei:
        f2.first.to_list()                                   # every branch of a Fork must be consumed
    assert ei.value.error_code == ErrorCode.BranchOfForkNotUsed
    k = ctx.from_enumerable(range(30), 3).fork_by_key(lambda x: x % 3, [0, 1, 2])
    r0, r1, r2 = ctx.submit_and_wait(k.by_key(0), k.by_key(1), k.by_key(2))
    assert sorted(r1.to_list()) == list(range(1, 30, 3))
    fp = ctx.from_enumerable(range(30), 3).fork_per_partition(lambda xs: ([x for x in xs if x < 5], [len(xs)]))
    a, b = ctx.submit_and_wait(fp[0], fp[1])
    assert sorted(a.to_list()) == [0, 1, 2, 3, 4] and b.to_list() == [10, 10, 10]


def test_shared_subquery_becomes_a_tee_and_runs_once(tmp_path):
    ctx = D.DryadLinqContext(2)
    calls = []
    base = ctx.from_enumerable(range(10), 2).select(lambda x: (calls.append(x), x * 2)[1])
    q1, q2 = base.where(lambda x: x > 5), base.select(lambda x: x + 1)
    r1, r2 = ctx.submit_and_wait(q1, q2)
    assert sorted(r1.to_list()) == [6, 8, 10, 12, 14, 16, 18] and sorted(r2.to_list()) == [x * 2 + 1 for x in range(10)]
    assert len([c for c in calls if isinstance(c, int)]) == 10      # (+1 symbolic call: the tracer)
    assert "Tee" in ctx.last_plan.explain()


def test_sliding_window_broadcast_do_while():
    data = list(range(20))
    check_both(lambda c, s: s(data).sliding_window(lambda w: sum(w), 3), sort=False)
    check_both(lambda c, s: s(data).sliding_window(lambda w: tuple(w), 7), parts=(1, 4, 9), sort=False)
    ctx = D.DryadLinqContext(3)
    b = ctx.from_enumerable([1, 2], 1).broadcast(3).partitions()
    assert b == [[1, 2]] * 3
    res = ctx.from_enumerable([1, 2, 3], 3).do_while(lambda q: q.select(lambda x: x * 2),
                                                     lambda old, new: new.sum_as_query().select(lambda s: s < 100))
    assert sorted(res.to_list()) == [32, 64, 96]


def test_rewrites_push_where_and_drop_reordering():
    ctx = D.DryadLinqContext(3)
    q = ctx.from_enumerable(VALS, 3).order_by(lambda x: x).where(lambda x: x > 100)
    plan = q.explain()
    assert plan.index("Where(") < plan.index("OrderBy(key=")          # filter runs before the sort
    assert q.to_list() == sorted(x for x in VALS if x > 100)
    c = ctx.from_enumerable(VALS, 3).order_by(lambda x: x).count_as_query()
    assert "OrderBy(key=" not in c.explain() and c.to_list() == [len(VALS)]
    u = ctx.from_enumerable(VALS, 2).concat(ctx.from_enumerable(VALS, 2)).where(lambda x: x < 10)
    assert u.explain().count("Where(") == 2 and sorted(u.to_list()) == sorted([x for x in VALS if x < 10] * 2)


def test_misc_regressions():
    # negative keys hash into a valid port; anonymous-type style composite keys group correctly
    check_both(lambda c, s: s(range(-50, 50)).hash_partition(lambda x: x, 4).select(lambda x: x))
    check_both(lambda c, s: s([(i % 3, i % 2, i) for i in range(60)]).group_by(lambda r: (r[0], r[1]))
               .select(lambda g: (g.key, g.count())))
    check_both(lambda c, s: s(["b", "a", "c", "a"]).distinct().order_by(lambda x: x), sort=False)
    check_both(lambda c, s: s([3, 1, 2]).group_by(lambda x: x % 2, comparer=None).select(lambda g: (g.key, g.max())))


def test_user_library_operators_of_the_reference(tmp_path):
    """MapReduce, CrossProduct, CheckOrderBy, BroadCast(destination), DoWhile with an unrolling count
    (LinqToDryad/DryadLinqExtension.cs:152-300)."""
    import dryad_b200 as D
    from dryad_b200.api.errors import DryadLinqException
    for ld in (True, False):
        c = D.DryadLinqContext(3, local_debug=ld, work_dir=str(tmp_path))
        a = c.from_enumerable(list(range(6)), 3)
        b = c.from_enumerable(["x", "y"], 2)
        assert sorted(a.CrossProduct(b, lambda i, s: f"{s}{i}")) == sorted(f"{s}{i}" for i in range(6) for s in "xy")
        mr = c.from_enumerable(["a b", "b c c"], 2).MapReduce(lambda l: l.split(" "), lambda w: w, lambda k, ws: (k, len(list(ws))))
        assert sorted(mr) == [("a", 1), ("b", 2), ("c", 2)]
        assert sorted(c.from_enumerable([3, 1, 2], 1).order_by(lambda x: x).CheckOrderBy(lambda x: x)) == [1, 2, 3]
        assert list(c.from_enumerable([3, 2, 1], 1).check_order_by(lambda x: x, descending=True)) == [3, 2, 1]
        with pytest.raises(DryadLinqException) as e:
            list(c.from_enumerable([3, 1, 2], 1).check_order_by(lambda x: x))
        assert "not ordered" in str(e.value)
        assert sorted(b.broadcast(a)) == sorted(["x", "y"] * (1 if ld else 3))      # one copy per partition of the destination
        r = c.from_enumerable([1], 1).DoWhile(lambda q: q.select(lambda x: x * 2),
                                              lambda o, n: n.select(lambda x: x < 60).any_as_query(lambda t: t), count=2)
        assert list(r) == [64]                                                       # two doublings per job
        assert list(c.from_enumerable([5], 1).do_while(lambda q: q, lambda o, n: n.any_as_query(), count=0)) == [5]
        assert c.from_enumerable(range(10), 2).LongCountAsQuery().to_list() == [10]
        assert c.from_enumerable([1, 2], 1).SequenceEqualAsQuery(c.from_enumerable([1, 2], 1)).to_list() == [True]
        assert sorted(c.from_enumerable(range(4), 2).ApplyWithPartitionIndex(lambda xs, i: [(i, len(list(xs)))]))[0][1] >= 2
