"""Decomposable GroupBy (mirrors DryadLinqTests/GroupByReduceTests.cs:71-1201)."""
import pytest

import dryad_b200 as D
from dryad_b200 import IDecomposable, decomposable
from util import check_both

DATA = [(i % 7, i) for i in range(200)]          # 2 x 100: enough per partition to trigger the combiner


def test_average_decomposition():
    check_both(lambda c, s: s(DATA).group_by(lambda r: r[0]).select(lambda g: (g.key, g.average(lambda r: r[1]))))


def test_builtin_count_sum_min_max_any_all_first():
    check_both(lambda c, s: s(DATA).group_by(lambda r: r[0]).select(
        lambda g: (g.key, g.count(), g.sum(lambda r: r[1]), g.min(lambda r: r[1]), g.max(lambda r: r[1]),
                   g.any(lambda r: r[1] > 190), g.all(lambda r: r[1] >= 0))))


def test_distributive_result_selector_form():
    check_both(lambda c, s: s(DATA).group_by(lambda r: r[0], result=lambda k, g: (k, g.sum(lambda r: r[1]) * 2 + 1)))


def test_same_aggregate_twice_and_complex_expression():
    def build(c, s):
        return s(DATA).group_by(lambda r: r[0]).select(
            lambda g: {"k": g.key, "a": g.sum(lambda r: r[1]) + g.sum(lambda r: r[1]), "b": g.count() * 3 - g.key})
    want = check_both(build)
    assert all(r["a"] % 2 == 0 for r in want)
    # the identical Sum is planned once
    ctx = D.DryadLinqContext(2)
    q = build(ctx, lambda d, n=2: ctx.from_enumerable(d, n))
    assert len(q.expr.params["decomp"].aggs) == 2


class SumOfSquares(IDecomposable):
    def seed(self, x): return x * x
    def accumulate(self, acc, x): return acc + x * x
    def recursive_accumulate(self, a, b): return a + b
    def final_reduce(self, acc): return ("ss", acc)


class NoFinal(IDecomposable):
    def seed(self, x): return [x]
    def accumulate(self, acc, x): return acc + [x] if len(acc) < 3 else acc
    def recursive_accumulate(self, a, b): return (a + b)[:3]


def test_custom_decomposable_with_and_without_finalizer():
    check_both(lambda c, s: s(DATA).group_by(lambda r: r[0]).select(
        lambda g: (g.key, g.aggregate(SumOfSquares, lambda r: r[1]))))
    got = check_both(lambda c, s: s(DATA).group_by(lambda r: r[0]).select(
        lambda g: (g.key, g.count(), g.aggregate(SumOfSquares, lambda r: r[1]))), parts=(1, 4))
    assert all(r[2][0] == "ss" for r in got)
    ctx = D.DryadLinqContext(2)
    r = ctx.from_enumerable(DATA, 2).group_by(lambda r: r[0]).select(lambda g: (g.key, g.aggregate(NoFinal, lambda r: r[1]))).to_list()
    assert all(len(v) == 3 for _, v in r)


@decomposable(SumOfSquares)
def sum_squares_of_group(group):
    return ("ss", sum(x * x for x in group))


def test_decomposable_attribute_on_a_function():
    check_both(lambda c, s: s([x for _, x in DATA]).group_by(lambda x: x % 5).select(sum_squares_of_group))


def test_non_decomposable_selector_still_works():
    check_both(lambda c, s: s(DATA).group_by(lambda r: r[0]).select(lambda g: (g.key, sorted(x[1] for x in g)[:3])))
    check_both(lambda c, s: s(DATA).group_by(lambda r: r[0], elem=lambda r: r[1]).select(lambda g: (g.key, len(g))))


def test_group_by_on_already_partitioned_input_skips_the_shuffle():
    ctx = D.DryadLinqContext(3)
    q = ctx.from_enumerable([x for _, x in DATA], 3).hash_partition(lambda x: x % 5, 3) \
        .group_by(lambda x: x % 5).select(lambda g: (g.key, g.count()))
    plan = q.explain()
    assert plan.count("HashPartition(") == 1        # only the explicit one
    assert "mode=full_acc" in plan
    assert sorted(q.to_list()) == [(k, 40) for k in range(5)]


def test_checked_sum_overflow_raises():
    big = [2**62, 2**62, 2**62]
    with pytest.raises(Exception) as ei:
        D.DryadLinqContext(1).from_enumerable(big).group_by(lambda x: 0).select(lambda g: g.sum()).to_list()
    assert "overflow" in str(ei.value).lower()


def test_pair_results_keep_their_type_through_decomposition_and_sorting():
    """`new Pair(g.Key, g.Count())` stays a Pair (`.key` / `.value`) after the decomposed GroupBy,
    so later operators can use its members (the WordCount sample's OrderByDescending(p => p.Value))."""
    import dryad_b200 as D
    words = ("a b c a b a d d d d".split()) * 3
    for ctx in (D.DryadLinqContext(1, local_debug=True), D.DryadLinqContext(3)):
        q = (ctx.from_enumerable(words, 3 if not ctx.local_debug else 1).group_by(lambda w: w)
             .select(lambda g: D.Pair(g.key, g.count())).order_by_descending(lambda p: p.value).take(2))
        got = q.to_list()
        assert got == [D.Pair("d", 12), D.Pair("a", 9)] and all(isinstance(p, D.Pair) for p in got)
    nums = list(range(100))
    r = D.DryadLinqContext(2).from_enumerable(nums, 2).select(lambda x: D.Pair(x % 5, x * 2)).where(lambda p: p.key == 3).to_list()
    assert r and all(isinstance(p, D.Pair) and p.key == 3 for p in r)
