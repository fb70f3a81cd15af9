"""Random queries over the less common operators (Fork, the Apply family, Assume*, SequenceEqual, Contains,
Single, indexed Take/SkipWhile, seeded Aggregate, chained Zip, wide SlidingWindow, explicit range keys, empty
inputs): every result is phrased so that it does not depend on how the data happens to be partitioned."""
import random  # noqa: F401


def build_operators(rng, c):
    n = rng.choice([1, 2, 3, 4])
    data = [rng.randrange(-20, 60) for _ in range(rng.choice([0, 1, 2, 9, 80, 400]))]
    q = c.from_enumerable(data, n)
    kind = rng.choice(["fork", "fork_key", "app_pp", "app_bin_first", "app_bin_full", "app_idx", "assume_hash", "assume_range", "assume_order",
                       "seq_equal", "contains", "single", "elem_defaults", "tw_idx", "sw_idx", "order_take", "order_skip_rev", "distinct_order",
                       "gb_nested", "join_self", "union_concat", "selmany_idx", "hash_counts", "range_keys", "agg_seed", "zip3", "window_big",
                       "group_comparer", "take_zero", "concat_empty"])
    if kind == "fork":
        a, b = q.fork(lambda x: (x if x % 2 == 0 else None, x * 10 if x % 3 == 0 else None))
        return kind, a.concat(b), False
    if kind == "fork_key":
        m = q.fork_by_key(lambda x: x % 3, [0, 1, 2])
        return kind, m[0].select(lambda x: ("z", x)).concat(m[2].select(lambda x: ("t", x))).concat(m[1].select(lambda x: ("o", x))), False
    if kind == "app_pp":
        return kind, q.apply_per_partition(lambda xs: [sum(xs)]).sum_as_query(), True
    if kind == "app_bin_first":
        o = c.from_enumerable([1, 2, 3], 2)
        return kind, q.apply_per_partition(lambda xs, ys: [x + sum(ys) for x in xs], o, True), False
    if kind == "app_bin_full":
        o = c.from_enumerable(list(range(len(data))), n)
        return kind, q.apply_per_partition(lambda xs, ys: [len(list(xs)) - len(list(ys))], o, False).sum_as_query(), True
    if kind == "app_idx":
        return kind, q.apply_with_partition_index(lambda xs, i: [i * 0 + len(list(xs))]).sum_as_query(), True
    if kind == "assume_hash":
        return kind, q.hash_partition(lambda x: x, n).assume_hash_partition(lambda x: x).group_by(lambda x: x).select(lambda g: (g.key, g.count())), False
    if kind == "assume_range":
        return kind, q.order_by(lambda x: x).assume_range_partition(lambda x: x, descending=False).select(lambda x: x + 1).order_by(lambda x: x), True
    if kind == "assume_order":
        return kind, q.order_by(lambda x: x).assume_order_by(lambda x: x, descending=False).order_by(lambda x: x).take(5), True
    if kind == "seq_equal":
        return kind, ("scalar", lambda: (q.sequence_equal(c.from_enumerable(list(data), max(1, n - 1))), q.sequence_equal(c.from_enumerable(data[:-1], 1)))), True
    if kind == "contains":
        return kind, ("scalar", lambda: (q.contains(data[0] if data else 5), q.contains(1000))), True
    if kind == "single":
        return kind, ("scalar", lambda: q.where(lambda x: x == 59).single_or_default()), True
    if kind == "elem_defaults":
        return kind, ("scalar", lambda: (q.first_or_default(lambda x: x > 50), q.last_or_default(lambda x: x < -10), q.count(lambda x: x > 0))), True
    if kind == "tw_idx":
        return kind, q.take_while_indexed(lambda x, i: i < 7 or x > 0), True
    if kind == "sw_idx":
        return kind, q.skip_while_indexed(lambda x, i: i < 3 and x < 100), True
    if kind == "order_take":
        return kind, q.select(lambda x: (x, 1)).distinct().order_by(lambda t: t[0]).take(4), True
    if kind == "order_skip_rev":
        return kind, q.distinct().order_by_descending(lambda x: x).skip(2).reverse(), True
    if kind == "distinct_order":
        return kind, q.select(lambda x: x % 5).distinct().order_by(lambda x: x), True
    if kind == "gb_nested":
        return kind, q.group_by(lambda x: x % 4).select(lambda g: (g.key, sorted(g)[:3], g.count())).where(lambda t: t[2] > 0), False
    if kind == "join_self":
        return kind, q.join(q.select(lambda x: x + 1), lambda a: a, lambda b: b, lambda a, b: (a, b)).distinct(), False
    if kind == "union_concat":
        return kind, q.union(q.select(lambda x: -x)).concat(c.from_enumerable([1000], 1)), False
    if kind == "selmany_idx":
        return kind, q.select_many_indexed(lambda x, i: [i] * (x % 3)), True
    if kind == "hash_counts":
        return kind, ("scalar", lambda: sorted(x for p in q.hash_partition(lambda x: x % 5, rng.choice([1, 3, 5])).partitions() for x in p)), True
    if kind == "range_keys":
        return kind, ("scalar", lambda: [x for p in q.range_partition(lambda x: x, range_keys=[0, 20, 40]).partitions() for x in sorted(p)]), True
    if kind == "agg_seed":
        return kind, ("scalar", lambda: q.aggregate(lambda a, x: a + x * 2, 7)), True
    if kind == "zip3":
        return kind, q.zip(q.select(lambda x: x * 2), lambda a, b: b - a).zip(c.from_enumerable(list(range(1000)), 3), lambda a, i: a * 1000 + i), True
    if kind == "window_big":
        return kind, q.sliding_window(lambda w: w[0] - w[-1], 50), True
    if kind == "group_comparer":
        return kind, q.group_by(lambda x: x % 6, None, lambda k, g: (k, max(g), min(g))), False
    if kind == "take_zero":
        return kind, q.take(0).concat(q.skip(10 ** 6)), True
    if kind == "concat_empty":
        return kind, c.from_enumerable([], 2).concat(q).concat(c.from_enumerable([], 1)), True

