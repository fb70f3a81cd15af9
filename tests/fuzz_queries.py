"""Random query generators for the differential tests: the same seed builds the same query on two
contexts, one of which is the LocalDebug oracle (the reference's tests validate every cluster result
against LocalDebug the same way, DryadLinqTests/Utils.cs Validate.Check)."""
import random


def rand_pipeline(rng, c, depth=0):
    n = rng.choice([1, 2, 3, 5])
    data = [rng.randrange(0, 40) for _ in range(rng.choice([0, 1, 7, 60, 200]))]
    q = c.from_enumerable(data, n)
    desc = [f"src(n={n},len={len(data)})"]
    ordered = True
    steps = rng.randrange(1, 5)
    for _ in range(steps):
        op = rng.choice(["select", "where", "distinct", "order", "orderd", "take", "skip", "reverse", "hashp", "rangep", "concat", "union",
                         "intersect", "except", "zip", "window", "groupcount", "selmany", "indexed", "join", "apply", "takewhile", "skipwhile"])
        desc.append(op)
        if not ordered and op in ("take", "skip", "takewhile", "skipwhile", "zip", "window", "indexed", "apply", "concat", "reverse", "selmany", "select", "where"):
            if op in ("take", "skip", "takewhile", "skipwhile", "zip", "window", "indexed"):
                return None, desc, ordered
        if op == "select": q = q.select(lambda x: x * 3 + 1)
        elif op == "where": q = q.where(lambda x: x % 3 != 0)
        elif op == "distinct": q = q.distinct(); ordered = False
        elif op == "order": q = q.order_by(lambda x: x); ordered = True
        elif op == "orderd": q = q.order_by_descending(lambda x: x); ordered = True
        elif op == "take": q = q.take(rng.choice([0, 1, 5, 50]))
        elif op == "skip": q = q.skip(rng.choice([0, 1, 5, 50]))
        elif op == "reverse": q = q.reverse()
        elif op == "hashp": q = q.hash_partition(lambda x: x, rng.choice([1, 2, 4])); ordered = False
        elif op == "rangep": q = q.range_partition(lambda x: x, rng.choice([1, 2, 4])); ordered = False
        elif op == "concat": q = q.concat(c.from_enumerable([rng.randrange(0, 40) for _ in range(rng.choice([0, 3, 30]))], rng.choice([1, 2, 3])))
        elif op == "union": q = q.union(c.from_enumerable([rng.randrange(0, 40) for _ in range(20)], rng.choice([1, 3]))); ordered = False
        elif op == "intersect": q = q.intersect(c.from_enumerable([rng.randrange(0, 40) for _ in range(20)], rng.choice([1, 3]))); ordered = False
        elif op == "except": q = q.except_(c.from_enumerable([rng.randrange(0, 40) for _ in range(20)], rng.choice([1, 3]))); ordered = False
        elif op == "zip": q = q.zip(c.from_enumerable(list(range(100)), rng.choice([1, 2, 4])), lambda a, b: a * 100 + b)
        elif op == "window": q = q.sliding_window(lambda w: sum(w), rng.choice([1, 2, 4]))
        elif op == "groupcount": q = q.group_by(lambda x: x % 7).select(lambda g: g.key * 1000 + g.count()); ordered = False
        elif op == "selmany": q = q.select_many(lambda x: [x, x + 1] if x % 2 else [])
        elif op == "indexed": q = q.select_indexed(lambda x, i: x * 1000 + i)
        elif op == "join": q = q.join(c.from_enumerable(list(range(0, 40, 3)), rng.choice([1, 2])), lambda x: x, lambda y: y, lambda x, y: x + y); ordered = False
        elif op == "apply": q = q.apply(lambda xs: [x for x in xs if x % 2 == 0])
        elif op == "takewhile": q = q.take_while(lambda x: x < 30)
        elif op == "skipwhile": q = q.skip_while(lambda x: x < 10)
    return q, desc, ordered



WORDS = ["ant", "bee", "cat", "dog", "eel", "fox", "gnu", "hen"]

def build(rng, c):
    n = rng.choice([1, 2, 4])
    m = rng.choice([0, 1, 9, 120])
    recs = [(rng.choice(WORDS), rng.randrange(0, 20), rng.random()) for _ in range(m)]
    q = c.from_enumerable(recs, n)
    kind = rng.choice(["gb_multi", "gb_str", "gb_elem", "order_then", "groupjoin", "agg_sum", "agg_minmax", "agg_avg", "agg_fn", "fork", "distinct_rec",
                       "join_rec", "count_pred", "first_last", "setops", "lookup", "gb_result", "anyall"])
    if kind == "gb_multi":
        return kind, q.group_by(lambda r: r[1] % 5).select(lambda g: (g.key, g.count(), g.sum(lambda r: r[1]), g.min(lambda r: r[1]), g.max(lambda r: r[1]))), False
    if kind == "gb_str":
        return kind, q.group_by(lambda r: r[0]).select(lambda g: (g.key, g.count())), False
    if kind == "gb_elem":
        return kind, q.group_by(lambda r: r[0], lambda r: r[1]).select(lambda g: (g.key, sorted(g))), False
    if kind == "gb_result":
        return kind, q.group_by(lambda r: r[1] % 3, lambda r: r[1], lambda k, g: (k, sum(g))), False
    if kind == "order_then":
        return kind, q.order_by(lambda r: r[0]).then_by_descending(lambda r: r[1]).then_by(lambda r: r[2]), True
    if kind == "groupjoin":
        other = c.from_enumerable([(w, len(w) + i) for i, w in enumerate(WORDS[:5])], rng.choice([1, 2]))
        return kind, other.group_join(q, lambda o: o[0], lambda r: r[0], lambda o, rs: (o[0], len(list(rs)))), False
    if kind == "agg_sum":
        return kind, ("scalar", lambda: q.sum(lambda r: r[1])), True
    if kind == "agg_minmax":
        return kind, ("scalar", lambda: (q.min(lambda r: r[1]), q.max(lambda r: r[1]))), True
    if kind == "agg_avg":
        return kind, ("scalar", lambda: round(q.average(lambda r: r[1]), 9)), True
    if kind == "agg_fn":
        import operator
        return kind, ("scalar", lambda: q.select(lambda r: r[1]).aggregate(operator.add)), True
    if kind == "fork":
        evens, odds = q.select(lambda r: r[1]).fork(lambda x: x % 2 == 0, lambda x: x % 2 == 1) if hasattr(q, "fork") else (None, None)
        return kind, evens.concat(odds), False
    if kind == "distinct_rec":
        return kind, q.select(lambda r: (r[0], r[1] % 3)).distinct(), False
    if kind == "join_rec":
        other = c.from_enumerable([(w, i) for i, w in enumerate(WORDS)], rng.choice([1, 3]))
        return kind, q.join(other, lambda r: r[0], lambda o: o[0], lambda r, o: (r[0], r[1], o[1])), False
    if kind == "count_pred":
        return kind, ("scalar", lambda: (q.count(lambda r: r[1] > 10), q.long_count())), True
    if kind == "first_last":
        return kind, ("scalar", lambda: (q.first_or_default(), q.last_or_default())), True
    if kind == "setops":
        a = q.select(lambda r: r[1]); b = c.from_enumerable(list(range(5, 15)), 2)
        return kind, a.union(b).except_(c.from_enumerable([7, 8], 1)), False
    if kind == "lookup":
        return kind, q.select(lambda r: r[1]).where(lambda x: x > 3).select_many(lambda x: [x] * (x % 3)), True
    if kind == "anyall":
        return kind, ("scalar", lambda: (q.any(), q.any(lambda r: r[1] == 19), q.all(lambda r: r[1] < 19))), True

