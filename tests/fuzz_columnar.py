"""Random TYPED queries (records of int/float columns, traced lambdas only): on the GPU backend every
stage of these runs as kernels; on any backend the result must equal the LocalDebug oracle's."""
import random


def build_columnar(rng: random.Random, c):
    """-> (description, query | ("scalar", thunk), ordered)."""
    n = rng.choice([1, 2, 3])
    m = rng.choice([0, 1, 5, 300, 2500])
    kmod = rng.choice([1, 3, 17, 400])
    neg = rng.choice([0, 50])
    recs = [{"k": rng.randrange(kmod) - (neg if rng.random() < 0.3 else 0), "v": rng.randrange(-1000, 1000),
             "f": rng.randrange(-50, 50) * 0.25} for _ in range(m)]
    q = c.from_enumerable(recs, n)
    desc = [f"recs(n={n},m={m},kmod={kmod})"]
    ordered = True
    for _ in range(rng.randrange(1, 4)):
        op = rng.choice(["where", "select", "selectf", "order", "orderd", "orderf", "distinctk", "take", "skip", "reverse", "hashp",
                         "rangep", "indexed", "whereidx", "zip", "window", "concat", "takewhile", "skipwhile"])
        if not ordered and op in ("take", "skip", "reverse", "indexed", "whereidx", "zip", "window", "takewhile", "skipwhile"):
            break
        desc.append(op)
        if op == "where":
            q = q.where(lambda r: (r.v + r.k) % 3 != 0)
        elif op == "select":
            q = q.select(lambda r: {"k": r.k, "v": r.v * 2 - r.k, "f": r.f})
        elif op == "selectf":
            q = q.select(lambda r: {"k": r.k, "v": r.v, "f": r.f * 0.5 + r.v})
        elif op == "order":
            q = q.order_by(lambda r: r.v * 4096 + r.k)      # near-unique key: a total order both sides agree on
            ordered = "keyed"
        elif op == "orderd":
            q = q.order_by_descending(lambda r: r.v * 4096 + r.k)
            ordered = "keyed"
        elif op == "orderf":
            q = q.order_by(lambda r: r.f)
            ordered = False                                   # many ties: what follows must not depend on their order
        elif op == "distinctk":
            q = q.select(lambda r: {"k": r.k, "v": r.v % 5, "f": 0.0}).distinct()
            ordered = False
        elif op == "take":
            q = q.take(rng.choice([0, 1, 7, 1000]))
        elif op == "skip":
            q = q.skip(rng.choice([0, 1, 7, 1000]))
        elif op == "reverse":
            q = q.reverse()
        elif op == "hashp":
            q = q.hash_partition(lambda r: r.k, rng.choice([1, 2, 4]))
            ordered = False
        elif op == "rangep":
            q = q.range_partition(lambda r: r.v, rng.choice([1, 2, 4]))
            ordered = False
        elif op == "indexed":
            q = q.select_indexed(lambda r, i: {"k": r.k, "v": r.v + i, "f": r.f})
        elif op == "whereidx":
            q = q.where_indexed(lambda r, i: (r.v + i) % 2 == 0)
        elif op == "zip":
            q = q.zip(q.select(lambda r: r.v + 1), lambda r, w: {"k": r.k, "v": r.v + w, "f": r.f})
        elif op == "window":
            q = q.select(lambda r: r.v).sliding_window(lambda w: w[0] + w[-1], rng.choice([2, 3]))
            desc.append("->ints")
            return _finish_ints(rng, desc, q, ordered)
        elif op == "concat":
            q = q.concat(c.from_enumerable([{"k": 1, "v": 5, "f": 0.5}, {"k": 2, "v": -5, "f": 1.5}], 1))
        elif op == "takewhile":
            q = q.take_while(lambda r: r.v < 900)
        elif op == "skipwhile":
            q = q.skip_while(lambda r: r.v < 0)
    fin = rng.choice(["rows", "group", "groupf", "join", "gjoin", "sum", "minmax", "avg", "count", "anyall", "firstlast", "setops", "rows"])
    desc.append(fin)
    if fin == "rows":
        return desc, q, ordered
    if fin == "group":
        return desc, q.group_by(lambda r: r.k).select(lambda g: {"k": g.key, "c": g.count(), "s": g.sum(lambda r: r.v),
                                                                 "mn": g.min(lambda r: r.v), "mx": g.max(lambda r: r.v)}), False
    if fin == "groupf":
        return desc, q.group_by(lambda r: r.k % 4).select(lambda g: {"k": g.key, "a": g.average(lambda r: r.f), "sf": g.sum(lambda r: r.f)}), False
    if fin == "join":
        dim = c.from_enumerable([{"id": i - 20, "w": i * 3} for i in range(0, 80, rng.choice([1, 2]))], rng.choice([1, 2]))
        return desc, q.join(dim, lambda r: r.k, lambda d: d.id, lambda r, d: {"k": r.k, "x": r.v + d.w}), False
    if fin == "gjoin":
        outer = c.from_enumerable([{"id": i - 5, "w": i} for i in range(30)], rng.choice([1, 2]))
        return desc, outer.group_join(q, lambda o: o.id, lambda r: r.k,
                                      lambda o, g: {"id": o.id, "n": g.count(), "s": g.sum(lambda r: r.v) + o.w}), False
    if fin == "sum":
        return desc, ("scalar", lambda: q.sum(lambda r: r.v)), True
    if fin == "minmax":
        return desc, ("scalar", lambda: (q.min(lambda r: r.v), q.max(lambda r: r.f))), True
    if fin == "avg":
        return desc, ("scalar", lambda: round(q.average(lambda r: r.v), 9)), True
    if fin == "count":
        return desc, ("scalar", lambda: (q.count(), q.count(lambda r: r.v > 0))), True
    if fin == "anyall":
        return desc, ("scalar", lambda: (q.any(lambda r: r.v == 7), q.all(lambda r: r.v > -1000))), True
    if fin == "firstlast":
        if not ordered:
            return desc, ("scalar", lambda: q.count()), True
        return desc, ("scalar", lambda: (q.select(lambda r: r.v).first_or_default(), q.select(lambda r: r.v).last_or_default())), True
    if fin == "setops":
        a = q.select(lambda r: r.k)
        b = c.from_enumerable(list(range(-3, 9)), 2)
        return desc, rng.choice([a.union(b), a.intersect(b), a.except_(b)]), False
    return desc, q, ordered


def _finish_ints(rng, desc, q, ordered):
    fin = rng.choice(["rows", "sum", "distinct"])
    desc.append(fin)
    if fin == "sum":
        return desc, ("scalar", lambda: q.sum()), True
    if fin == "distinct":
        return desc, q.distinct(), False
    return desc, q, ordered
