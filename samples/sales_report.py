"""A star-schema report: join a fact table with a small dimension, aggregate per group, order the result.

  python samples/sales_report.py [--mode local|debug|cluster|gpu] [--rows 20000]

  orders.Join(products, o => o.product, p => p.id, (o, p) => new { p.category, revenue = o.qty * p.price })
        .GroupBy(r => r.category).Select(g => new { category = g.Key, revenue = g.Sum(r => r.revenue), n = g.Count() })
        .OrderByDescending(r => r.revenue)

The planner picks a broadcast join for the small dimension (no shuffle of the fact table), decomposes the
aggregation into partial -> shuffle -> final, and range-partitions for the OrderBy; `--mode cluster` runs
it in a graph-manager process with vertex-host processes, `--mode gpu` on one B200 with columnar data.
"""
import argparse
import os
import random
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dryad_b200 as D  # noqa: E402
from dryad_b200.api.context import PlatformKind  # noqa: E402


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", choices=["local", "debug", "cluster", "gpu"], default="local")
    ap.add_argument("--rows", type=int, default=20000)
    ap.add_argument("--partitions", type=int, default=4)
    a = ap.parse_args(argv)
    rng = random.Random(11)
    n_products, n_cat = 200, 8
    product = [rng.randrange(n_products) for _ in range(a.rows)]
    qty = [rng.randrange(1, 6) for _ in range(a.rows)]
    pid = list(range(n_products))
    category = [i % n_cat for i in pid]
    price = [5 + (i * 7) % 90 for i in pid]
    ctx = D.DryadLinqContext(a.partitions, local_debug=(a.mode == "debug"), device="cuda:0" if a.mode == "gpu" else "cpu",
                             platform=PlatformKind.LOCAL_CLUSTER if a.mode == "cluster" else PlatformKind.LOCAL)
    orders = ctx.from_columns({"product": product, "qty": qty}, a.partitions)
    products = ctx.from_columns({"id": pid, "category": category, "price": price}, 1)
    report = (orders.join(products, lambda o: o.product, lambda p: p.id,
                          lambda o, p: {"category": p.category, "revenue": o.qty * p.price})
                    .group_by(lambda r: r.category)
                    .select(lambda g: {"category": g.key, "revenue": g.sum(lambda r: r.revenue), "n": g.count()})
                    .order_by_descending(lambda r: r.revenue))
    rows = report.to_list()
    want = {}
    for p_, q_ in zip(product, qty):
        c = category[p_]
        r, n = want.get(c, (0, 0))
        want[c] = (r + q_ * price[p_], n + 1)
    ok = sorted((r.category, r.revenue, r.n) for r in rows) == sorted((c, r, n) for c, (r, n) in want.items())
    ok = ok and [r.revenue for r in rows] == sorted((r.revenue for r in rows), reverse=True)
    print(f"{'category':>8} {'revenue':>12} {'orders':>8}")
    for r in rows:
        print(f"{r.category:>8} {r.revenue:>12} {r.n:>8}")
    print("plan:\n" + ctx.last_plan.explain() if getattr(ctx, "last_plan", None) is not None else "")
    return 0 if ok else 1


if __name__ == "__main__":
    raise SystemExit(main())
