"""k-means as a query loop — the DryadLINQ idiom: centroids are a small dataset, every iteration is one
job (assign points to the nearest centroid, GroupBy the assignment, average per group), `do_while`
repeats it until the centroids stop moving (reference: DryadLinqQueryable.DoWhile,
LinqToDryad/DryadLinqQueryable.cs:1280-1306).

  python samples/kmeans.py [--mode local|debug|cluster] [--points 3000] [--k 4] [--dim 2]

On B200s the same loop runs on the tcgen05 k-means kernel: see `dryad_b200.models.kmeans.KMeansStep`
and `bench.py --workload kmeans` (the distance GEMM + fp32-exact assignment + NVLS centroid all-reduce).
"""
import argparse
import os
import random
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dryad_b200 as D  # noqa: E402
from dryad_b200.api.context import PlatformKind  # noqa: E402


def nearest(p, cents):
    best, bi = None, -1
    for i, c in cents:
        d = sum((a - b) * (a - b) for a, b in zip(p, c))
        if best is None or d < best:
            best, bi = d, i
    return bi


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", choices=["local", "debug", "cluster"], default="local")
    ap.add_argument("--points", type=int, default=3000)
    ap.add_argument("--k", type=int, default=4)
    ap.add_argument("--dim", type=int, default=2)
    ap.add_argument("--partitions", type=int, default=4)
    ap.add_argument("--max-iter", type=int, default=20)
    a = ap.parse_args(argv)
    rng = random.Random(7)
    true = [tuple(rng.uniform(-10, 10) for _ in range(a.dim)) for _ in range(a.k)]
    pts = [tuple(c + rng.gauss(0, 0.5) for c in true[i % a.k]) for i in range(a.points)]
    ctx = D.DryadLinqContext(a.partitions, local_debug=(a.mode == "debug"),
                             platform=PlatformKind.LOCAL_CLUSTER if a.mode == "cluster" else PlatformKind.LOCAL)
    points = ctx.from_enumerable(pts, a.partitions)
    cents0 = ctx.from_enumerable([(i, pts[i * 37 % len(pts)]) for i in range(a.k)], 1)
    dim = a.dim
    iters = [0]

    def step(cents):
        cur = cents.to_list()                         # the small side: shipped with the query, like a broadcast side input
        iters[0] += 1
        return (points.group_by(lambda p: nearest(p, cur))
                      .select(lambda g: (g.key, tuple(sum(p[j] for p in g) / g.count() for j in range(dim)))))

    def moved(old, new):
        o = dict(old.to_list())
        limit = a.max_iter
        n_iter = iters[0]
        return new.select(lambda kc: sum((x - y) ** 2 for x, y in zip(kc[1], o.get(kc[0], kc[1])))) \
                  .sum_as_query().select(lambda s: s > 1e-9 and n_iter < limit)

    final = cents0.do_while(step, moved)
    got = sorted(c for _, c in final.to_list())
    err = max(min(sum((x - y) ** 2 for x, y in zip(c, t)) ** 0.5 for t in true) for c in got)
    print(f"{iters[0]} iterations, {len(got)} centroids, worst distance to a true centre {err:.3f}")
    for c in got:
        print("  " + ", ".join(f"{x:8.3f}" for x in c))
    return 0 if err < 1.0 else 1


if __name__ == "__main__":
    raise SystemExit(main())
