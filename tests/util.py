"""Differential-test harness: run the same query in cluster mode (partitioned host execution)
and in LocalDebug (LINQ-to-Objects oracle) and compare after sorting — the reference's
Validate.Check pattern (DryadLinqTests/Utils.cs:301-360)."""
import dryad_b200 as D


def _key(x):
    return repr(x)


def check_both(build, parts=(1, 3, 4), sort=True, data=None):
    """build(ctx, source_fn) -> Query.  source_fn(data, n) makes a source with n partitions."""
    ldc = D.DryadLinqContext(1, local_debug=True)
    want = build(ldc, lambda d, n=1: ldc.from_enumerable(d, 1)).to_list()
    for p in parts:
        ctx = D.DryadLinqContext(p)
        got = build(ctx, lambda d, n=None, _p=p: ctx.from_enumerable(d, n or _p)).to_list()
        if sort:
            assert sorted(got, key=_key) == sorted(want, key=_key), f"partitions={p}"
        else:
            assert list(got) == list(want), f"partitions={p}"
    return want
