"""Differential-test harness: run the same query in cluster mode (partitioned host execution)
and in LocalDebug (LINQ-to-Objects oracle) and compare after sorting — the reference's
Validate.Check pattern (DryadLinqTests/Utils.cs:301-360).

`DRYAD_TEST_PLATFORM=cluster python -m pytest tests/test_api_*.py` runs the same suites through a
graph-manager process and vertex-host processes, the way the reference's own tests run
(`new DryadLinqContext(numProcesses)` -> LocalJobSubmission, LocalJobSubmission.cs:97-147)."""
import dryad_b200 as D


def _key(x):
    return repr(x)


def check_both(build, parts=(1, 3, 4), sort=True, data=None):
    """build(ctx, source_fn) -> Query.  source_fn(data, n) makes a source with n partitions."""
    ldc = D.DryadLinqContext(1, local_debug=True)
    want = build(ldc, lambda d, n=1: ldc.from_enumerable(d, 1)).to_list()
    import os
    import tempfile
    cluster = os.environ.get("DRYAD_TEST_PLATFORM", "") == "cluster"     # the same suites through a graph-manager process
    if cluster:                                                          # and vertex hosts (the reference's local test mode)
        parts = tuple(parts)[-1:]
    for p in parts:
        if cluster:
            ctx = D.DryadLinqContext(p, platform=D.PlatformKind.LOCAL_CLUSTER, work_dir=tempfile.mkdtemp(prefix="dryad_t_"))
            ctx.job_max_nodes = 2
        else:
            ctx = D.DryadLinqContext(p)
        got = build(ctx, lambda d, n=None, _p=p: ctx.from_enumerable(d, n or _p)).to_list()
        if sort:
            assert sorted(got, key=_key) == sorted(want, key=_key), f"partitions={p}"
        else:
            assert list(got) == list(want), f"partitions={p}"
    return want
