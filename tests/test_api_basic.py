"""Basic API behaviour (mirrors DryadLinqTests/BasicAPITests.cs and SimpleTests.cs)."""
import os

import pytest

import dryad_b200 as D
from dryad_b200 import DryadLinqException, ErrorCode
from util import check_both

DATA = list(range(1, 13))


def test_where_select_pipeline():
    check_both(lambda c, src: src(DATA).where(lambda x: x % 2 == 0).select(lambda x: x * 10))


def test_select_many_and_indexed_ops():
    check_both(lambda c, src: src(DATA).select_many(lambda x: [x] * (x % 3)))
    check_both(lambda c, src: src(DATA).select_indexed(lambda x, i: (x, i)), sort=False)
    check_both(lambda c, src: src(DATA).where_indexed(lambda x, i: i % 2 == 0), sort=False)
    check_both(lambda c, src: src(DATA).take_while_indexed(lambda x, i: i < 7), sort=False)
    check_both(lambda c, src: src(DATA).skip_while_indexed(lambda x, i: i < 7), sort=False)


def test_order_take_skip_reverse_zip():
    check_both(lambda c, src: src([5, 3, 9, 1, 7, 2, 8]).order_by(lambda x: x), sort=False)
    check_both(lambda c, src: src([5, 3, 9, 1, 7, 2, 8]).order_by_descending(lambda x: x), sort=False)
    check_both(lambda c, src: src(DATA).order_by(lambda x: -x).take(4), sort=False)
    check_both(lambda c, src: src(DATA).skip(5), sort=False)
    check_both(lambda c, src: src(DATA).take(5), sort=False)
    check_both(lambda c, src: src(DATA).reverse(), sort=False)
    check_both(lambda c, src: src(DATA).zip(src(DATA).select(lambda x: x * x), lambda a, b: (a, b)), sort=False)
    # the two sides of a Zip may be partitioned differently: the right side is re-aligned to the left's boundaries
    check_both(lambda c, src: c.from_enumerable(DATA, 4).zip(c.from_enumerable([x * x for x in DATA], 3), lambda a, b: (a, b)), sort=False)
    check_both(lambda c, src: src(DATA).take_while(lambda x: x < 6), sort=False)
    check_both(lambda c, src: src(DATA).skip_while(lambda x: x < 6), sort=False)


def test_set_operators_and_concat():
    a, b = [1, 2, 2, 3, 4, 5, 5], [4, 5, 6, 6, 7]
    check_both(lambda c, src: src(a).distinct())
    check_both(lambda c, src: src(a).union(src(b)))
    check_both(lambda c, src: src(a).intersect(src(b)))
    check_both(lambda c, src: src(a).except_(src(b)))
    check_both(lambda c, src: src(a).concat(src(b)))


def test_joins():
    left = [(i, f"L{i}") for i in range(10)]
    right = [(i % 5, f"R{i}") for i in range(12)]
    check_both(lambda c, src: src(left).join(src(right), lambda l: l[0], lambda r: r[0], lambda l, r: (l[1], r[1])))
    check_both(lambda c, src: src(left).group_join(src(right), lambda l: l[0], lambda r: r[0],
                                                   lambda l, rs: (l[1], sorted(x[1] for x in rs))))


def test_scalar_aggregates():
    for p in (1, 3):
        c = D.DryadLinqContext(p)
        q = c.from_enumerable(DATA, p)
        assert q.count() == 12 and q.sum() == 78 and q.min() == 1 and q.max() == 12
        assert q.average() == 6.5
        assert q.any(lambda x: x > 11) and not q.all(lambda x: x > 1)
        assert q.contains(7) and not q.contains(70)
        assert q.first() == 1 and q.last() == 12
        assert q.first_or_default(lambda x: x > 100) is None
        assert q.single(lambda x: x == 5) == 5
        assert q.aggregate(lambda a, b: a + b) == 78
        assert q.aggregate(lambda a, b: a + b, seed=100) == 178
        assert q.sequence_equal(c.from_enumerable(DATA, 2))
        with pytest.raises(DryadLinqException) as ei:
            q.where(lambda x: x > 100).min()
        assert ei.value.error_code == ErrorCode.MinNoElements
        with pytest.raises(DryadLinqException):
            q.single()


def test_unsupported_operators_raise():
    c = D.DryadLinqContext(2)
    q = c.from_enumerable(DATA, 2)
    for fn in (lambda: q.order_by(lambda x: x).then_by(lambda x: x), lambda: q.element_at(3),
               lambda: q.default_if_empty(), lambda: q.of_type(int)):
        with pytest.raises(DryadLinqException):
            fn()


def test_to_store_submit_and_reload(tmp_path):
    c = D.DryadLinqContext(3)
    uri = f"partfile://{tmp_path}/out.pt"
    q = c.from_enumerable(DATA, 3).select(lambda x: (x, str(x))).to_store(uri)
    job = q.submit()
    res = job.wait()[0]
    assert sorted(res.to_list()) == sorted((x, str(x)) for x in DATA)
    assert os.path.exists(f"{tmp_path}/out.pt")
    back = c.from_store(uri)
    assert back.expr.params["partition_count"] == 3
    assert sorted(back.to_list()) == sorted((x, str(x)) for x in DATA)
    # the output exists now: writing again without delete_if_exists is an error
    with pytest.raises(DryadLinqException) as ei:
        c.from_enumerable(DATA).to_store(uri)
    assert ei.value.error_code == ErrorCode.StreamAlreadyExists
    # gzip output round-trips
    uri2 = f"partfile://{tmp_path}/out_gz.pt"
    c.from_enumerable(DATA, 2).to_store(uri2, compression="Gzip").submit_and_wait()
    assert sorted(c.from_store(uri2).to_list()) == DATA


def test_multi_query_submit_is_one_job(tmp_path):
    c = D.DryadLinqContext(2)
    src = c.from_enumerable(DATA, 2)
    q1 = src.where(lambda x: x > 6).to_store(f"partfile://{tmp_path}/a.pt")
    q2 = src.where(lambda x: x <= 6).to_store(f"partfile://{tmp_path}/b.pt")
    r1, r2 = c.submit_and_wait(q1, q2)
    assert sorted(r1.to_list()) == [7, 8, 9, 10, 11, 12] and sorted(r2.to_list()) == [1, 2, 3, 4, 5, 6]
    with pytest.raises(DryadLinqException):
        c.submit(q1, q1)
    other = D.DryadLinqContext(2)
    with pytest.raises(DryadLinqException):
        c.submit(other.from_enumerable(DATA).select(lambda x: x))


def test_output_rules(tmp_path):
    c = D.DryadLinqContext(2)
    src = c.from_enumerable(DATA, 2)
    u = f"partfile://{tmp_path}/same.pt"
    with pytest.raises(DryadLinqException) as ei:
        c.submit(src.select(lambda x: x).to_store(u), src.select(lambda x: x + 1).to_store(u))
    assert ei.value.error_code == ErrorCode.MultipleOutputsWithSameDscUri
    c.from_enumerable(DATA, 2).to_store(u).submit_and_wait()
    stored = c.from_store(u)
    with pytest.raises(DryadLinqException) as ei:
        stored.select(lambda x: x).to_store(u, delete_if_exists=True).submit()
    assert ei.value.error_code == ErrorCode.OutputUriAlsoQueryInput


def test_context_rules():
    c = D.DryadLinqContext(2)
    q = c.from_enumerable(DATA, 2).select(lambda x: x)
    assert q.to_list() is not None
    with pytest.raises(DryadLinqException):
        c.local_debug = True                      # settings are frozen after a submit
    c.dispose()
    with pytest.raises(DryadLinqException) as ei:
        c.from_enumerable(DATA)
    assert ei.value.error_code == ErrorCode.ContextDisposed
    a, b = D.DryadLinqContext(4), D.DryadLinqContext(4)
    assert a.settings_equal(b)
    b.enable_speculative_duplication = False
    assert not a.settings_equal(b)


def test_resubmit_returns_same_job_and_data_backed_requery():
    c = D.DryadLinqContext(2)
    q = c.from_enumerable(DATA, 2).select(lambda x: x + 1)
    j1 = q.submit()
    j2 = q.submit()
    assert j1 is j2
    res = j1.wait()[0]
    assert sorted(res.where(lambda x: x > 10).to_list()) == [11, 12, 13]     # data-backed: no recompute


def test_empty_inputs_and_empty_partitions():
    check_both(lambda c, src: src([]).select(lambda x: x))
    check_both(lambda c, src: src([1]).where(lambda x: x > 5).group_by(lambda x: x).select(lambda g: g.count()))
    c = D.DryadLinqContext(4)
    assert c.from_enumerable([1, 2], 4).order_by(lambda x: x).to_list() == [1, 2]
    assert c.from_enumerable([], 3).count() == 0


def test_wordcount_local_text_file(tmp_path):
    p = tmp_path / "in.txt"
    p.write_text("the quick brown fox\nthe lazy dog\nthe end\n")
    for dbg in (True, False):
        c = D.DryadLinqContext(2, local_debug=dbg)
        lines = c.from_store(f"file://{p}")
        counts = (lines.select_many(lambda l: l.split(" ")).group_by(lambda w: w)
                  .select(lambda g: D.Pair(g.key, g.count())))
        got = dict(counts.to_list())
        assert got == {"the": 3, "quick": 1, "brown": 1, "fox": 1, "lazy": 1, "dog": 1, "end": 1}


def test_context_equality_is_over_all_settings():
    """DryadLinqContext.Equals (DryadLinqContext.cs:1283-1308): equal when every setting is equal."""
    import dryad_b200 as D
    a, b = D.DryadLinqContext(3), D.DryadLinqContext(3)
    assert a.equals(b) and a.Equals(b)
    b.enable_speculative_duplication = not b.enable_speculative_duplication
    assert not a.equals(b)
    c = D.DryadLinqContext(4)
    assert not a.equals(c) and not a.equals("x")
    a.register_azure_account("acct", "key")          # recorded only: no Azure provider on this engine
