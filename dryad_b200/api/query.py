"""The LINQ-style lazy query surface.

Reference: the ~110 DryadLINQ-specific operators of LinqToDryad/DryadLinqQueryable.cs:54-4265 plus
the standard System.Linq.Queryable operators the provider intercepts
(LinqToDryad/DryadLinqQuery.cs:137-210, dispatcher DryadLinqQueryGen.cs:3439-4787).

Operators only build `QExpr` nodes (lazy).  Execution is triggered by iteration, `submit()`,
`submit_and_wait()`, `to_list()`, or a scalar aggregate (`count()`, `sum()` …) — the same
triggers as the reference (DryadLinqQuery.cs:171-209, 563-596).  Every operator has a LocalDebug
branch: with `context.local_debug` the query runs through the CPU oracle (LINQ-to-Objects).

Method names are snake_case; PascalCase aliases (Select, Where, GroupBy, …) exist for users
coming from the reference.
"""
from __future__ import annotations

from typing import Dict, Any, Callable, Iterable, List, Optional, Sequence

from ..expr import ir as X
from ..plan.decomposition import try_decompose
from ..plan.logical import Lam, QExpr, Schema, _records_out
from .errors import DryadLinqException, ErrorCode


def _lam(fn, schema: Schema, nargs: int = 1) -> Lam:
    if isinstance(fn, Lam):
        return fn
    if isinstance(fn, X.Expr) or isinstance(fn, dict):
        # users may pass column expressions directly: q.select({"a": col("x") + 1})
        ir = fn if isinstance(fn, X.Expr) else {k: X.wrap(v) for k, v in fn.items()}
        return Lam(_fn_from_ir(ir), ir, 1)
    if isinstance(fn, str):
        kind = (schema or {}).get(fn, "o")
        ir = X.col(fn, kind) if kind != "o" else None
        return Lam(lambda r, _f=fn: r[_f] if isinstance(r, dict) else getattr(r, _f), ir, 1)
    if not callable(fn):
        raise DryadLinqException(ErrorCode.UnsupportedExpressionType, f"expected a function, got {fn!r}")
    return Lam.traced(fn, schema, nargs)


IDX = "__idx"          # symbolic GLOBAL record index in traced two-argument lambdas ((x, i) => ...)


def _lam_indexed(fn, schema: Schema) -> Lam:
    """(x, i) => ... : traced with the index as the column `__idx`; the GPU backend binds it to
    rowid + the partition's start offset (device count -> scan, K16)."""
    ir = None
    if schema is not None and callable(fn) and all(k != "o" for k in schema.values()):
        ir = X.trace(fn, schema, X.col(IDX, "i"))
    return Lam(_records_out(fn) if callable(fn) else fn, ir, 2)


def _lam_pair(fn, sa: Schema, sb: Schema) -> Lam:
    """(a, b) => ... over two row-aligned datasets (Zip): columns are prefixed `a.` / `b.`."""
    ir = None
    if sa is not None and sb is not None and callable(fn) and all(k != "o" for k in list(sa.values()) + list(sb.values())):
        def sym(schema, prefix):
            if list(schema) == ["value"]:
                return X.col(prefix + "value", schema["value"])
            return X.RowProxy(schema, prefix=prefix)
        try:
            ir = X.normalise_traced(fn(sym(sa, "a."), sym(sb, "b.")))
        except Exception:  # noqa: BLE001  (opaque lambda: host path)
            ir = None
    return Lam(_records_out(fn), ir, 2)


def _fn_from_ir(ir):
    from .types import Record

    def row_of(r):
        return r if isinstance(r, dict) else {"value": r}
    if isinstance(ir, X.Expr):
        return lambda r: X.eval_row(ir, row_of(r))
    if isinstance(ir, X.TupleFields):
        return lambda r: ir.build(X.eval_row(e, row_of(r)) for e in ir.values())
    return lambda r: Record((k, X.eval_row(e, row_of(r))) for k, e in ir.items())


class Query:
    """IQueryable<T>: a lazy, partitioned dataset."""

    def __init__(self, context, expr: QExpr):
        self.context = context
        self.expr = expr
        self._job = None          # set once executed (data-backed afterwards)

    # ---- plumbing --------------------------------------------------------------------------
    @property
    def schema(self) -> Schema:
        return self.expr.schema

    def _new(self, op: str, sources: Sequence["Query"] = (), schema: Schema = None, **params) -> "Query":
        for s in sources:
            if s.context is not self.context:
                raise DryadLinqException(ErrorCode.MustStartFromContext,
                                         "all inputs of a query must come from the same context")
        self.context._check_not_disposed()
        return Query(self.context, QExpr(op, [self.expr] + [s.expr for s in sources], schema, **params))

    # ---- element-wise ----------------------------------------------------------------------
    def select(self, fn) -> "Query":
        if self.expr.op == "group_by" and "result" not in self.expr.params and "decomp" not in self.expr.params:
            src = self.expr
            dec = try_decompose(fn, src.params["elem_schema"], src.params["key"]) if callable(fn) else None
            if dec is not None:
                q = Query(self.context, QExpr("group_by", src.sources, dec.out_schema(), key=src.params["key"],
                                              elem=src.params.get("elem"), elem_schema=src.params["elem_schema"],
                                              comparer=src.params.get("comparer"), decomp=dec))
                return q
        lam = _lam(fn, self.schema)
        return self._new("select", schema=lam.out_schema(), fn=lam, indexed=False)

    def select_indexed(self, fn) -> "Query":
        """Select((x, i) => …) with the GLOBAL record index."""
        lam = _lam_indexed(fn, self.schema)
        return self._new("select", schema=lam.out_schema(), fn=lam, indexed=True)

    long_select = select_indexed

    def where(self, pred) -> "Query":
        return self._new("where", schema=self.schema, fn=_lam(pred, self.schema), indexed=False)

    def where_indexed(self, pred) -> "Query":
        return self._new("where", schema=self.schema, fn=_lam_indexed(pred, self.schema), indexed=True)

    long_where = where_indexed

    def select_many(self, fn, result: Optional[Callable] = None) -> "Query":
        # `l => l.Split(' ')` over a string column traces to a `split` expression: the GPU backend runs
        # it as the tokenizer kernel and the result is again a string column
        ir, schema = None, None
        if result is None and self.schema is not None and all(k != "o" for k in self.schema.values()):
            t = X.trace(fn, self.schema)
            if isinstance(t, X.Expr) and t.op == "split":
                ir, schema = t, {"value": "s"}
        return self._new("select_many", schema=schema, fn=Lam(fn, ir, 1),
                         result=Lam(result, None, 2) if result else None, indexed=False)

    def select_many_indexed(self, fn, result: Optional[Callable] = None) -> "Query":
        return self._new("select_many", schema=None, fn=Lam(fn, None, 2),
                         result=Lam(result, None, 2) if result else None, indexed=True)

    long_select_many = select_many_indexed

    # ---- grouping / ordering ---------------------------------------------------------------
    def group_by(self, key, elem=None, result=None, comparer=None) -> "Query":
        klam = _lam(key, self.schema)
        elam = _lam(elem, self.schema) if elem is not None else None
        elem_schema = elam.out_schema() if elam is not None else self.schema
        params = dict(key=klam, elem=elam, elem_schema=elem_schema, comparer=comparer)
        if result is not None:
            dec = try_decompose(result, elem_schema, klam, two_arg=True)
            if dec is not None:
                return self._new("group_by", schema=dec.out_schema(), decomp=dec, **params)
            return self._new("group_by", schema=None, result=Lam(result, None, 2), **params)
        return self._new("group_by", schema=None, **params)

    def order_by(self, key, comparer=None, descending: bool = False) -> "Query":
        return self._new("order_by", schema=self.schema, key=_lam(key, self.schema), comparer=comparer,
                         descending=descending)

    def order_by_descending(self, key, comparer=None) -> "Query":
        return self.order_by(key, comparer, True)

    def then_by(self, *_a, **_k):
        raise DryadLinqException(ErrorCode.ThenByNotSupported,
                                 "ThenBy is not supported; order by a composite key instead")

    then_by_descending = then_by

    def distinct(self, comparer=None) -> "Query":
        return self._new("distinct", schema=self.schema, comparer=comparer)

    def reverse(self) -> "Query":
        return self._new("reverse", schema=self.schema)

    # ---- binary ----------------------------------------------------------------------------
    def join(self, inner: "Query", outer_key, inner_key, result, comparer=None) -> "Query":
        ok, ik = _lam(outer_key, self.schema), _lam(inner_key, inner.schema)
        rlam = _join_result_lam(result, self.schema, inner.schema)
        return self._new("join", [inner], schema=rlam.out_schema(), outer_key=ok, inner_key=ik,
                         result=rlam, comparer=comparer)

    def group_join(self, inner: "Query", outer_key, inner_key, result, comparer=None) -> "Query":
        # (outer, group) => ... selectors that use the group only through decomposable aggregates are
        # traced: the GPU backend then aggregates the inner side by key and probes it once per outer row
        from ..plan.decomposition import try_decompose_group_join
        dec = try_decompose_group_join(result, self.schema, inner.schema) if callable(result) else None
        return self._new("group_join", [inner], schema=dec.out_schema() if dec is not None else None,
                         outer_key=_lam(outer_key, self.schema), inner_key=_lam(inner_key, inner.schema),
                         result=Lam(_records_out(result) if callable(result) else result, None, 2), comparer=comparer, gj_decomp=dec)

    def concat(self, other: "Query") -> "Query":
        return self._new("concat", [other], schema=self.schema if self.schema == other.schema else None)

    def union(self, other: "Query", comparer=None) -> "Query":
        return self._new("union", [other], schema=self.schema if self.schema == other.schema else None, comparer=comparer)

    def intersect(self, other: "Query", comparer=None) -> "Query":
        return self._new("intersect", [other], schema=self.schema, comparer=comparer)

    def except_(self, other: "Query", comparer=None) -> "Query":
        return self._new("except", [other], schema=self.schema, comparer=comparer)

    def zip(self, other: "Query", fn) -> "Query":
        lam = _lam_pair(fn, self.schema, other.schema)
        return self._new("zip", [other], schema=lam.out_schema(), fn=lam)

    # ---- prefix operators ------------------------------------------------------------------
    def take(self, n: int) -> "Query":
        return self._new("take", schema=self.schema, n=int(n))

    def skip(self, n: int) -> "Query":
        return self._new("skip", schema=self.schema, n=int(n))

    def take_while(self, pred) -> "Query":
        return self._new("take_while", schema=self.schema, fn=_lam(pred, self.schema), indexed=False)

    def skip_while(self, pred) -> "Query":
        return self._new("skip_while", schema=self.schema, fn=_lam(pred, self.schema), indexed=False)

    def take_while_indexed(self, pred) -> "Query":
        return self._new("take_while", schema=self.schema, fn=_lam_indexed(pred, self.schema), indexed=True)

    def skip_while_indexed(self, pred) -> "Query":
        return self._new("skip_while", schema=self.schema, fn=_lam_indexed(pred, self.schema), indexed=True)

    long_take_while, long_skip_while = take_while_indexed, skip_while_indexed

    # ---- partitioning (DryadLinqQueryable.cs:274-863, 3407-3677) ---------------------------
    def hash_partition(self, key=None, count: Optional[int] = None, comparer=None) -> "Query":
        klam = _lam(key, self.schema) if key is not None else _identity_lam(self.schema)
        if count is not None and count < 1:
            raise DryadLinqException(ErrorCode.IndexTooSmall, "partition count must be positive")
        return self._new("hash_partition", schema=self.schema, key=klam, count=count, comparer=comparer)

    def range_partition(self, key=None, count: Optional[int] = None, range_keys: Optional[Sequence] = None,
                        comparer=None, descending: Optional[bool] = None) -> "Query":
        klam = _lam(key, self.schema) if key is not None else _identity_lam(self.schema)
        return self._new("range_partition", schema=self.schema, key=klam, count=count,
                         range_keys=list(range_keys) if range_keys is not None else None,
                         comparer=comparer, descending=descending)

    def assume_hash_partition(self, key=None, comparer=None) -> "Query":
        klam = _lam(key, self.schema) if key is not None else _identity_lam(self.schema)
        return self._new("assume_hash_partition", schema=self.schema, key=klam, comparer=comparer)

    def assume_range_partition(self, key=None, range_keys: Optional[Sequence] = None, comparer=None,
                               descending: Optional[bool] = None) -> "Query":
        klam = _lam(key, self.schema) if key is not None else _identity_lam(self.schema)
        return self._new("assume_range_partition", schema=self.schema, key=klam,
                         range_keys=list(range_keys) if range_keys is not None else None,
                         comparer=comparer, descending=descending)

    def assume_order_by(self, key=None, comparer=None, descending: bool = False) -> "Query":
        klam = _lam(key, self.schema) if key is not None else _identity_lam(self.schema)
        return self._new("assume_order_by", schema=self.schema, key=klam, comparer=comparer, descending=descending)

    # ---- Apply / Fork (DryadLinqQueryable.cs:929-1208, 3716-3851) --------------------------
    def apply(self, fn, other: Optional["Query"] = None) -> "Query":
        """Non-homomorphic Apply: fn sees the WHOLE dataset (merged to one partition)."""
        srcs = [other] if other is not None else []
        return self._new("apply", srcs, schema=None, fn=Lam(fn, None, 1 + len(srcs)), per_partition=False,
                         first_only=False)

    def apply_per_partition(self, fn, other: Optional["Query"] = None, is_first_only: bool = False) -> "Query":
        """Homomorphic Apply: fn runs once per partition.  Binary form: is_first_only=True
        broadcasts `other` to every partition of self (left-homomorphic); False pairs partition i
        with partition i (fully homomorphic; equal partition counts required)."""
        srcs = [other] if other is not None else []
        return self._new("apply", srcs, schema=None, fn=Lam(fn, None, 1 + len(srcs)), per_partition=True,
                         first_only=is_first_only)

    def apply_with_partition_index(self, fn) -> "Query":
        return self._new("apply", schema=None, fn=Lam(fn, None, 2), per_partition=True, first_only=False,
                         with_index=True)

    def sliding_window(self, fn, window: int) -> "Query":
        # the window function sees a sequence of `window` records; it is traced with one symbolic
        # record per window slot (columns w0.*, w1.*, ...), so `lambda w: sum(w)` or
        # `lambda w: w[-1] - w[0]` become one fused kernel over shifted views of the column
        w = int(window)
        ir = None
        if self.schema is not None and w > 0 and all(k != "o" for k in self.schema.values()):
            if list(self.schema) == ["value"]:
                syms = [X.col(f"w{j}.value", self.schema["value"]) for j in range(w)]
            else:
                syms = [X.RowProxy(self.schema, prefix=f"w{j}.") for j in range(w)]
            try:
                ir = X.normalise_traced(fn(syms))
            except Exception:  # noqa: BLE001  (opaque lambda: host path)
                ir = None
        lam = Lam(fn, ir, 1)
        return self._new("sliding_window", schema=lam.out_schema(), fn=lam, window=w)

    def fork(self, fn, n: int = 2) -> "MultiQuery":
        """Mapper returns a tuple of n optional values; branch i receives the non-None i-th values.
        `fn` may also be a per-partition function returning n sequences (fork_per_partition)."""
        fk = self._new("fork", schema=None, fn=Lam(fn, None, 1), n=n, per_partition=False)
        return MultiQuery(self.context, fk, n)

    def fork_per_partition(self, fn, n: int = 2) -> "MultiQuery":
        fk = self._new("fork", schema=None, fn=Lam(fn, None, 1), n=n, per_partition=True)
        return MultiQuery(self.context, fk, n)

    def fork_by_key(self, key_fn, keys: Sequence) -> "KeyedMultiQuery":
        keys = list(keys)
        idx = {k: i for i, k in enumerate(keys)}

        def mapper(x, _idx=idx, _n=len(keys), _kf=key_fn):
            out = [None] * _n
            i = _idx.get(_kf(x))
            if i is not None:
                out[i] = x
            return tuple(out)
        fk = self._new("fork", schema=None, fn=Lam(mapper, None, 1), n=len(keys), per_partition=False)
        return KeyedMultiQuery(self.context, fk, keys)

    def broadcast(self, n=None) -> "Query":
        """BroadCast: every one of n partitions receives the whole dataset; `n` may also be a
        destination dataset, whose partition count is used (DryadLinqExtension.cs:264-281)."""
        if isinstance(n, Query):
            return n.apply_per_partition(lambda x, y: y, self, is_first_only=True)
        return self._new("broadcast", schema=self.schema, count=n)

    # ---- the user-library operators of the reference (LinqToDryad/DryadLinqExtension.cs:152-300) --------
    def map_reduce(self, mapper, key, reducer) -> "Query":
        """The standard MapReduce: SelectMany(mapper).GroupBy(key, reducer) (DryadLinqExtension.cs:166-174)."""
        return self.select_many(mapper).group_by(key, None, reducer)

    def cross_product(self, other: "Query", fn) -> "Query":
        """fn over every pair of the two datasets: `other` is broadcast to every partition of self
        (DryadLinqExtension.cs:187-193, DryadLinqHelper.Cross)."""
        return self.apply_per_partition(lambda xs, ys: [fn(x, y) for x in xs for y in ys], other, is_first_only=True)

    def check_order_by(self, key, comparer=None, descending: bool = False) -> "Query":
        """Pass-through that fails the job when a partition is not ordered by `key`
        (DryadLinqExtension.cs:293-300, DryadLinqHelper.CheckSort)."""
        kf = key.fn if isinstance(key, Lam) else key
        import functools

        def check(xs):
            prev, have = None, False
            for x in xs:
                k = kf(x)
                if have:
                    c = comparer(prev, k) if comparer is not None else (-1 if prev < k else (1 if k < prev else 0))
                    if (c > 0 and not descending) or (c < 0 and descending):
                        raise DryadLinqException(ErrorCode.FailureInUserApplyFunction,
                                                 f"partition is not ordered: {prev!r} before {k!r}")
                prev, have = k, True
                yield x
        functools.update_wrapper(check, kf) if hasattr(kf, "__name__") else None
        return self.apply_per_partition(lambda xs: list(check(xs)))

    def do_while(self, body: Callable[["Query"], "Query"], cond: Callable[["Query", "Query"], "Query"], count: int = 1) -> "Query":
        """Client-side loop (DryadLinqQueryable.cs:1280-1306): each iteration submits body and
        cond as one job; cond must produce a single bool.  `count` unrolls the body that many times per
        job (DryadLinqExtension.cs:204-231); 0 returns the source."""
        if count < 0:
            raise ValueError("count")
        if count == 0:
            return self
        inner = body
        if count > 1:
            def body(q, _inner=inner, _n=count):     # noqa: F811  (the unrolled body)
                for _ in range(_n):
                    q = _inner(q)
                return q
        cur = self
        while True:
            nxt = body(cur)
            c = cond(cur, nxt)
            nxt_m, c_m = self.context.submit_and_wait(nxt.materialize_lazy(), c.materialize_lazy())
            ok = bool(c_m.to_list()[0])
            cur = nxt_m
            if not ok:
                return cur

    # ---- scalar aggregates as datasets (*AsQuery, DryadLinqQueryable.cs:1384-3368) ---------
    def _agg(self, kind: str, **params) -> "Query":
        return self._new("aggregate", schema=None, kind=kind, **params)

    def count_as_query(self, pred=None): return self._agg("count", fn=_lam(pred, self.schema) if pred else None)
    def long_count_as_query(self, pred=None): return self.count_as_query(pred)
    def sum_as_query(self, sel=None): return self._agg("sum", fn=_lam(sel, self.schema) if sel else None)
    def min_as_query(self, sel=None): return self._agg("min", fn=_lam(sel, self.schema) if sel else None)
    def max_as_query(self, sel=None): return self._agg("max", fn=_lam(sel, self.schema) if sel else None)
    def average_as_query(self, sel=None): return self._agg("average", fn=_lam(sel, self.schema) if sel else None)
    def any_as_query(self, pred=None): return self._agg("any", fn=_lam(pred, self.schema) if pred else None)
    def all_as_query(self, pred): return self._agg("all", fn=_lam(pred, self.schema))
    def contains_as_query(self, value): return self._agg("contains", value=value)
    def first_as_query(self, pred=None): return self._agg("first", fn=_lam(pred, self.schema) if pred else None, default=False)
    def first_or_default_as_query(self, pred=None): return self._agg("first", fn=_lam(pred, self.schema) if pred else None, default=True)
    def last_as_query(self, pred=None): return self._agg("last", fn=_lam(pred, self.schema) if pred else None, default=False)
    def last_or_default_as_query(self, pred=None): return self._agg("last", fn=_lam(pred, self.schema) if pred else None, default=True)
    def single_as_query(self, pred=None): return self._agg("single", fn=Lam(pred, None, 1) if pred else None, default=False)
    def single_or_default_as_query(self, pred=None): return self._agg("single", fn=Lam(pred, None, 1) if pred else None, default=True)

    def sequence_equal_as_query(self, other: "Query"):
        return self._new("aggregate", [other], schema=None, kind="sequence_equal")

    def aggregate_as_query(self, func, seed=None, has_seed: Optional[bool] = None, result=None, seed_fn=None):
        hs = has_seed if has_seed is not None else (seed is not None or seed_fn is not None)
        return self._agg("aggregate", fn=Lam(func, None, 2), seed=seed, seed_fn=seed_fn, has_seed=hs,
                         result=Lam(result, None, 1) if result else None)

    # ---- scalar aggregates (execute immediately) --------------------------------------------
    def _scalar(self, q: "Query"):
        vals = q.to_list()
        return vals[0] if vals else None

    def count(self, pred=None): return self._scalar(self.count_as_query(pred))
    long_count = count
    def sum(self, sel=None): return self._scalar(self.sum_as_query(sel))
    def min(self, sel=None): return self._scalar(self.min_as_query(sel))
    def max(self, sel=None): return self._scalar(self.max_as_query(sel))
    def average(self, sel=None): return self._scalar(self.average_as_query(sel))
    def any(self, pred=None): return self._scalar(self.any_as_query(pred))
    def all(self, pred): return self._scalar(self.all_as_query(pred))
    def contains(self, value): return self._scalar(self.contains_as_query(value))
    def first(self, pred=None): return self._scalar(self.first_as_query(pred))
    def first_or_default(self, pred=None): return self._scalar(self.first_or_default_as_query(pred))
    def last(self, pred=None): return self._scalar(self.last_as_query(pred))
    def last_or_default(self, pred=None): return self._scalar(self.last_or_default_as_query(pred))
    def single(self, pred=None): return self._scalar(self.single_as_query(pred))
    def single_or_default(self, pred=None): return self._scalar(self.single_or_default_as_query(pred))
    def sequence_equal(self, other): return self._scalar(self.sequence_equal_as_query(other))

    def aggregate(self, func, seed=None, has_seed: Optional[bool] = None, result=None, seed_fn=None):
        return self._scalar(self.aggregate_as_query(func, seed, has_seed, result, seed_fn))

    def element_at(self, *_a):
        raise DryadLinqException(ErrorCode.OperatorNotSupported, "ElementAt is not supported")

    def default_if_empty(self, *_a):
        raise DryadLinqException(ErrorCode.OperatorNotSupported, "DefaultIfEmpty is not supported")

    def of_type(self, *_a):
        raise DryadLinqException(ErrorCode.OperatorNotSupported, "OfType is not supported")

    # ---- outputs / execution ---------------------------------------------------------------
    def to_store(self, uri: str, delete_if_exists: bool = False, compression: Optional[str] = None) -> "Query":
        """Lazy: names the output table; run it with submit()/submit_and_wait()."""
        self.context._check_store_target(uri, delete_if_exists)
        return self._new("to_store", schema=self.schema, uri=uri, compression=compression, temporary=False)

    def materialize_lazy(self) -> "Query":
        """A temporary output (ToTemporary role, DryadLinqQuery.cs:586-596)."""
        if self.expr.op == "to_store":
            return self
        return self._new("to_store", schema=self.schema, uri=None, compression=None, temporary=True)

    def submit(self):
        return self.context.submit(self)

    def submit_and_wait(self):
        return self.context.submit_and_wait(self)[0]

    def to_list(self) -> List:
        return self.context._execute_to_list(self)

    def __iter__(self):
        return iter(self.to_list())

    def to_columns(self) -> Dict[str, Any]:
        """Columnar egress (the counterpart of `from_columns`): runs the query and returns
        name -> tensor for the partitions this process holds, concatenated in partition order.  On
        the GPU backend the tensors stay in HBM — nothing is converted to Python objects."""
        from ..runtime.columnar import ColumnBatch
        res = self.context._execute(self)
        t = res.expr.params["table"]
        parts = [p for p in t.partitions if p is not None]
        batches = []
        for p in parts:
            if isinstance(p, ColumnBatch):
                batches.append(p)
            elif len(p):
                batches.append(ColumnBatch.from_records(list(p), t.schema, "cpu"))
        if not batches:
            return {}
        return dict(ColumnBatch.concat(batches).cols)

    def partitions(self) -> List[List]:
        """Materialised per-partition contents (structural partition checks in tests)."""
        return self.context._execute_to_partitions(self)

    def explain(self) -> str:
        from ..plan.querygen import QueryGen
        return QueryGen(self.context, [self.materialize_lazy()]).explain()

    def to_xml(self) -> str:
        from ..plan.querygen import QueryGen
        return QueryGen(self.context, [self.materialize_lazy()]).to_xml()

    # PascalCase aliases for users of the reference API
    Select, Where, SelectMany, GroupBy, OrderBy, OrderByDescending = select, where, select_many, group_by, order_by, order_by_descending
    Distinct, Reverse, Join, GroupJoin, Concat, Union, Intersect, Except, Zip = distinct, reverse, join, group_join, concat, union, intersect, except_, zip
    Take, Skip, TakeWhile, SkipWhile = take, skip, take_while, skip_while
    HashPartition, RangePartition, Apply, ApplyPerPartition, Fork, SlidingWindow = hash_partition, range_partition, apply, apply_per_partition, fork, sliding_window
    AssumeHashPartition, AssumeRangePartition, AssumeOrderBy = assume_hash_partition, assume_range_partition, assume_order_by
    LongSelect, LongWhere, LongSelectMany, LongTakeWhile, LongSkipWhile = select_indexed, where_indexed, select_many_indexed, take_while_indexed, skip_while_indexed
    Count, LongCount, Sum, Min, Max, Average, Any, All, Contains = count, count, sum, min, max, average, any, all, contains
    First, FirstOrDefault, Last, LastOrDefault, Single, SingleOrDefault = first, first_or_default, last, last_or_default, single, single_or_default
    Aggregate, SequenceEqual, ToStore, Submit, SubmitAndWait, DoWhile, BroadCast = aggregate, sequence_equal, to_store, submit, submit_and_wait, do_while, broadcast
    MapReduce, CrossProduct, CheckOrderBy, ApplyWithPartitionIndex = map_reduce, cross_product, check_order_by, apply_with_partition_index
    LongCountAsQuery, SequenceEqualAsQuery = long_count_as_query, sequence_equal_as_query
    CountAsQuery, SumAsQuery, MinAsQuery, MaxAsQuery, AverageAsQuery = count_as_query, sum_as_query, min_as_query, max_as_query, average_as_query
    AnyAsQuery, AllAsQuery, ContainsAsQuery, FirstAsQuery, LastAsQuery, SingleAsQuery, AggregateAsQuery = any_as_query, all_as_query, contains_as_query, first_as_query, last_as_query, single_as_query, aggregate_as_query


def _identity_lam(schema: Schema) -> Lam:
    ir = None
    if schema is not None and list(schema) == ["value"] and schema["value"] != "o":
        ir = X.col("value", schema["value"])
    return Lam(lambda x: x, ir, 1, "identity")


def _join_result_lam(result, outer_schema: Schema, inner_schema: Schema) -> Lam:
    """Trace a (outer, inner) => record selector: outer columns are prefixed 'o.', inner 'i.'."""
    if isinstance(result, Lam):
        return result
    ir = None
    if outer_schema is not None and inner_schema is not None and \
            all(k != "o" for k in list(outer_schema.values()) + list(inner_schema.values())):
        try:
            o = X.RowProxy(outer_schema, "o.") if list(outer_schema) != ["value"] else X.col("o.value", outer_schema["value"])
            i = X.RowProxy(inner_schema, "i.") if list(inner_schema) != ["value"] else X.col("i.value", inner_schema["value"])
            ir = X.normalise_traced(result(o, i))
        except Exception:
            ir = None
    return Lam(_records_out(result), ir, 2)       # a selector that builds `{...}` yields Records on the host path too


class MultiQuery:
    """IMultiQueryable: handles to the branches of a Fork (LinqToDryad/MultiQueryable.cs:38-389)."""

    def __init__(self, context, fork_query: Query, n: int):
        self.context, self._fork, self.n = context, fork_query, n
        self._branches = [fork_query._new("fork_branch", schema=None, index=i) for i in range(n)]

    def __getitem__(self, i: int) -> Query:
        if not 0 <= i < self.n:
            raise DryadLinqException(ErrorCode.IndexOutOfRange, f"fork branch {i} out of range")
        return self._branches[i]

    def __iter__(self):
        return iter(self._branches)

    first = property(lambda s: s[0])
    second = property(lambda s: s[1])
    third = property(lambda s: s[2])
    First, Second, Third = first, second, third


class KeyedMultiQuery(MultiQuery):
    """IKeyedMultiQueryable: branches addressed by key."""

    def __init__(self, context, fork_query: Query, keys: Sequence):
        super().__init__(context, fork_query, len(keys))
        self.keys = list(keys)

    def by_key(self, key) -> Query:
        try:
            return self._branches[self.keys.index(key)]
        except ValueError:
            raise DryadLinqException(ErrorCode.MultiQueryableKeyOutOfRange, f"no fork branch for key {key!r}") from None
