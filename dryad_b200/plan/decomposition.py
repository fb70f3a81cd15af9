"""Decomposition of group result selectors into (seed, accumulate, recursive-accumulate,
final-reduce) — reference: LinqToDryad/DryadLinqDecomposition.cs:34-922.

`group_by(k).select(lambda g: f(g))` is traced with a symbolic group.  Every aggregate the lambda
asks of the group (Count, Sum, Min, Max, Average, Any, All, First, Last, a user IDecomposable)
becomes an `agg` leaf; the rest of the lambda is ordinary expression structure over those leaves
and `g.key`.  That is exactly the information the reference extracts to build the partial /
combine / final lambdas (built-ins :292-593, user [Decomposable] types :597-677, several
aggregates combined in one selector :681-760), and it maps 1:1 onto hash-aggregate lanes on the GPU.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional

from ..expr import ir as X
from ..oracle.linq import checked_add
from .logical import Lam, Schema


@dataclass
class AggSpec:
    kind: str                     # count | sum | min | max | any | all | first | last | user
    selector: Optional[Lam]       # element -> value (None = element itself / not needed)
    out_kind: str                 # 'i' | 'f' | 'b' | 'o'
    decomposer: Any = None        # IDecomposable instance for kind == 'user'

    # ---- host semantics ---------------------------------------------------------------------
    def _val(self, x):
        return self.selector.fn(x) if self.selector is not None else x

    def seed(self, x):
        k = self.kind
        if k == "count":
            return 1
        if k in ("sum", "min", "max", "first", "last"):
            return self._val(x)
        if k in ("any", "all"):
            return bool(self._val(x)) if self.selector is not None else True
        if k == "user":
            return self.decomposer.seed(self._val(x))
        raise ValueError(k)

    def accumulate(self, acc, x):
        k = self.kind
        if k == "count":
            return acc + 1
        if k == "sum":
            v = self._val(x)
            return checked_add(acc, v) if isinstance(v, int) and isinstance(acc, int) else acc + v
        if k == "min":
            return min(acc, self._val(x))
        if k == "max":
            return max(acc, self._val(x))
        if k == "any":
            return acc or (bool(self._val(x)) if self.selector is not None else True)
        if k == "all":
            return acc and bool(self._val(x))
        if k == "first":
            return acc
        if k == "last":
            return self._val(x)
        if k == "user":
            return self.decomposer.accumulate(acc, self._val(x))
        raise ValueError(k)

    def combine(self, a, b):
        k = self.kind
        if k == "count":
            return a + b
        if k == "sum":
            return checked_add(a, b) if isinstance(a, int) and isinstance(b, int) else a + b
        if k == "min":
            return min(a, b)
        if k == "max":
            return max(a, b)
        if k == "any":
            return a or b
        if k == "all":
            return a and b
        if k == "first":
            return a
        if k == "last":
            return b
        if k == "user":
            return self.decomposer.recursive_accumulate(a, b)
        raise ValueError(k)

    def finalize(self, acc):
        return self.decomposer.final_reduce(acc) if self.kind == "user" else acc

    # ---- GPU mapping ------------------------------------------------------------------------
    def gpu_lane(self) -> Optional[str]:
        """Hash-aggregate lane op, or None when this aggregate needs the host path."""
        if self.kind == "count":
            return "count"
        if self.selector is not None and not isinstance(self.selector.ir, X.Expr):
            return None
        vk = self.out_kind
        if self.kind == "sum":
            return "sum_f64" if vk == "f" else "sum_i64_checked"
        if self.kind == "min":
            return "min_f64" if vk == "f" else "min_i64"
        if self.kind == "max":
            return "max_f64" if vk == "f" else "max_i64"
        return None

    def describe(self) -> str:
        s = self.selector.describe() if self.selector is not None else ""
        return f"{self.kind.capitalize()}({s})"


@dataclass
class DecompositionInfo:
    aggs: List[AggSpec]
    result: Any                    # Expr or dict[str, Expr] over `agg` leaves and key columns
    key_schema: Schema

    def final(self, key, accs: List[Any]):
        """Apply the final reducer on the host: evaluate `result` with leaves bound."""
        env = _LeafEnv(key, [a.finalize(v) for a, v in zip(self.aggs, accs)], self.key_schema)
        if isinstance(self.result, X.TupleFields):
            return self.result.build(_eval_with_leaves(e, env) for e in self.result.values())
        if isinstance(self.result, dict):
            from ..api.types import Record
            return Record((k, _eval_with_leaves(e, env)) for k, e in self.result.items())
        return _eval_with_leaves(self.result, env)

    def out_schema(self) -> Schema:
        if isinstance(self.result, dict):
            return {k: e.kind for k, e in self.result.items()}
        return {"value": self.result.kind}

    def gpu_lanes(self) -> Optional[List[str]]:
        lanes = [a.gpu_lane() for a in self.aggs]
        return None if any(l is None for l in lanes) else lanes  # type: ignore[return-value]

    def describe(self) -> str:
        if isinstance(self.result, dict):
            body = "new {" + ", ".join(f"{k} = {X.to_string(e)}" for k, e in self.result.items()) + "}"
        else:
            body = X.to_string(self.result)
        return f"{body} where " + ", ".join(f"agg{i} = {a.describe()}" for i, a in enumerate(self.aggs))


class _LeafEnv:
    def __init__(self, key, vals, key_schema):
        self.key, self.vals, self.key_schema = key, vals, key_schema


def _eval_with_leaves(e: X.Expr, env: _LeafEnv):
    if e.op == "agg":
        return env.vals[e.args[0]]
    if e.op == "col":
        name = e.args[0]
        if name == "key":
            return env.key
        if name.startswith("key."):
            k = env.key
            f = name[4:]
            return k[f] if isinstance(k, dict) else getattr(k, f)
        raise X.TraceError(f"unbound column {name}")
    if e.op in ("lit", "rowid"):
        return X.eval_row(e, {})
    # generic: evaluate children then re-dispatch through eval_row on literals
    args = tuple(X.lit(_eval_with_leaves(a, env)) if isinstance(a, X.Expr) else a for a in e.args)
    return X.eval_row(X.Expr(e.op, args, e.kind), {})


class GroupProxy:
    """Symbolic IGrouping handed to group result selectors while tracing."""

    def __init__(self, elem_schema: Schema, key_ir, aggs: List[AggSpec]):
        self._schema, self._key_ir, self._aggs = elem_schema, key_ir, aggs

    # -- key -------------------------------------------------------------------------------------
    @property
    def key(self):
        k = self._key_ir
        if isinstance(k, X.Expr):
            return X.col("key", k.kind)
        if isinstance(k, dict):
            return X.RowProxy({n: e.kind for n, e in k.items()}, prefix="key.")
        return X.col("key", "o")

    Key = key

    # -- helpers ----------------------------------------------------------------------------------
    def _lam(self, fn) -> Optional[Lam]:
        if fn is None:
            if self._schema is not None and list(self._schema) == ["value"]:
                return Lam(lambda x: x, X.col("value", self._schema["value"]))
            return None
        return Lam.traced(fn, self._schema)

    def _kind_of(self, lam: Optional[Lam], default="o") -> str:
        if lam is not None and isinstance(lam.ir, X.Expr):
            return lam.ir.kind
        return default

    def _leaf(self, kind: str, sel: Optional[Lam], out_kind: str, decomposer=None) -> X.Expr:
        # reuse an identical aggregate (same kind + same selector expression)
        for i, a in enumerate(self._aggs):
            if a.kind == kind and a.decomposer is decomposer:
                same = (a.selector is None and sel is None) or (
                    a.selector is not None and sel is not None and isinstance(a.selector.ir, X.Expr)
                    and isinstance(sel.ir, X.Expr) and a.selector.ir.same_as(sel.ir))
                if same:
                    return X.Expr("agg", (i,), a.out_kind)
        self._aggs.append(AggSpec(kind, sel, out_kind, decomposer))
        return X.Expr("agg", (len(self._aggs) - 1,), out_kind)

    # -- built-in decomposable aggregates (DryadLinqDecomposition.cs:292-593) ----------------
    def count(self, pred=None):
        if pred is not None:
            lam = self._lam(pred)
            if lam is None or not isinstance(lam.ir, X.Expr):
                raise X.TraceError("Count(predicate) needs a traceable predicate")
            one = Lam(lambda x, _p=pred: 1 if _p(x) else 0, X.where(lam.ir, 1, 0))
            return self._leaf("sum", one, "i")
        return self._leaf("count", None, "i")

    long_count = count
    Count = count
    LongCount = count

    def sum(self, sel=None):
        lam = self._lam(sel)
        return self._leaf("sum", lam, self._kind_of(lam))

    def min(self, sel=None):
        lam = self._lam(sel)
        return self._leaf("min", lam, self._kind_of(lam))

    def max(self, sel=None):
        lam = self._lam(sel)
        return self._leaf("max", lam, self._kind_of(lam))

    def average(self, sel=None):
        lam = self._lam(sel)
        s = self._leaf("sum", lam, self._kind_of(lam))
        c = self._leaf("count", None, "i")
        if s.kind == "o":
            return X.Expr("div", (s, c), "o")
        return X.cast(s, "f") / X.cast(c, "f")

    def any(self, pred=None):
        lam = self._lam(pred) if pred is not None else None
        return self._leaf("any", lam, "b")

    def all(self, pred):
        return self._leaf("all", self._lam(pred), "b")

    def first(self):
        lam = self._lam(None)
        return self._leaf("first", lam, self._kind_of(lam))

    def last(self):
        lam = self._lam(None)
        return self._leaf("last", lam, self._kind_of(lam))

    def aggregate(self, decomposer, sel=None):
        """User IDecomposable instance (or class) applied to the group (optionally to sel(x))."""
        from ..api.decomposable import IDecomposable
        inst = decomposer() if isinstance(decomposer, type) else decomposer
        if not isinstance(inst, IDecomposable):
            from ..api.errors import DryadLinqException, ErrorCode
            raise DryadLinqException(ErrorCode.DecomposerTypeDoesNotImplementInterface,
                                     "decomposer must implement IDecomposable")
        inst.initialize(None)
        lam = Lam(sel, None) if sel is not None else None
        return self._leaf("user", lam, "o", inst)

    Sum, Min, Max, Average, Any, All, First, Last, Aggregate = sum, min, max, average, any, all, first, last, aggregate

    def __iter__(self):
        raise X.TraceError("enumerating a group is not decomposable")

    def __len__(self):
        raise X.TraceError("len(group) is not traceable; use g.count()")


def try_decompose(fn: Callable, elem_schema: Schema, key_lam: Lam, two_arg: bool = False) -> Optional[DecompositionInfo]:
    """Trace a group result selector.  `two_arg` = the (key, group) => ... form.  Also honours
    functions marked with @decomposable(T): g => fn(g) becomes a single user aggregate."""
    dec_type = getattr(fn, "__dryad_decomposable__", None)
    aggs: List[AggSpec] = []
    g = GroupProxy(elem_schema, key_lam.ir, aggs)
    try:
        if dec_type is not None and not two_arg:
            out = g.aggregate(dec_type)
        else:
            out = fn(g.key, g) if two_arg else fn(g)
    except X.TraceError:
        return None
    except Exception:
        return None
    if isinstance(out, GroupProxy):
        return None
    norm = X.normalise_traced(out)
    if norm is None or not aggs and not _uses_only_key(norm):
        return None
    if not aggs:
        return None
    key_schema = key_lam.out_schema()
    return DecompositionInfo(aggs, norm, key_schema)


def try_decompose_group_join(fn: Callable, outer_schema: Schema, inner_schema: Schema) -> Optional[DecompositionInfo]:
    """Trace a GroupJoin result selector (outer, group) => ...: outer fields appear as columns
    `o.<name>` (`o.value` for scalars), the matching inner records only through decomposable
    aggregates of the group.  None when the selector enumerates the group or is opaque."""
    if outer_schema is None or inner_schema is None:
        return None
    if any(k == "o" for k in list(outer_schema.values()) + list(inner_schema.values())):
        return None
    aggs: List[AggSpec] = []
    g = GroupProxy(inner_schema, None, aggs)
    osym = X.col("o.value", outer_schema["value"]) if list(outer_schema) == ["value"] else X.RowProxy(outer_schema, prefix="o.")
    try:
        out = fn(osym, g)
    except Exception:  # noqa: BLE001  (TraceError or anything the lambda does to a symbol)
        return None
    norm = X.normalise_traced(out)
    if norm is None or not aggs:
        return None
    exprs = norm.values() if isinstance(norm, dict) else [norm]
    if any(c == "key" or c.startswith("key.") for e in exprs for c in e.columns()):
        return None
    return DecompositionInfo(aggs, norm, None)


def _uses_only_key(norm) -> bool:
    exprs = norm.values() if isinstance(norm, dict) else [norm]
    return all(all(c == "key" or c.startswith("key.") for c in e.columns()) for e in exprs)
