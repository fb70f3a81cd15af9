"""Logical query expressions: what the LINQ-style operators build (lazily) before planning.

Role of the C# expression trees + QueryNodeInfo graph (LinqToDryad/DryadLinqQueryGen.cs:1002-1143):
operators only record structure; sharing of a sub-query by several consumers is visible as the same
`QExpr` object reachable twice, which the planner turns into a Tee.
"""
from __future__ import annotations

import itertools
from typing import Any, Callable, Dict, List, Optional, Sequence

from ..expr import ir as X

_ids = itertools.count(1)

Schema = Optional[Dict[str, str]]   # column name -> kind ('i','f','b','o'); None = opaque objects


def _records_out(fn: Callable) -> Callable:
    """A lambda that builds `{...}` (the anonymous-type idiom, `new { k, v }` in the reference's C#)
    yields Records — dicts with attribute access — so later lambdas can write `r.k`."""
    import types
    from ..api.types import Record
    if not isinstance(fn, (types.FunctionType, types.MethodType)):
        return fn                     # callable objects (RowKey, comparers, ...) keep their identity

    def wrapped(*a):
        out = fn(*a)
        return Record(out) if type(out) is dict else out
    wrapped.__name__ = getattr(fn, "__name__", "λ")
    wrapped.__wrapped__ = fn
    return wrapped


class Lam:
    """A user lambda plus its traced form.  `ir` is an Expr (scalar result), a dict of Exprs
    (record construction) or None (opaque Python: host path only)."""

    __slots__ = ("fn", "ir", "nargs", "name")

    def __init__(self, fn: Optional[Callable], ir=None, nargs: int = 1, name: str = ""):
        self.fn, self.ir, self.nargs, self.name = fn, ir, nargs, name or getattr(fn, "__name__", "λ")

    @staticmethod
    def traced(fn: Callable, schema: Schema, nargs: int = 1) -> "Lam":
        ir = None
        if schema is not None and nargs == 1 and all(k != "o" for k in schema.values()):
            ir = X.trace(fn, schema)
        return Lam(_records_out(fn), ir, nargs, name=getattr(fn, "__name__", "λ"))

    @property
    def is_traced(self) -> bool:
        return self.ir is not None

    def out_schema(self) -> Schema:
        if isinstance(self.ir, X.Expr):
            return {"value": self.ir.kind}
        if isinstance(self.ir, dict):
            return {k: v.kind for k, v in self.ir.items()}
        return None

    def describe(self) -> str:
        if isinstance(self.ir, X.Expr):
            return X.to_string(self.ir)
        if isinstance(self.ir, dict):
            return "new {" + ", ".join(f"{k} = {X.to_string(v)}" for k, v in self.ir.items()) + "}"
        return f"<{self.name}>"


class QExpr:
    """One logical operator application."""

    def __init__(self, op: str, sources: Sequence["QExpr"] = (), schema: Schema = None, **params):
        self.id = next(_ids)
        self.op = op
        self.sources: List[QExpr] = list(sources)
        self.params: Dict[str, Any] = params
        self.schema: Schema = schema

    def __repr__(self):
        return f"QExpr#{self.id}({self.op})"

    def walk(self, seen=None):
        seen = set() if seen is None else seen
        if self.id in seen:
            return
        seen.add(self.id)
        for s in self.sources:
            yield from s.walk(seen)
        yield self


def infer_schema(sample: Any) -> Schema:
    """Schema of python records: numbers -> scalar 'value' column; dict / Record / namedtuple of
    numbers -> one column per field; anything else is opaque."""
    import numbers
    from .type_system import kind_of_value

    def kind(v):
        if isinstance(v, (bool, int, float)):
            return kind_of_value(v)          # ints outside int64 are opaque (host path)
        if isinstance(v, numbers.Integral):
            return "i"
        if isinstance(v, numbers.Real):
            return "f"
        return "o"

    if isinstance(sample, (bool, numbers.Integral, numbers.Real)):
        return {"value": kind(sample)}
    if isinstance(sample, dict) and sample and all(isinstance(k, str) for k in sample):
        return {k: kind(v) for k, v in sample.items()}
    if hasattr(sample, "_fields"):
        return {k: kind(getattr(sample, k)) for k in sample._fields}
    if type(sample) is tuple and sample and all(isinstance(v, (bool, numbers.Integral, numbers.Real)) for v in sample):
        return {f"item{i + 1}": kind(v) for i, v in enumerate(sample)}
    return None
