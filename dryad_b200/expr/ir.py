"""Column-expression IR — the framework's counterpart of the C# expression trees DryadLINQ
inspects (LinqToDryad/DryadLinqExpression.cs, ExpressionMatcher.cs, ExpressionSimplifier.cs).

User lambdas are *traced*: they are called once with a symbolic row whose fields are `Col`
expressions, and whatever arithmetic / comparison structure they build becomes an `Expr` tree the
planner can reason about (key equality for shuffle elision, decomposable aggregates, pushing Where
through projections) and the GPU can run as a fused register program.  A lambda that cannot be
traced (data-dependent Python control flow, arbitrary objects) stays an opaque Python UDF and runs
on the host path — the documented slow path.
"""
from __future__ import annotations

import math
import operator
from typing import Any, Callable, Dict, Iterable, Mapping, Optional, Sequence, Tuple

import numpy as np


class TraceError(Exception):
    """Raised when a lambda does something the IR cannot represent."""


_ARITH = {"add", "sub", "mul", "div", "mod", "floordiv"}
_CMP = {"eq", "ne", "lt", "le", "gt", "ge"}
_LOGIC = {"and", "or"}
_BIT = {"band", "bor", "bxor", "shl", "shr"}


class Expr:
    """Immutable expression node.  `kind` is 'i' (int64), 'f' (float64), 'b' (bool), 's' (string
    column: views into a byte buffer, runtime/columnar.py StrColumn), 'S' (sequence of strings per
    row: the result of `split`) or 'o' (opaque object, host only)."""

    __slots__ = ("op", "args", "kind", "_key")

    def __init__(self, op: str, args: tuple, kind: str):
        self.op, self.args, self.kind = op, args, kind
        self._key = None

    # -- structural identity (ExpressionMatcher role) ------------------------------------------
    def key(self) -> tuple:
        if self._key is None:
            self._key = (self.op, self.kind) + tuple(a.key() if isinstance(a, Expr) else ("#", repr(a)) for a in self.args)
        return self._key

    def same_as(self, other: "Expr") -> bool:
        return isinstance(other, Expr) and self.key() == other.key()

    def __hash__(self):
        return hash(self.key())

    def columns(self) -> set:
        out = set()
        if self.op == "col":
            out.add(self.args[0])
        for a in self.args:
            if isinstance(a, Expr):
                out |= a.columns()
        return out

    def substitute(self, mapping: Mapping[str, "Expr"]) -> "Expr":
        """Replace column references by expressions (pushes an expression through a projection;
        DataSetInfo.Rewrite role, LinqToDryad/DataSetInfo.cs:226-449)."""
        if self.op == "col":
            return mapping.get(self.args[0], self)
        return Expr(self.op, tuple(a.substitute(mapping) if isinstance(a, Expr) else a for a in self.args), self.kind)

    # -- operators -----------------------------------------------------------------------------
    def _bin(self, other, op, rev=False):
        o = wrap(other)
        a, b = (o, self) if rev else (self, o)
        if op in _CMP:
            return Expr(op, (a, b), "b")
        if op in _LOGIC:
            return Expr(op, (a, b), "b")
        if op in _BIT:
            return Expr(op, (a, b), "i")
        if a.kind == "o" or b.kind == "o":
            # opaque operands (strings, objects, untyped aggregates): host evaluation only
            return Expr("truediv_o" if op == "truediv" else op, (a, b), "o")
        if op == "floordiv":                              # `//`: floored, on integers and on floats
            if "f" in (a.kind, b.kind):
                return Expr("floor", (Expr("div", (cast(a, "f"), cast(b, "f")), "f"),), "f")
            return Expr("div", (a, b), "i")
        if op == "div":                                   # an existing node being rebuilt: integer = floored, float = true division
            k = "f" if "f" in (a.kind, b.kind) else "i"
        elif op == "truediv":
            return Expr("div", (cast(a, "f"), cast(b, "f")), "f")
        else:
            k = "f" if "f" in (a.kind, b.kind) else "i"
        if k == "f":
            a, b = cast(a, "f"), cast(b, "f")
        return Expr(op, (a, b), k)

    def __add__(self, o): return self._bin(o, "add")
    def __radd__(self, o): return self._bin(o, "add", True)
    def __sub__(self, o): return self._bin(o, "sub")
    def __rsub__(self, o): return self._bin(o, "sub", True)
    def __mul__(self, o): return self._bin(o, "mul")
    def __rmul__(self, o): return self._bin(o, "mul", True)
    def __truediv__(self, o): return self._bin(o, "truediv")
    def __rtruediv__(self, o): return self._bin(o, "truediv", True)
    def __floordiv__(self, o): return self._bin(o, "floordiv")
    def __rfloordiv__(self, o): return self._bin(o, "floordiv", True)
    def __mod__(self, o): return self._bin(o, "mod")
    def __rmod__(self, o): return self._bin(o, "mod", True)
    def __neg__(self): return Expr("neg", (self,), self.kind)
    def __abs__(self): return Expr("abs", (self,), self.kind)
    def __eq__(self, o): return self._bin(o, "eq")      # type: ignore[override]
    def __ne__(self, o): return self._bin(o, "ne")      # type: ignore[override]
    def __lt__(self, o): return self._bin(o, "lt")
    def __le__(self, o): return self._bin(o, "le")
    def __gt__(self, o): return self._bin(o, "gt")
    def __ge__(self, o): return self._bin(o, "ge")
    def __and__(self, o):
        o = wrap(o)
        return self._bin(o, "and" if self.kind == "b" and o.kind == "b" else "band")
    def __rand__(self, o): return wrap(o).__and__(self)
    def __or__(self, o):
        o = wrap(o)
        return self._bin(o, "or" if self.kind == "b" and o.kind == "b" else "bor")
    def __ror__(self, o): return wrap(o).__or__(self)
    def __xor__(self, o): return self._bin(o, "bxor")
    def split(self, sep: str = " "):
        """`line.split(" ")` on a string column: a sequence of strings per row (kind 'S'), the
        SelectMany body of WordCount (samples/WordCount.cs.pp:72).  One single-byte separator."""
        if self.kind != "s" or not isinstance(sep, str) or len(sep.encode("utf-8")) != 1:
            raise TraceError("split needs a string column and a one-byte separator")
        return Expr("split", (self, sep), "S")

    def __lshift__(self, o): return self._bin(o, "shl")
    def __rshift__(self, o): return self._bin(o, "shr")
    def __invert__(self):
        return Expr("not", (self,), "b") if self.kind == "b" else Expr("bnot", (self,), "i")

    def __bool__(self):
        raise TraceError("data-dependent Python control flow cannot be traced; use where()/select()")

    def __iter__(self):
        raise TraceError("symbolic values are not iterable")

    # -- helpers -------------------------------------------------------------------------------
    def cast(self, kind: str) -> "Expr":
        return cast(self, kind)

    def hash(self) -> "Expr":
        return Expr("hash", (self,), "i")

    def sqrt(self) -> "Expr":
        return Expr("sqrt", (cast(self, "f"),), "f")

    def floor(self) -> "Expr":
        return Expr("floor", (cast(self, "f"),), "f")

    def minimum(self, o) -> "Expr":
        return self._bin(o, "min")

    def maximum(self, o) -> "Expr":
        return self._bin(o, "max")

    def is_in(self, values: Iterable) -> "Expr":
        e: Optional[Expr] = None
        for v in values:
            t = self == v
            e = t if e is None else (e | t)
        return e if e is not None else lit(False)

    def __repr__(self):
        return to_string(self)


def col(name: str, kind: str = "i") -> Expr:
    return Expr("col", (name,), kind)


def lit(v: Any) -> Expr:
    if isinstance(v, Expr):
        return v
    if isinstance(v, (bool, np.bool_)):
        return Expr("lit", (bool(v),), "b")
    if isinstance(v, (int, np.integer)):
        return Expr("lit", (int(v),), "i")
    if isinstance(v, (float, np.floating)):
        return Expr("lit", (float(v),), "f")
    return Expr("lit", (v,), "o")


def wrap(v: Any) -> Expr:
    return v if isinstance(v, Expr) else lit(v)


def cast(e: Expr, kind: str) -> Expr:
    e = wrap(e)
    if e.kind == kind:
        return e
    if e.op == "lit" and kind in "if" and e.kind in "ifb":
        return lit(float(e.args[0]) if kind == "f" else int(e.args[0]))
    return Expr("cast", (e,), kind)


def where(cond, a, b) -> Expr:
    c, a, b = wrap(cond), wrap(a), wrap(b)
    k = "f" if "f" in (a.kind, b.kind) else a.kind
    if k == "f":
        a, b = cast(a, "f"), cast(b, "f")
    return Expr("select", (c, a, b), k)


def retype(e: Expr, col_kinds: Mapping[str, str]) -> Expr:
    """Rebuild `e` with column kinds taken from `col_kinds` (kinds of interior nodes follow)."""
    if not isinstance(e, Expr):
        return e
    if e.op == "col":
        return col(e.args[0], col_kinds.get(e.args[0], e.kind))
    if e.op in ("lit", "rowid"):
        return e
    a = [retype(x, col_kinds) if isinstance(x, Expr) else x for x in e.args]
    if e.op in _ARITH | _CMP | _LOGIC | _BIT | {"min", "max"}:
        return a[0]._bin(a[1], e.op)
    if e.op == "select":
        return where(a[0], a[1], a[2])
    if e.op == "cast":
        return cast(a[0], e.kind)
    if e.op in ("neg", "abs"):
        return Expr(e.op, (a[0],), a[0].kind)
    return Expr(e.op, tuple(a), e.kind)


def probe(table: str, key) -> Expr:
    """Row of the build side of hash table `table` whose key equals `key`, or -1 (GPU-only node:
    the probe of a unique-key join fused into the consuming pipeline, expr/codegen.py)."""
    return Expr("probe", (table, wrap(key)), "i")


def probe_slot(table: str, key) -> Expr:
    """Like `probe`, but returns the table SLOT of the key (or -1): side arrays permuted into slot
    order (`BroadcastJoin.slot_payload`) are then gathered without the extra slot -> row hop."""
    return Expr("probe_slot", (table, wrap(key)), "i")


def gather(array: str, index, kind: str = "i") -> Expr:
    """array[index] for a side array that is not row-aligned with the inputs (0 where index < 0)."""
    return Expr("gather", (array, wrap(index)), kind)


def rowid() -> Expr:
    """Position of the row inside its partition (indexed Select/Where support)."""
    return Expr("rowid", (), "i")


_SYM = {"add": "+", "sub": "-", "mul": "*", "div": "/", "mod": "%", "eq": "==", "ne": "!=",
        "lt": "<", "le": "<=", "gt": ">", "ge": ">=", "and": "&&", "or": "||", "band": "&",
        "bor": "|", "bxor": "^", "shl": "<<", "shr": ">>", "concat": "+"}


def to_string(e: Expr) -> str:
    """C#-flavoured rendering used by plan explain (DryadLinqExpression.ToCSharpString role)."""
    if e.op == "col":
        return str(e.args[0])
    if e.op == "lit":
        return repr(e.args[0])
    if e.op in _SYM:
        return f"({to_string(e.args[0])} {_SYM[e.op]} {to_string(e.args[1])})"
    if e.op == "select":
        return f"({to_string(e.args[0])} ? {to_string(e.args[1])} : {to_string(e.args[2])})"
    if e.op == "cast":
        return f"(({ {'i': 'long', 'f': 'double', 'b': 'bool', 'o': 'object'}[e.kind] }){to_string(e.args[0])})"
    if e.op == "rowid":
        return "index"
    return f"{e.op}({', '.join(to_string(a) if isinstance(a, Expr) else repr(a) for a in e.args)})"


# ---- evaluation (host) ---------------------------------------------------------------------
# The lambdas are Python: `//` and `%` keep Python's meaning on every backend (floored division, a
# remainder with the divisor's sign), so a traced query and the same query run as opaque Python agree
# on negative operands.  (The reference's C# truncates; its users wrote C#.)
def _trunc_div(a, b):
    if isinstance(a, float) or isinstance(b, float):
        return a / b
    if b == 0:
        raise ZeroDivisionError("integer division by zero")
    return a // b


def _trunc_mod(a, b):
    if isinstance(a, float) or isinstance(b, float):
        return a % b if b != 0 else math.nan
    if b == 0:
        raise ZeroDivisionError("integer modulo by zero")
    return a % b


_PY = {
    "add": operator.add, "sub": operator.sub, "mul": operator.mul, "div": _trunc_div, "mod": _trunc_mod,
    "eq": operator.eq, "ne": operator.ne, "lt": operator.lt, "le": operator.le, "gt": operator.gt,
    "ge": operator.ge, "band": operator.and_, "bor": operator.or_, "bxor": operator.xor,
    "shl": operator.lshift, "shr": operator.rshift, "min": min, "max": max, "concat": operator.add,
    "truediv_o": operator.truediv,
}


def eval_row(e: Expr, row: Mapping[str, Any], index: int = 0) -> Any:
    """Scalar evaluation; integer division and remainder are Python's (floored)."""
    op = e.op
    if op == "col":
        return row[e.args[0]]
    if op == "lit":
        return e.args[0]
    if op == "rowid":
        return index
    if op in _PY:
        return _PY[op](eval_row(e.args[0], row, index), eval_row(e.args[1], row, index))
    if op == "and":
        return bool(eval_row(e.args[0], row, index)) and bool(eval_row(e.args[1], row, index))
    if op == "or":
        return bool(eval_row(e.args[0], row, index)) or bool(eval_row(e.args[1], row, index))
    if op == "not":
        return not eval_row(e.args[0], row, index)
    if op == "bnot":
        return ~eval_row(e.args[0], row, index)
    if op == "neg":
        return -eval_row(e.args[0], row, index)
    if op == "abs":
        return abs(eval_row(e.args[0], row, index))
    if op == "cast":
        v = eval_row(e.args[0], row, index)
        return {"i": int, "f": float, "b": bool, "o": lambda x: x}[e.kind](v)
    if op == "select":
        return eval_row(e.args[1], row, index) if eval_row(e.args[0], row, index) else eval_row(e.args[2], row, index)
    if op == "hash":
        from ..utils.hashing import key_bits, mix64
        h = mix64(key_bits(eval_row(e.args[0], row, index)))
        return h - (1 << 64) if h >= (1 << 63) else h
    if op == "sqrt":
        return math.sqrt(eval_row(e.args[0], row, index))
    if op == "floor":
        return float(math.floor(eval_row(e.args[0], row, index)))
    raise TraceError(f"cannot evaluate op {op}")


def eval_numpy(e: Expr, cols: Mapping[str, np.ndarray], n: int) -> np.ndarray:
    """Vectorised host evaluation (used by the CPU execution mode for traced stages)."""
    op = e.op
    if op == "col":
        return np.asarray(cols[e.args[0]])
    if op == "lit":
        dt = {"i": np.int64, "f": np.float64, "b": np.bool_}.get(e.kind, object)
        return np.full(n, e.args[0], dtype=dt)
    if op == "rowid":
        return np.arange(n, dtype=np.int64)
    A = [eval_numpy(a, cols, n) for a in e.args if isinstance(a, Expr)]
    if op == "add": return A[0] + A[1]
    if op == "sub": return A[0] - A[1]
    if op == "mul": return A[0] * A[1]
    if op == "div":
        if e.kind == "f":
            return A[0] / A[1]
        return np.where(A[1] == 0, 0, np.floor_divide(A[0], np.where(A[1] == 0, 1, A[1])))
    if op == "mod":
        if e.kind == "f":
            return np.mod(A[0], A[1])
        return np.where(A[1] == 0, 0, np.mod(A[0], np.where(A[1] == 0, 1, A[1])))
    if op == "eq": return A[0] == A[1]
    if op == "ne": return A[0] != A[1]
    if op == "lt": return A[0] < A[1]
    if op == "le": return A[0] <= A[1]
    if op == "gt": return A[0] > A[1]
    if op == "ge": return A[0] >= A[1]
    if op == "and": return A[0].astype(bool) & A[1].astype(bool)
    if op == "or": return A[0].astype(bool) | A[1].astype(bool)
    if op == "not": return ~A[0].astype(bool)
    if op == "band": return A[0] & A[1]
    if op == "bor": return A[0] | A[1]
    if op == "bxor": return A[0] ^ A[1]
    if op == "bnot": return ~A[0]
    if op == "shl": return A[0] << A[1]
    if op == "shr": return A[0] >> A[1]
    if op == "neg": return -A[0]
    if op == "abs": return np.abs(A[0])
    if op == "min": return np.minimum(A[0], A[1])
    if op == "max": return np.maximum(A[0], A[1])
    if op == "cast":
        return A[0].astype({"i": np.int64, "f": np.float64, "b": np.bool_}[e.kind])
    if op == "select": return np.where(A[0].astype(bool), A[1], A[2])
    if op == "sqrt": return np.sqrt(A[0])
    if op == "floor": return np.floor(A[0])
    if op == "hash":
        from ..utils.hashing import mix64_np
        x = A[0]
        bits = x.view(np.uint64) if x.dtype in (np.int64, np.float64) else x.astype(np.int64).view(np.uint64)
        return mix64_np(bits).view(np.int64)
    raise TraceError(f"cannot vectorise op {op}")


# ---- tracing -------------------------------------------------------------------------------
class RowProxy:
    """Symbolic record handed to user lambdas: attribute or item access yields column Exprs."""

    def __init__(self, schema: Mapping[str, str], prefix: str = ""):
        object.__setattr__(self, "_schema", dict(schema))
        object.__setattr__(self, "_prefix", prefix)

    def __getattr__(self, name):
        schema = object.__getattribute__(self, "_schema")
        if name in schema:
            return col(object.__getattribute__(self, "_prefix") + name, schema[name])
        raise TraceError(f"record has no field {name!r}")

    def __getitem__(self, name):
        if isinstance(name, int):
            schema = object.__getattribute__(self, "_schema")
            names = list(schema)
            return self.__getattr__(names[name])
        return self.__getattr__(name)


def trace(fn: Callable, schema: Mapping[str, str], *extra_args) -> Optional[Any]:
    """Try to trace `fn(row, *extra)`.  Returns an Expr, a dict of Exprs (record construction),
    or None when the lambda is opaque.  A scalar schema is {'value': kind}: the lambda then
    receives the column Expr itself rather than a record."""
    try:
        if list(schema) == ["value"]:
            sym = col("value", schema["value"])
        else:
            sym = RowProxy(schema)
        out = fn(sym, *extra_args)
    except TraceError:
        return None
    except Exception:
        return None
    return normalise_traced(out)


class TupleFields(dict):
    """A traced tuple result: field dict that materialises back into a Python tuple — or into the
    tuple SUBCLASS the lambda built (`Pair(k, v)`, `ForkTuple(a, b)`), kept in `cls`."""

    cls = tuple

    def build(self, values):
        values = tuple(values)
        return values if self.cls is tuple else self.cls(*values)


def normalise_traced(out: Any) -> Optional[Any]:
    if isinstance(out, Expr):
        return out
    if isinstance(out, RowProxy):
        schema = object.__getattribute__(out, "_schema")
        return {k: col(k, v) for k, v in schema.items()}
    if isinstance(out, dict):
        res = {}
        for k, v in out.items():
            if isinstance(v, RowProxy):
                return None
            res[k] = wrap(v)
        return res
    if isinstance(out, tuple) and out and all(isinstance(v, (Expr, int, float, bool)) for v in out):
        tf = TupleFields((f"item{i + 1}", wrap(v)) for i, v in enumerate(out))
        if type(out) is not tuple:
            tf.cls = type(out)
        return tf
    if isinstance(out, (int, float, bool)):
        return lit(out)
    return None
