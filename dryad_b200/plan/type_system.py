"""Reflection helpers of the planner — the role of `TypeSystem` / `IdentityFunction`
(LinqToDryad/TypeSystem.cs:37-1399): which callables are query operators, what the default
comparer of a key type is, whether a value is an anonymous type, the identity function test, and
which non-system modules a job drags along as resources (the reference ships every loaded
non-system assembly with the job, DryadLinqJobExecutor / TypeSystem.GetLoadedNonSystemAssemblies)."""
from __future__ import annotations

import sys
import sysconfig
from typing import Any, Callable, Dict, List, Optional

QUERY_OPERATORS = frozenset("""select where select_many group_by order_by order_by_descending then_by
then_by_descending distinct reverse join group_join concat union intersect except_ zip take skip take_while
skip_while hash_partition range_partition apply apply_per_partition fork sliding_window assume_hash_partition
assume_range_partition assume_order_by select_indexed where_indexed count sum min max average any all contains
first first_or_default last last_or_default single single_or_default aggregate sequence_equal to_store
do_while broadcast default_if_empty element_at of_type cast long_count""".split())


def is_query_operator(name: str) -> bool:
    """snake_case or the reference's PascalCase operator name."""
    import re
    snake = re.sub(r"(?<!^)(?=[A-Z])", "_", name).lower()
    return snake in QUERY_OPERATORS or snake.rstrip("_") in QUERY_OPERATORS


def identity(x):
    return x


def is_identity_function(fn: Optional[Callable]) -> bool:
    """True for None, the canonical identity, and lambdas whose traced form is their argument."""
    if fn is None or fn is identity:
        return True
    try:
        from ..expr import ir as X
        out = X.trace(fn, {"value": "i"})
        return isinstance(out, X.Expr) and out.op == "col" and out.args[0] == "value"
    except Exception:  # noqa: BLE001
        return False


def is_identity_ir(ir: Any, schema: Optional[Dict[str, str]]) -> bool:
    """True when a TRACED selector returns its argument unchanged: the `value` column for scalar
    datasets, or a record whose every field is the same-named input column."""
    from ..expr import ir as X
    if schema is None:
        return False
    if isinstance(ir, X.Expr):
        return list(schema) == ["value"] and ir.op == "col" and ir.args[0] == "value"
    if isinstance(ir, dict) and not isinstance(ir, X.TupleFields):
        return list(ir) == list(schema) and all(isinstance(e, X.Expr) and e.op == "col" and e.args[0] == k for k, e in ir.items())
    return False


def is_anonymous_type(v: Any) -> bool:
    """`new { a, b }` in the reference = a plain dict / Record here."""
    return isinstance(v, dict)


def kind_of_value(v: Any) -> str:
    """Column kind of a Python value: 'b', 'i', 'f' or 'o' (opaque: host path only)."""
    if isinstance(v, bool):
        return "b"
    if isinstance(v, int):
        return "i" if -(1 << 63) <= v < (1 << 63) else "o"
    if isinstance(v, float):
        return "f"
    return "o"


def default_comparer(kind: str) -> Callable[[Any, Any], int]:
    """Comparer<T>.Default: natural order; strings compare ordinally by code point."""
    def cmp(a, b):
        return (a > b) - (a < b)
    return cmp


def default_equality(kind: str) -> Callable[[Any, Any], bool]:
    return lambda a, b: a == b


def loaded_non_system_modules() -> List[str]:
    """Files of the imported modules that live outside the interpreter's own library trees — what
    a job submission has to ship to remote workers (`resources_to_add` default set)."""
    roots = {sysconfig.get_paths().get(k) for k in ("stdlib", "platstdlib", "purelib", "platlib")}
    roots = tuple(r for r in roots if r)
    out = []
    for m in list(sys.modules.values()):
        f = getattr(m, "__file__", None)
        if f and not f.startswith(roots) and f.endswith((".py", ".so")):
            out.append(f)
    return sorted(set(out))


def schema_of_record(rec: Any) -> Optional[Dict[str, str]]:
    if isinstance(rec, dict):
        return {k: kind_of_value(v) for k, v in rec.items()}
    if isinstance(rec, tuple):
        return {f"item{i + 1}": kind_of_value(v) for i, v in enumerate(rec)}
    k = kind_of_value(rec)
    return {"value": k}
