"""Pre-planning algebraic rewrites (reference: LinqToDryad/SimpleRewriter.cs:37-327, rules
:246-324): push Where below OrderBy / Distinct / HashPartition / RangePartition and through
Concat; push Select / SelectMany through Concat; swap Take below Select; drop partitioning and
sorting under Contains / Any / All / Count (order does not affect them).

Rewrites build NEW QExpr nodes and never mutate the user's tree; shared sub-queries (a node with
several consumers) are left where they are so the Tee the planner inserts still sees one producer.
"""
from __future__ import annotations

from typing import Dict

from .logical import QExpr

_ORDER_INSENSITIVE_AGGS = {"count", "sum", "min", "max", "average", "any", "all", "contains"}
_REORDERING = {"order_by", "hash_partition", "range_partition"}


def _use_counts(root: QExpr) -> Dict[int, int]:
    counts: Dict[int, int] = {}
    seen = set()

    def rec(e: QExpr):
        for s in e.sources:
            counts[s.id] = counts.get(s.id, 0) + 1
            if s.id not in seen:
                seen.add(s.id)
                rec(s)
    rec(root)
    return counts


def _clone(e: QExpr, sources) -> QExpr:
    return QExpr(e.op, sources, e.schema, **e.params)


def simple_rewrite(root: QExpr) -> QExpr:
    uses = _use_counts(root)
    memo: Dict[int, QExpr] = {}

    def single(e: QExpr) -> bool:
        return uses.get(e.id, 0) <= 1

    def rw(e: QExpr) -> QExpr:
        if e.id in memo:
            return memo[e.id]
        srcs = [rw(s) for s in e.sources]
        out = e if all(a is b for a, b in zip(srcs, e.sources)) else _clone(e, srcs)
        s0 = out.sources[0] if out.sources else None
        if out.op == "select" and not out.params.get("indexed") and s0 is not None:
            # Select(x => x) does nothing: the planner never sees it (IdentityFunction, TypeSystem.cs)
            from .type_system import is_identity_ir
            lam = out.params.get("fn")
            if lam is not None and is_identity_ir(getattr(lam, "ir", None), s0.schema):
                memo[e.id] = s0
                return s0
        if out.op == "where" and not out.params.get("indexed") and s0 is not None and single(e.sources[0]):
            # Where below OrderBy / Distinct / partition operators
            if s0.op in ("order_by", "distinct", "hash_partition", "range_partition") and \
                    not (s0.op == "range_partition" and s0.params.get("range_keys") is None and False):
                inner = QExpr("where", [s0.sources[0]], s0.sources[0].schema, **out.params)
                out = _clone(s0, [rw_once(inner)] + s0.sources[1:])
            elif s0.op == "concat":
                out = QExpr("concat", [rw_once(QExpr("where", [c], c.schema, **out.params)) for c in s0.sources],
                            s0.schema)
        elif out.op in ("select", "select_many") and not out.params.get("indexed") and s0 is not None \
                and s0.op == "concat" and single(e.sources[0]):
            out = QExpr("concat", [QExpr(out.op, [c], out.schema, **out.params) for c in s0.sources], out.schema)
        elif out.op == "take" and s0 is not None and s0.op == "select" and not s0.params.get("indexed") \
                and single(e.sources[0]):
            inner = QExpr("take", [s0.sources[0]], s0.sources[0].schema, **out.params)
            out = _clone(s0, [inner])
        elif out.op == "aggregate" and out.params.get("kind") in _ORDER_INSENSITIVE_AGGS and s0 is not None:
            cur = s0
            while cur.op in _REORDERING and single(cur):
                cur = cur.sources[0]
            if cur is not s0:
                out = _clone(out, [cur] + out.sources[1:])
        memo[e.id] = out
        return out

    def rw_once(e: QExpr) -> QExpr:
        # a freshly created Where may itself sit above another pushable operator
        uses.setdefault(e.id, 1)
        for s in e.sources:
            uses[s.id] = uses.get(s.id, 1)
        return rw(e)

    return rw(root)
