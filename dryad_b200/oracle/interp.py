"""LocalDebug execution: interpret a logical query tree with LINQ-to-Objects semantics.

Reference: DryadLinqLocalProvider / DryadLinqLocalQuery (LinqToDryad/DryadLinqQuery.cs:55-134) and
the `IsLocalDebugSource` branch every operator has (LinqToDryad/DryadLinqQueryable.cs:41).  Results
are memoised per QExpr so shared sub-queries run once, and `to_store` outputs really are written.
"""
from __future__ import annotations

from typing import Any, Dict, List

from ..plan.logical import QExpr
from . import linq as L


def _group_final(dec, groups):
    out = []
    for g in groups:
        accs = None
        for x in g:
            accs = [a.seed(x) for a in dec.aggs] if accs is None else [a.accumulate(v, x) for a, v in zip(dec.aggs, accs)]
        out.append(dec.final(g.key, accs))
    return out


class LocalDebugInterpreter:
    def __init__(self, context):
        self.context = context
        self.memo: Dict[int, Any] = {}

    def run(self, e: QExpr) -> List:
        if e.id in self.memo:
            return self.memo[e.id]
        r = self._eval(e)
        self.memo[e.id] = r
        return r

    def _eval(self, e: QExpr):
        op, p = e.op, e.params
        S = [self.run(s) for s in e.sources] if op != "fork_branch" else None
        if op == "source":
            return self.context._read_source_flat(e)
        if op == "select":
            return L.select(S[0], p["fn"].fn, p["indexed"])
        if op == "where":
            return L.where(S[0], p["fn"].fn, p["indexed"])
        if op == "select_many":
            return L.select_many(S[0], p["fn"].fn, p["result"].fn if p.get("result") else None, p["indexed"])
        if op == "group_by":
            elem = p["elem"].fn if p.get("elem") else None
            groups = L.group_by(S[0], p["key"].fn, elem)
            if "decomp" in p:
                return _group_final(p["decomp"], groups)
            if "result" in p:
                return [p["result"].fn(g.key, g) for g in groups]
            return groups
        if op == "order_by":
            return L.order_by(S[0], p["key"].fn, p["descending"], p.get("comparer"))
        if op == "distinct":
            return L.distinct(S[0])
        if op == "reverse":
            return L.reverse(S[0])
        if op == "join":
            return L.join(S[0], S[1], p["outer_key"].fn, p["inner_key"].fn, p["result"].fn)
        if op == "group_join":
            return L.group_join(S[0], S[1], p["outer_key"].fn, p["inner_key"].fn, p["result"].fn)
        if op == "concat":
            return L.concat(S[0], S[1])
        if op == "union":
            return L.union(S[0], S[1])
        if op == "intersect":
            return L.intersect(S[0], S[1])
        if op == "except":
            return L.except_(S[0], S[1])
        if op == "zip":
            return L.zip_(S[0], S[1], p["fn"].fn)
        if op == "take":
            return L.take(S[0], p["n"])
        if op == "skip":
            return L.skip(S[0], p["n"])
        if op == "take_while":
            return L.take_while(S[0], p["fn"].fn, p["indexed"])
        if op == "skip_while":
            return L.skip_while(S[0], p["fn"].fn, p["indexed"])
        if op in ("hash_partition", "range_partition", "assume_hash_partition", "assume_range_partition",
                  "assume_order_by", "broadcast"):
            return list(S[0])      # identity in LocalDebug (DryadLinqEnumerable.cs:41-168)
        if op == "apply":
            if p.get("with_index"):
                return list(p["fn"].fn(list(S[0]), 0))
            if len(S) == 2:
                return L.apply2(S[0], S[1], p["fn"].fn)
            return L.apply(S[0], p["fn"].fn)
        if op == "sliding_window":
            return L.sliding_window(S[0], p["fn"].fn, p["window"])
        if op == "fork":
            if p["per_partition"]:
                outs = p["fn"].fn(list(S[0]))
                return tuple(list(o) for o in outs)
            return L.fork(S[0], p["fn"].fn, p["n"])
        if op == "fork_branch":
            return list(self.run(e.sources[0])[p["index"]])
        if op == "aggregate":
            return [self._aggregate(S, p)]
        if op == "to_store":
            data = S[0]
            if p.get("uri"):
                self.context._write_store(p["uri"], [list(data)], e.sources[0].schema, p.get("compression"))
            return data
        raise NotImplementedError(f"LocalDebug: operator {op}")

    def _aggregate(self, S, p):
        from ..api.errors import DryadLinqException, ErrorCode
        k = p["kind"]
        src = S[0]
        fn = p["fn"].fn if p.get("fn") else None
        try:
            if k == "count": return L.count(src, fn)
            if k == "sum": return L.sum_(src, fn)
            if k == "min": return L.min_(src, fn)
            if k == "max": return L.max_(src, fn)
            if k == "average": return L.average(src, fn)
            if k == "any": return L.any_(src, fn)
            if k == "all": return L.all_(src, fn)
            if k == "contains": return L.contains(src, p["value"])
            if k == "first": return L.first(src, fn, p["default"])
            if k == "last": return L.last(src, fn, p["default"])
            if k == "single": return L.single(src, fn, p["default"])
            if k == "sequence_equal": return L.sequence_equal(S[0], S[1])
            if k == "aggregate":
                seed = p["seed_fn"]() if p.get("seed_fn") else p.get("seed")
                return L.aggregate(src, fn, seed, p["has_seed"], p["result"].fn if p.get("result") else None)
        except ValueError as ex:
            code = {"min": ErrorCode.MinNoElements, "max": ErrorCode.MaxNoElements,
                    "average": ErrorCode.AverageNoElements, "first": ErrorCode.FirstNoElementsFirst,
                    "last": ErrorCode.LastNoElements, "aggregate": ErrorCode.AggregateNoElements,
                    "single": ErrorCode.SingleNoElements}.get(k, ErrorCode.FailureInOperator)
            if k == "single" and "more than one" in str(ex):
                code = ErrorCode.SingleMoreThanOneElement
            raise DryadLinqException(code, str(ex)) from ex
        raise NotImplementedError(k)
