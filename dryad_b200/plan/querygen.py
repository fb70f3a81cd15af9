"""The planner: logical query trees -> physical stage DAG.

Reference: LinqToDryad/DryadLinqQueryGen.cs.  Same three phases:
  Phase 1 (:269-388)  rewrite (SimpleRewriter) then one Visit* per operator creates physical nodes,
                      inserting shuffles / merges / sorts only where the propagated DataSetInfo
                      says the data is not already partitioned / ordered as needed;
  Phase 2 (:391-456)  chains of pipelinable operators are fused into Super nodes; forked outputs
                      get a Tee;
  Phase 3 (:459-522)  useless Tee / Merge nodes are removed, broadcast managers attached to Tees;
then AssignUniqueId (:634-690, children first) and the XML plan (:814-974) with the element
vocabulary QueryPlanParser reads (DryadLinqGraphManager/QueryParser.cs:129-347).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple
from xml.sax.saxutils import escape

from ..api.errors import DryadLinqException, ErrorCode
from ..expr import ir as X
from . import static_config as SC
from .dataset_info import (DataSetInfo, DistinctInfo, HashPartition, KeySelector, OrderByInfo, RandomPartition,
                           RangePartition, compute_is_descending)
from .logical import Lam, QExpr
from .qnodes import (XML_VERTEX_TYPE, ChannelType, ConnOp, DLinqQueryNode, DynamicManagerType, NodeType,
                     make_super)
from .rewriter import simple_rewrite

PW, CP = ConnOp.Pointwise, ConnOp.CrossProduct


def _ks(lam: Lam) -> KeySelector:
    return KeySelector(lam.fn, lam.ir if isinstance(lam.ir, X.Expr) else None, lam.name)


class QueryGen:
    def __init__(self, context, outputs: Sequence):
        """`outputs`: Query objects whose root is a to_store expression."""
        self.context = context
        self.outputs = [q.expr for q in outputs]
        for e in self.outputs:
            if e.op != "to_store":
                raise DryadLinqException(ErrorCode.Internal, "plan roots must be outputs")
        self._memo: Dict[int, DLinqQueryNode] = {}
        # DynamicOptLevel (DryadLinqGlobals.cs:40-52): which stages get a dynamic-manager tag
        self._dyn_level = int(getattr(context, "dynamic_opt_level", SC.DYNAMIC_OPT_LEVEL))
        self.roots: List[DLinqQueryNode] = []
        self.nodes: List[DLinqQueryNode] = []
        self._built = False

    def _spmd_width(self, count):
        """One process per GPU: a stage has one partition per rank (or a single one), so an explicit partition
        count of a repartitioning is a hint there — any other value becomes the number of ranks.  (Explicit range
        KEYS define their own count and are left alone.)"""
        world = int(getattr(self.context, "world", 1) or 1)
        if count and world > 1 and count not in (1, world):
            return world
        return count

    # ---- driver ------------------------------------------------------------------------------
    def build(self) -> "QueryGen":
        if self._built:
            return self
        self._check_output_rules()
        exprs = [simple_rewrite(e) for e in self.outputs]
        self.roots = [self.visit(e) for e in exprs]          # phase 1
        for node, orig in zip(self.roots, self.outputs):     # results are keyed by the user's expression
            node.params["qexpr_id"] = orig.id
        self._phase2()
        self._phase3()
        self._assign_unique_ids()
        self._built = True
        return self

    def _check_output_rules(self):
        uris, inputs = set(), set()
        for e in self.outputs:
            for n in e.walk():
                if n.op == "source" and n.params.get("uri"):
                    inputs.add(n.params["uri"])
        for e in self.outputs:
            u = e.params.get("uri")
            if u is None:
                continue
            if u in uris:
                raise DryadLinqException(ErrorCode.MultipleOutputsWithSameDscUri, f"two outputs share {u}")
            if u in inputs:
                raise DryadLinqException(ErrorCode.OutputUriAlsoQueryInput, f"{u} is also a query input")
            uris.add(u)
        # every fork branch must be consumed (DryadLinqQueryGen.cs:411-424)
        forks: Dict[int, set] = {}
        n_of: Dict[int, int] = {}
        for e in self.outputs:
            for n in e.walk():
                if n.op == "fork":
                    n_of[n.id] = n.params["n"]
                    forks.setdefault(n.id, set())
                if n.op == "fork_branch":
                    forks.setdefault(n.sources[0].id, set()).add(n.params["index"])
        for fid, used in forks.items():
            if fid in n_of and len(used) < n_of[fid]:
                raise DryadLinqException(ErrorCode.BranchOfForkNotUsed, "every branch of a Fork must be consumed")

    # ---- phase 1: visitors -------------------------------------------------------------------
    def visit(self, e: QExpr) -> DLinqQueryNode:
        n = self._memo.get(e.id)
        if n is None:
            fn = getattr(self, "_visit_" + e.op, None)
            if fn is None:
                raise DryadLinqException(ErrorCode.OperatorNotSupported, f"operator {e.op} is not supported")
            n = fn(e)
            self._memo[e.id] = n
        return n

    def _node(self, t: NodeType, children, pc, info=None, schema=None, name="", **p) -> DLinqQueryNode:
        return DLinqQueryNode(t, children, pc, info, schema, name, **p)

    def _merge(self, child: DLinqQueryNode, pc: int, conn: ConnOp, info: Optional[DataSetInfo] = None,
               sort_key: Optional[Lam] = None, descending=False, comparer=None) -> DLinqQueryNode:
        return self._node(NodeType.Merge, [(child, conn)], pc, info or DataSetInfo(RandomPartition(pc)),
                          child.schema, keep_port_order=True, sort_key=sort_key, descending=descending,
                          comparer=comparer)

    def _to_one(self, child: DLinqQueryNode, ordered_by: Optional[OrderByInfo] = None) -> DLinqQueryNode:
        if child.partition_count == 1:
            return child
        info = DataSetInfo(RandomPartition(1), ordered_by or OrderByInfo())
        return self._merge(child, 1, PW, info)

    def _hash_shuffle(self, child: DLinqQueryNode, key: Lam, pc: int, comparer=None) -> DLinqQueryNode:
        hp = self._node(NodeType.HashPartition, [(child, PW)], child.partition_count, None, child.schema,
                        key=key, nports=pc, comparer=comparer)
        hp.n_out_ports = pc
        if self._dyn_level & SC.DYN_HASH_PARTITION:
            hp.dynamic_manager = DynamicManagerType.HashDistributor
        info = DataSetInfo(HashPartition(_ks(key), pc, comparer))
        return self._merge(hp, pc, CP, info)

    # sources / sinks
    def _visit_source(self, e: QExpr):
        pc = max(1, int(e.params["partition_count"]))
        info = e.params.get("info") or DataSetInfo(RandomPartition(pc))   # stored tables load as Random
        return self._node(NodeType.InputTable, [], pc, info, e.schema, table=e.params["table"],
                          uri=e.params.get("uri"))

    def _visit_to_store(self, e: QExpr):
        child = self.visit(e.sources[0])
        if child.type == NodeType.InputTable:
            # input -> output needs a vertex in between (dummy Apply, DryadLinqQueryGen.cs:330-345)
            child = self._node(NodeType.Dummy, [(child, PW)], child.partition_count, child.info, child.schema)
        return self._node(NodeType.OutputTable, [(child, PW)], child.partition_count, child.info, child.schema,
                          uri=e.params.get("uri"), temporary=e.params.get("temporary", False),
                          compression=e.params.get("compression"), qexpr_id=e.id)

    # element-wise
    def _offsets_side_input(self, child: DLinqQueryNode) -> DLinqQueryNode:
        """CreateOffset (DryadLinqQueryGen.cs:1225-1248): per-partition count -> one vertex ->
        exclusive scan -> one start offset back to every partition."""
        pc = child.partition_count
        cnt = self._node(NodeType.BasicAggregate, [(child, PW)], pc, None, {"value": "i"}, kind="count", stage=1, fn=None)
        one = self._to_one(cnt)
        off = self._node(NodeType.Offsets, [(one, PW)], 1, None, {"value": "i"}, nports=pc)
        off.n_out_ports = pc
        return self._merge(off, pc, CP)

    def _visit_select(self, e: QExpr):
        child = self.visit(e.sources[0])
        lam: Lam = e.params["fn"]
        if e.params["indexed"]:
            side = self._offsets_side_input(child)
            return self._node(NodeType.Select, [(child, PW), (side, PW)], child.partition_count,
                              DataSetInfo(RandomPartition(child.partition_count)), e.schema, fn=lam, indexed=True)
        proj = lam.ir if isinstance(lam.ir, dict) else ({"value": lam.ir} if isinstance(lam.ir, X.Expr) else None)
        info = child.info.rewrite(proj) if proj is not None else DataSetInfo(RandomPartition(child.partition_count))
        if proj is None and self.context.select_order_preserving:
            info = DataSetInfo(RandomPartition(child.partition_count))
        return self._node(NodeType.Select, [(child, PW)], child.partition_count, info, e.schema, fn=lam, indexed=False)

    def _visit_where(self, e: QExpr):
        child = self.visit(e.sources[0])
        if e.params["indexed"]:
            side = self._offsets_side_input(child)
            return self._node(NodeType.Where, [(child, PW), (side, PW)], child.partition_count, child.info,
                              e.schema, fn=e.params["fn"], indexed=True)
        return self._node(NodeType.Where, [(child, PW)], child.partition_count, child.info, e.schema,
                          fn=e.params["fn"], indexed=False)

    def _visit_select_many(self, e: QExpr):
        child = self.visit(e.sources[0])
        kids = [(child, PW)]
        if e.params["indexed"]:
            kids.append((self._offsets_side_input(child), PW))
        return self._node(NodeType.SelectMany, kids, child.partition_count,
                          DataSetInfo(RandomPartition(child.partition_count)), e.schema, fn=e.params["fn"],
                          result=e.params.get("result"), indexed=e.params["indexed"])

    # grouping (VisitGroupBy, DryadLinqQueryGen.cs:2007-2357)
    def _visit_group_by(self, e: QExpr):
        child = self.visit(e.sources[0])
        p = e.params
        key: Lam = p["key"]
        ks = _ks(key)
        comparer = p.get("comparer")
        pc = child.partition_count
        dec = p.get("decomp")
        partitioned = child.info.partition.is_partitioned_by(ks, comparer) or pc == 1
        out_info = DataSetInfo(HashPartition(KeySelector(None, None, "key"), pc, comparer))
        if dec is not None:
            if not partitioned and not SC.GROUP_BY_LOCAL_AGGREGATION_IS_PARTIAL:
                # no local pre-aggregation: shuffle the rows by key, then one full aggregation
                child = self._hash_shuffle(child, key, pc, comparer)
                partitioned = True
            if partitioned:
                g = self._node(NodeType.GroupBy, [(child, PW)], pc, child.info.keep_partition_only(), None,
                               key=key, elem=p.get("elem"), decomp=dec, mode="full_acc", comparer=comparer)
            else:
                part = self._node(NodeType.GroupBy, [(child, PW)], pc, None, None, key=key, elem=p.get("elem"),
                                  decomp=dec, mode="partial_acc", comparer=comparer)
                part.dynamic_manager = DynamicManagerType.PartialAggregator \
                    if ((self._dyn_level & SC.DYN_AGGREGATE) or SC.GROUP_BY_DYNAMIC_REDUCE) else DynamicManagerType.NoManager
                kir = X.col("key", key.ir.kind) if isinstance(key.ir, X.Expr) and key.ir.kind in "ifbs" else None
                pkey = Lam(lambda kv: kv[0], kir, 1, "Key")
                merged = self._hash_shuffle(part, pkey, pc, comparer)
                g = self._node(NodeType.GroupBy, [(merged, PW)], pc, out_info, None, key=pkey, decomp=dec,
                               mode="final_acc", comparer=comparer)
            return self._node(NodeType.Select, [(g, PW)], pc, g.info, e.schema, fn=Lam(None, None, 1, "FinalReduce"),
                              finalize=dec, indexed=False, name="Select_FinalReduce")
        if not partitioned:
            child = self._hash_shuffle(child, key, pc, comparer)
        algo = "hash"
        if child.info.order.is_ordered_by(ks, comparer):
            algo = "ordered"            # OrderedGroupBy on already-sorted input
        return self._node(NodeType.GroupBy, [(child, PW)], pc, child.info.keep_partition_only(), e.schema, key=key,
                          elem=p.get("elem"), result=p.get("result"), mode="full", algorithm=algo, comparer=comparer)

    # ordering (CreateRangePartition :2362-2474, VisitOrderBy :2476-2521)
    def _range_shuffle(self, child: DLinqQueryNode, key: Lam, pc: int, descending: Optional[bool],
                       comparer=None, range_keys: Optional[list] = None) -> DLinqQueryNode:
        ks = _ks(key)
        if range_keys is not None:
            if len(range_keys) > pc - 1 and pc > 0:
                pass
            is_desc = compute_is_descending(range_keys, comparer)
            if descending is not None and len(range_keys) > 1 and is_desc != descending:
                raise DryadLinqException(ErrorCode.IsDescendingIsInconsistent,
                                         "isDescending disagrees with the order of the range keys")
            desc = descending if descending is not None else bool(is_desc)
            rp = self._node(NodeType.RangePartition, [(child, PW)], child.partition_count, None, child.schema,
                            key=key, nports=pc, descending=desc, comparer=comparer, range_keys=list(range_keys))
            rp.n_out_ports = pc
            info = DataSetInfo(RangePartition(ks, pc, list(range_keys), desc, comparer))
            return self._merge(rp, pc, CP, info)
        desc = bool(descending)
        # branch A: sample -> one vertex -> separators (broadcast to every distributor)
        # sampled keys and separators are a `value` column of the key's kind when the key is traced
        kschema = {"value": key.ir.kind} if isinstance(key.ir, X.Expr) and key.ir.kind in "if" and comparer is None else None
        smp = self._node(NodeType.Sampler, [(child, PW)], child.partition_count, None, kschema, key=key)
        one = self._to_one(smp)
        rs = self._node(NodeType.RangeSampler, [(one, PW)], 1, None, kschema, pcount=pc, descending=desc, comparer=comparer)
        # branch B: distribute with the separators as a side input
        rp = self._node(NodeType.RangePartition, [(child, PW), (rs, PW)], child.partition_count, None, child.schema,
                        key=key, nports=pc, descending=desc, comparer=comparer, range_keys=None)
        rp.n_out_ports = pc
        if self._dyn_level & SC.DYN_RANGE_PARTITION:
            rp.dynamic_manager = DynamicManagerType.RangeDistributor
        info = DataSetInfo(RangePartition(ks, pc, None, desc, comparer))
        return self._merge(rp, pc, CP, info)

    def _visit_order_by(self, e: QExpr):
        child = self.visit(e.sources[0])
        p = e.params
        key, desc, comparer = p["key"], p["descending"], p.get("comparer")
        ks = _ks(key)
        pc = child.partition_count
        part = child.info.partition
        ranged = isinstance(part, RangePartition) and part.is_partitioned_by(ks, comparer) and part.descending == desc
        if not ranged and pc > 1:
            child = self._range_shuffle(child, key, pc, desc, comparer)
        if child.info.order.is_ordered_by(ks, comparer, desc):
            return child                                       # already sorted: nothing to do
        info = DataSetInfo(child.info.partition if pc > 1 else RangePartition(ks, 1, [], desc, comparer),
                           OrderByInfo(ks, desc, comparer))
        return self._node(NodeType.OrderBy, [(child, PW)], pc, info, e.schema, key=key, descending=desc, comparer=comparer)

    def _visit_hash_partition(self, e: QExpr):
        child = self.visit(e.sources[0])
        pc = self._spmd_width(e.params["count"]) or child.partition_count
        return self._hash_shuffle(child, e.params["key"], pc, e.params.get("comparer"))

    def _visit_range_partition(self, e: QExpr):
        child = self.visit(e.sources[0])
        p = e.params
        rk = p.get("range_keys")
        pc = (p["count"] if rk is not None else self._spmd_width(p["count"])) or (len(rk) + 1 if rk is not None else child.partition_count)
        if rk is not None and len(rk) > pc - 1:
            raise DryadLinqException(ErrorCode.RangePartitionInputOutputMismatch,
                                     f"{len(rk)} range keys for {pc} partitions")
        return self._range_shuffle(child, p["key"], pc, p.get("descending"), p.get("comparer"), rk)

    def _visit_assume_hash_partition(self, e: QExpr):
        child = self.visit(e.sources[0])
        info = DataSetInfo(HashPartition(_ks(e.params["key"]), child.partition_count, e.params.get("comparer")),
                           child.info.order, child.info.distinct)
        return self._node(NodeType.Dummy, [(child, PW)], child.partition_count, info, child.schema, name="AssumeHashPartition")

    def _visit_assume_range_partition(self, e: QExpr):
        child = self.visit(e.sources[0])
        p = e.params
        rk = p.get("range_keys")
        if rk is not None and len(rk) != child.partition_count - 1:
            raise DryadLinqException(ErrorCode.BadSeparatorCount,
                                     "the number of separators must equal the number of partitions - 1")
        desc = p.get("descending")
        if desc is None:
            desc = bool(compute_is_descending(rk, p.get("comparer"))) if rk else False
        info = DataSetInfo(RangePartition(_ks(p["key"]), child.partition_count, rk, desc, p.get("comparer")),
                           child.info.order, child.info.distinct)
        return self._node(NodeType.Dummy, [(child, PW)], child.partition_count, info, child.schema, name="AssumeRangePartition")

    def _visit_assume_order_by(self, e: QExpr):
        child = self.visit(e.sources[0])
        p = e.params
        info = DataSetInfo(child.info.partition, OrderByInfo(_ks(p["key"]), p["descending"], p.get("comparer")),
                           child.info.distinct)
        return self._node(NodeType.Dummy, [(child, PW)], child.partition_count, info, child.schema, name="AssumeOrderBy")

    # joins (VisitJoin, DryadLinqQueryGen.cs:1419-1609)
    def _co_partition(self, outer, inner, okey: Lam, ikey: Lam, comparer):
        oks, iks = _ks(okey), _ks(ikey)
        op, ip = outer.info.partition, inner.info.partition
        o_ok = op.is_partitioned_by(oks, comparer)
        i_ok = ip.is_partitioned_by(iks, comparer)
        if o_ok and i_ok and op.is_same_partition(ip):
            return outer, inner
        if o_ok and isinstance(op, HashPartition):
            return outer, self._hash_shuffle(inner, ikey, op.count, comparer)
        if i_ok and isinstance(ip, HashPartition):
            return self._hash_shuffle(outer, okey, ip.count, comparer), inner
        def co_range(side, key, part):
            # the other side is re-partitioned with the SAME separators (co-range); with DYN_CO_RANGE
            # the distributor is tagged so the runtime may re-balance both sides together
            merged = self._range_shuffle(side, key, part.count, part.descending, comparer, list(part.separators))
            if self._dyn_level & SC.DYN_CO_RANGE and merged.children:
                merged.children[0][0].dynamic_manager = DynamicManagerType.RangeDistributor
            return merged
        if o_ok and isinstance(op, RangePartition) and op.has_keys():
            return outer, co_range(inner, ikey, op)
        if i_ok and isinstance(ip, RangePartition) and ip.has_keys():
            return co_range(outer, okey, ip), inner
        pc = max(outer.partition_count, inner.partition_count)
        if pc == 1:
            return outer, inner
        return self._hash_shuffle(outer, okey, pc, comparer), self._hash_shuffle(inner, ikey, pc, comparer)

    def _visit_join(self, e: QExpr, ntype=NodeType.Join):
        outer, inner = self.visit(e.sources[0]), self.visit(e.sources[1])
        p = e.params
        okey, ikey, comparer = p["outer_key"], p["inner_key"], p.get("comparer")
        # broadcast join: planner rule added for the "large fact x small dim" config — the small
        # side is replicated to every partition of the big side instead of shuffling both.
        small = self.context.broadcast_join_rows
        in_rows = e.sources[1].params.get("row_estimate") if e.sources[1].op == "source" else None
        if ntype == NodeType.Join and small and in_rows is not None and in_rows <= small and outer.partition_count > 1:
            bc = self._broadcast(inner, outer.partition_count)
            return self._node(ntype, [(outer, PW), (bc, PW)], outer.partition_count,
                              DataSetInfo(RandomPartition(outer.partition_count)), e.schema, outer_key=okey,
                              inner_key=ikey, result=p["result"], algorithm="hash", broadcast=True, comparer=comparer)
        outer, inner = self._co_partition(outer, inner, okey, ikey, comparer)
        oks, iks = _ks(okey), _ks(ikey)
        o_sorted = outer.info.order.is_ordered_by(oks, comparer)
        i_sorted = inner.info.order.is_ordered_by(iks, comparer)
        algo = "hash"
        if o_sorted or i_sorted:
            algo = "merge"
            desc = outer.info.order.descending if o_sorted else inner.info.order.descending
            if not o_sorted:
                outer = self._node(NodeType.OrderBy, [(outer, PW)], outer.partition_count,
                                   DataSetInfo(outer.info.partition, OrderByInfo(oks, desc, comparer)), outer.schema,
                                   key=okey, descending=desc, comparer=comparer)
            if not i_sorted:
                inner = self._node(NodeType.OrderBy, [(inner, PW)], inner.partition_count,
                                   DataSetInfo(inner.info.partition, OrderByInfo(iks, desc, comparer)), inner.schema,
                                   key=ikey, descending=desc, comparer=comparer)
        pc = outer.partition_count
        return self._node(ntype, [(outer, PW), (inner, PW)], pc, DataSetInfo(RandomPartition(pc)), e.schema,
                          outer_key=okey, inner_key=ikey, result=p["result"], algorithm=algo, broadcast=False,
                          comparer=comparer, descending=(outer.info.order.descending if algo == "merge" else False),
                          gj_decomp=p.get("gj_decomp"))

    def _visit_group_join(self, e: QExpr):
        return self._visit_join(e, NodeType.GroupJoin)

    def _broadcast(self, child: DLinqQueryNode, pc: int) -> DLinqQueryNode:
        """Tee + CrossProduct + Merge(parCount) (DryadLinqQueryGen.cs:2972-2978)."""
        one = self._to_one(child)
        tee = self._node(NodeType.Tee, [(one, PW)], 1, one.info, one.schema, nports=pc, replicate=True)
        tee.n_out_ports = pc
        tee.dynamic_manager = DynamicManagerType.Broadcast if (self._dyn_level & SC.DYN_BROADCAST) else DynamicManagerType.NoManager
        return self._merge(tee, pc, CP)

    def _visit_broadcast(self, e: QExpr):
        child = self.visit(e.sources[0])
        return self._broadcast(child, self._spmd_width(e.params["count"]) or child.partition_count)

    # distinct / sets (VisitDistinct :1611, VisitSetOperation :1722)
    def _visit_distinct(self, e: QExpr):
        child = self.visit(e.sources[0])
        ident = Lam(lambda x: x, X.col("value", e.schema["value"]) if e.schema and list(e.schema) == ["value"] and e.schema["value"] != "o" else None, 1, "identity")
        pc = child.partition_count
        if child.info.distinct.is_distinct:
            return child
        if not (child.info.partition.is_partitioned_by(_ks(ident), e.params.get("comparer")) or pc == 1):
            if SC.USE_PARTIAL_DISTINCT:
                child = self._node(NodeType.Distinct, [(child, PW)], pc, None, child.schema, partial=True)
            child = self._hash_shuffle(child, ident, pc, e.params.get("comparer"))
        info = DataSetInfo(child.info.partition, OrderByInfo(), DistinctInfo(True, e.params.get("comparer")))
        return self._node(NodeType.Distinct, [(child, PW)], pc, info, e.schema, partial=False)

    def _set_op(self, e: QExpr, op: str):
        a, b = self.visit(e.sources[0]), self.visit(e.sources[1])
        sch = e.sources[0].schema
        ident = Lam(lambda x: x, X.col("value", sch["value"]) if sch and list(sch) == ["value"] and sch["value"] != "o" else None, 1, "identity")
        a, b = self._co_partition(a, b, ident, ident, e.params.get("comparer"))
        pc = a.partition_count
        info = DataSetInfo(a.info.partition, OrderByInfo(), DistinctInfo(True))
        return self._node(NodeType.SetOperation, [(a, PW), (b, PW)], pc, info, e.schema, op=op)

    def _visit_union(self, e): return self._set_op(e, "union")
    def _visit_intersect(self, e): return self._set_op(e, "intersect")
    def _visit_except(self, e): return self._set_op(e, "except")

    def _visit_concat(self, e: QExpr):
        a, b = self.visit(e.sources[0]), self.visit(e.sources[1])
        kids = []
        for c in (a, b):      # flatten nested Concat (PromoteConcat :1250)
            if c.type == NodeType.Concat and not c.is_forked:
                kids.extend(c.children)
                for cc, _ in c.children:
                    cc.parents.remove(c)
            else:
                kids.append((c, PW))
        pc = sum(c.partition_count for c, _ in kids)
        n = self._node(NodeType.Concat, kids, pc, DataSetInfo(RandomPartition(pc)), e.schema)
        if pc > SC.MAX_PARTITION_COUNT:      # halve by indexed repartitioning (:1652-1720)
            n = self._node(NodeType.Merge, [(n, PW)], (pc + 1) // 2, None, e.schema, keep_port_order=True, group=2)
        return n

    def _visit_zip(self, e: QExpr):
        a, b = self.visit(e.sources[0]), self.visit(e.sources[1])
        return self._node(NodeType.Zip, [(a, PW), (b, PW)], a.partition_count,
                          DataSetInfo(RandomPartition(a.partition_count)), e.schema, fn=e.params["fn"])

    def _visit_reverse(self, e: QExpr):
        child = self.visit(e.sources[0])
        return self._node(NodeType.Reverse, [(child, PW)], child.partition_count,
                          DataSetInfo(RandomPartition(child.partition_count)), e.schema)

    def _visit_sliding_window(self, e: QExpr):
        child = self.visit(e.sources[0])
        return self._node(NodeType.SlidingWindow, [(child, PW)], child.partition_count,
                          DataSetInfo(RandomPartition(child.partition_count)), e.schema, fn=e.params["fn"],
                          window=e.params["window"])

    # prefix operators (VisitPartitionOp :2546-2624)
    def _partition_op(self, e: QExpr, op: str):
        child = self.visit(e.sources[0])
        p = e.params
        pc = child.partition_count
        if op == "take" and pc > 1:
            child = self._node(NodeType.PartitionOp, [(child, PW)], pc, child.info, child.schema, op="take",
                               n=p["n"], stage=1)
        one = self._to_one(child, child.info.order)
        kids = [(one, PW)]
        return self._node(NodeType.PartitionOp, kids, 1, DataSetInfo(RandomPartition(1), child.info.order), e.schema,
                          op=op, n=p.get("n"), fn=p.get("fn"), indexed=p.get("indexed", False), stage=2)

    def _visit_take(self, e): return self._partition_op(e, "take")
    def _visit_skip(self, e): return self._partition_op(e, "skip")
    def _visit_take_while(self, e): return self._partition_op(e, "take_while")
    def _visit_skip_while(self, e): return self._partition_op(e, "skip_while")

    # aggregates (VisitBasicAggregate :1954-2005, VisitAggregate :1925-1952)
    def _visit_aggregate(self, e: QExpr):
        from ..api.decomposable import is_associative
        p = e.params
        kind = p["kind"]
        child = self.visit(e.sources[0])
        pc = child.partition_count
        if kind == "sequence_equal":
            other = self.visit(e.sources[1])
            a, b = self._to_one(child), self._to_one(other)
            return self._node(NodeType.BasicAggregate, [(a, PW), (b, PW)], 1, None, None, kind=kind, stage=2)
        if kind == "aggregate":
            assoc = is_associative(p["fn"].fn) and not p["has_seed"] and p.get("result") is None
            if assoc and pc > 1:
                s1 = self._node(NodeType.Aggregate, [(child, PW)], pc, None, None, fn=p["fn"], stage=1, has_seed=False)
                s1.dynamic_manager = DynamicManagerType.FullAggregator if (self._dyn_level & SC.DYN_AGGREGATE) else DynamicManagerType.NoManager
                one = self._to_one(s1)
                return self._node(NodeType.Aggregate, [(one, PW)], 1, None, None, fn=p["fn"], stage=2, has_seed=False)
            one = self._to_one(child)
            return self._node(NodeType.Aggregate, [(one, PW)], 1, None, None, fn=p["fn"], stage=0, has_seed=p["has_seed"],
                              seed=p.get("seed"), seed_fn=p.get("seed_fn"), result=p.get("result"))
        if kind == "contains":
            s1 = self._node(NodeType.Contains, [(child, PW)], pc, None, {"value": "b"}, value=p["value"], stage=1)
            one = self._to_one(s1)
            return self._node(NodeType.BasicAggregate, [(one, PW)], 1, None, {"value": "b"}, kind="any_true", stage=2)
        # schemas of the partials and of the result, where the types are known: they let the partials
        # travel between ranks as tensors and the scalar come back as a column
        fn = p.get("fn")
        vk = None
        if fn is not None and isinstance(getattr(fn, "ir", None), X.Expr):
            vk = fn.ir.kind
        elif fn is None and child.schema is not None and list(child.schema) == ["value"]:
            vk = child.schema["value"]
        s1_schema = s2_schema = None
        if kind == "count":
            s1_schema = s2_schema = {"value": "i"}
        elif kind in ("any", "all"):
            s1_schema = s2_schema = {"value": "b"}
        elif kind in ("sum", "min", "max") and vk in ("i", "f"):
            s1_schema = s2_schema = {"value": vk}
        elif kind == "average" and vk in ("i", "f"):
            s1_schema, s2_schema = {"item1": vk, "item2": "i"}, {"value": "f"}
        elif kind in ("first", "last") and child.schema is not None:
            s1_schema = s2_schema = dict(child.schema)
        s1 = self._node(NodeType.BasicAggregate, [(child, PW)], pc, None, s1_schema, kind=kind, stage=1, fn=p.get("fn"),
                        default=p.get("default", False))
        if self._dyn_level & SC.DYN_AGGREGATE:
            s1.dynamic_manager = DynamicManagerType.FullAggregator
        one = self._to_one(s1)
        return self._node(NodeType.BasicAggregate, [(one, PW)], 1, None, s2_schema, kind=kind, stage=2, fn=p.get("fn"),
                          default=p.get("default", False))

    # apply / fork (VisitApply :2925-3081, VisitFork :3083-3100)
    def _visit_apply(self, e: QExpr):
        p = e.params
        child = self.visit(e.sources[0])
        other = self.visit(e.sources[1]) if len(e.sources) > 1 else None
        if not p["per_partition"]:
            kids = [(self._to_one(child), PW)]
            if other is not None:
                kids.append((self._to_one(other), PW))
            return self._node(NodeType.Apply, kids, 1, DataSetInfo(RandomPartition(1)), e.schema, fn=p["fn"],
                              per_partition=False)
        pc = child.partition_count
        kids = [(child, PW)]
        if other is not None:
            if p["first_only"]:
                kids.append((self._broadcast(other, pc), PW))
            else:
                if other.partition_count != pc:
                    raise DryadLinqException(ErrorCode.HomomorphicApplyNeedsSamePartitionCount,
                                             "a fully homomorphic Apply needs inputs with equal partition counts")
                kids.append((other, PW))
        return self._node(NodeType.Apply, kids, pc, DataSetInfo(RandomPartition(pc)), e.schema, fn=p["fn"],
                          per_partition=True, with_index=p.get("with_index", False))

    def _visit_fork(self, e: QExpr):
        child = self.visit(e.sources[0])
        n = self._node(NodeType.Fork, [(child, PW)], child.partition_count, None, None, fn=e.params["fn"],
                       n=e.params["n"], per_partition=e.params["per_partition"])
        n.n_out_ports = e.params["n"]
        return n

    def _visit_fork_branch(self, e: QExpr):
        fk = self.visit(e.sources[0])
        return self._node(NodeType.ForkBranch, [(fk, PW)], fk.partition_count,
                          DataSetInfo(RandomPartition(fk.partition_count)), None, index=e.params["index"])

    # ---- phase 2: Tee insertion + pipeline fusion ------------------------------------------
    def _all_nodes(self) -> List[DLinqQueryNode]:
        seen, order = set(), []

        def rec(n):
            if id(n) in seen:
                return
            seen.add(id(n))
            for c, _ in n.children:
                rec(c)
            order.append(n)
        for r in self.roots:
            rec(r)
        return order

    def _phase2(self):
        # Tee where an output is consumed by several parents (except Fork, whose branches select ports)
        for n in self._all_nodes():
            if n.is_forked and n.type not in (NodeType.Tee, NodeType.Fork, NodeType.InputTable):
                parents = list(n.parents)
                tee = DLinqQueryNode(NodeType.Tee, [(n, PW)], n.partition_count, n.info, n.schema, replicate=False)
                n.parents = [tee]
                for par in parents:
                    par.children = [(tee if c is n else c, op) for c, op in par.children]
                    tee.parents.append(par)
        # greedy pipeline reduction, children first
        for n in self._all_nodes():
            if n.super is not None or n.type == NodeType.Super:
                continue
            if n.can_pipeline_into_child():
                child = n.children[0][0]
                if child.type == NodeType.Super:
                    members = list(child.members) + [n]
                    self._dissolve(child)
                else:
                    members = [child, n]
                sup = make_super(members)
                self.roots = [sup if r is n else r for r in self.roots]

    @staticmethod
    def _dissolve(sup: DLinqQueryNode):
        head, tail = sup.members[0], sup.members[-1]
        for c, _ in sup.children:
            c.parents = [head if x is sup else x for x in c.parents]
        for par in sup.parents:
            par.children = [(tail if c is sup else c, op) for c, op in par.children]
        for m in sup.members:
            m.super = None

    # ---- phase 3: remove useless Tee / Merge ------------------------------------------------
    def _phase3(self):
        for n in self._all_nodes():
            if n.type == NodeType.Tee and not n.params.get("replicate") and len(n.parents) <= 1:
                self._splice_out(n)
            elif n.type == NodeType.Merge and n.children[0][1] == PW and not n.params.get("group") \
                    and n.children[0][0].partition_count == n.partition_count and n.params.get("sort_key") is None:
                self._splice_out(n)

        self._assign_fifo_channels()

    def _assign_fifo_channels(self):
        """StaticConfig.UseMemoryFIFO: the inputs of multi-input vertices (Join, set operations,
        multi-source Apply) stream through in-memory FIFOs instead of files, unless the child is
        an input table, is forked, has one partition, or both ends are stateful
        (LinqToDryad/DryadLinqQueryNode.cs:2094-2107, 3225-3238, 4245-4258).  A FIFO edge puts
        producer and consumer into one start clique of the graph manager."""
        from . import static_config
        from .qnodes import ChannelType
        if not getattr(self.context, "use_memory_fifo", static_config.USE_MEMORY_FIFO):
            return
        multi = (NodeType.Join, NodeType.GroupJoin, NodeType.SetOperation, NodeType.Apply)
        for n in self._all_nodes():
            head = n.members[0].type if n.type == NodeType.Super and n.members else n.type
            if head not in multi or len(n.children) < 2:
                continue
            stateful = n.is_stateful
            for child, _ in n.children:
                if child.type != NodeType.InputTable and not child.is_forked and child.partition_count > 1 \
                        and not (stateful and child.is_stateful):
                    child.channel_type = ChannelType.MemoryFIFO
                stateful = stateful or child.is_stateful

    def _splice_out(self, n: DLinqQueryNode):
        child, _ = n.children[0]
        child.parents = [x for x in child.parents if x is not n]
        for par in n.parents:
            par.children = [(child if c is n else c, op) for c, op in par.children]
            child.parents.append(par)
        self.roots = [child if r is n else r for r in self.roots]

    def _assign_unique_ids(self):
        self.nodes = self._all_nodes()
        uid = 0
        for n in self.nodes:
            n.uid = uid
            uid += 1
            for m in n.members:
                m.uid = uid
                uid += 1

    # ---- explain / XML -----------------------------------------------------------------------
    def explain(self) -> str:
        self.build()
        lines = []
        for n in self.nodes:
            kids = ", ".join(f"#{c.uid}{'×' if op == CP else ''}" for c, op in n.children)
            lines.append(f"#{n.uid} {n.type.value} x{n.partition_count} <- [{kids}]  {{{n.info.describe()}}}")
            if n.type == NodeType.Super:
                for m in n.members:
                    lines.append(f"      | {m.describe()}")
            else:
                lines.append(f"      {n.describe()}")
        return "\n".join(lines)

    def _collect_objects(self):
        """Closure objects a plan refers to by index (comparers, explicit range keys, aggregate seeds):
        DryadLinqObjectStore role — the client puts them, the plan carries the indices, the job
        directory carries the serialised store (LinqToDryad/DryadLinqObjectStore.cs:43-173)."""
        from .object_store import ObjectStore
        store = ObjectStore()
        refs = {}
        for n in self.nodes:
            members = n.members if n.type == NodeType.Super else [n]
            for m in members:
                for name in ("comparer", "range_keys", "seed"):
                    v = m.params.get(name)
                    if v is not None:
                        refs.setdefault(n.uid, []).append((name, store.put(v)))
        self.object_store, self.object_refs = store, refs
        return store, refs

    def save_object_store(self, path: str) -> bool:
        """Serialise the plan's closure objects next to the plan file; False when one of them cannot be
        pickled (a lambda comparer), in which case the closures only live in this process."""
        store, _ = self._collect_objects()
        if not len(store):
            return False
        try:
            store.save(path)
            return True
        except Exception:  # noqa: BLE001
            return False

    def to_xml(self) -> str:
        from .query_doc import QueryDoc
        self.build()
        ctx = self.context
        _store, refs = self._collect_objects()
        doc = QueryDoc("Query")
        doc.element("DryadLinqVersion", ctx.version).element("ClusterName", ctx.cluster_name)
        doc.element("Visualization", "none")
        doc.element("IntermediateDataCompression", ctx.intermediate_data_compression_scheme)
        doc.element("EnableSpeculativeDuplication", str(ctx.enable_speculative_duplication).lower())
        doc.open("XmlExecHostArgs")
        for a in ctx.exec_host_args:
            doc.element("Argument", a)
        doc.close("XmlExecHostArgs")
        doc.element("ObjectStoreCount", len(_store))
        doc.open("QueryPlan")
        for n in self.nodes:
            doc.open("Vertex")
            doc.element("UniqueId", n.uid).element("Type", XML_VERTEX_TYPE.get(n.type, "WHERE")).element("Name", n.name)
            doc.cdata("Explain", n.describe() if n.type != NodeType.Super else " | ".join(m.describe() for m in n.members))
            doc.element("Partitions", n.partition_count).element("ChannelType", n.channel_type.value.upper())
            conn = CP if any(op == CP for _, op in n.children) else PW
            doc.element("ConnectionOperator", conn.value.upper())
            doc.open("DynamicManager").element("Type", n.dynamic_manager.value)
            if n.dynamic_manager in (DynamicManagerType.PartialAggregator, DynamicManagerType.FullAggregator):
                doc.element("AggregationLevels", n.aggregation_levels)
            if n.dynamic_manager == DynamicManagerType.RangeDistributor:
                doc.element("SampleRate", SC.SAMPLE_RATE)
            doc.close("DynamicManager")
            if n.type in (NodeType.InputTable, NodeType.OutputTable):
                uri = n.params.get("uri") or ""
                doc.open("StorageSet").element("Type", "PartitionedFile" if uri else "Memory")
                doc.element("SourceURI" if n.type == NodeType.InputTable else "SinkURI", uri)
                if n.type == NodeType.OutputTable:
                    doc.element("IsTemporary", str(bool(n.params.get("temporary"))).lower())
                    doc.element("OutputCompressionScheme", n.params.get("compression") or "None")
                doc.close("StorageSet")
            else:
                doc.open("Entry").element("AssemblyName", "dryad_b200.runtime.vertex_ops")
                doc.element("ClassName", "DryadLinq__Vertex").element("MethodName", f"{n.type.value}__{n.uid}")
                doc.close("Entry")
            if n.uid in refs:
                doc.open("ObjectStoreRefs")
                for name, idx in refs[n.uid]:
                    doc.element("Ref", idx, {"name": name})
                doc.close("ObjectStoreRefs")
            doc.open("Children")
            for c, op in n.children:
                doc.open("Child").element("UniqueId", c.uid).element("ConnectionOperator", op.value.upper())
                doc.element("AffinityConstraint", "UseDefault").close("Child")
            doc.close("Children")
            doc.close("Vertex")
        doc.close("QueryPlan")
        return '<?xml version="1.0" encoding="utf-8"?>\n' + doc.to_string()
