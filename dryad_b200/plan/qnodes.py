"""Physical plan IR — the DLinqQueryNode hierarchy (reference:
LinqToDryad/DryadLinqQueryNode.cs:39-4920: QueryNodeType :39, AggregateOpType :75, ChannelType :93,
ConnectionOpType :100, base node :115-835, pipelining rules :523-638, XML emission :728-827,
DLinqSuperNode :4001-4227, DLinqTeeNode :4678) plus the dynamic-manager tag
(LinqToDryad/DynamicManager.cs:35-202)."""
from __future__ import annotations

import itertools
from enum import Enum
from typing import Any, Dict, List, Optional, Tuple

from .dataset_info import DataSetInfo, RandomPartition


class NodeType(str, Enum):
    InputTable = "InputTable"
    OutputTable = "OutputTable"
    Where = "Where"
    Select = "Select"
    SelectMany = "SelectMany"
    Zip = "Zip"
    OrderBy = "OrderBy"
    GroupBy = "GroupBy"
    PartitionOp = "PartitionOp"
    Join = "Join"
    GroupJoin = "GroupJoin"
    Distinct = "Distinct"
    Contains = "Contains"
    BasicAggregate = "BasicAggregate"
    Aggregate = "Aggregate"
    Concat = "Concat"
    SetOperation = "SetOperation"
    Merge = "Merge"
    HashPartition = "HashPartition"
    RangePartition = "RangePartition"
    Super = "Super"
    Apply = "Apply"
    Fork = "Fork"
    DoWhile = "DoWhile"
    Tee = "Tee"
    Dynamic = "Dynamic"
    Dummy = "Dummy"
    # explicit forms of helper lambdas the reference hides inside Apply/Select nodes
    Sampler = "Sampler"              # Phase1Sampling (DryadLinqSampler.cs:42-106)
    RangeSampler = "RangeSampler"    # RangeSamplerCore (DryadLinqSampler.cs:139-246)
    Offsets = "Offsets"              # DryadLinqEnumerable.Offsets :1131
    Reverse = "Reverse"
    SlidingWindow = "SlidingWindow"
    ForkBranch = "ForkBranch"


class ChannelType(str, Enum):
    DiskFile = "DiskFile"        # default: a materialised partition buffer (HBM / host / file)
    MemoryFIFO = "MemoryFIFO"
    TCPPipe = "TCPPipe"


class ConnOp(str, Enum):
    Pointwise = "Pointwise"
    CrossProduct = "CrossProduct"


class DynamicManagerType(str, Enum):
    NoManager = "None"
    Splitter = "Splitter"
    PartialAggregator = "PartialAggregator"
    FullAggregator = "FullAggregator"
    HashDistributor = "HashDistributor"
    RangeDistributor = "RangeDistributor"
    Broadcast = "Broadcast"


# The XML <Type> vocabulary of the plan file (DryadLinqGraphManager/Query.cs:33-161).
XML_VERTEX_TYPE = {
    NodeType.InputTable: "INPUTTABLE", NodeType.OutputTable: "OUTPUTTABLE", NodeType.Where: "WHERE",
    NodeType.Join: "JOIN", NodeType.Fork: "FORK", NodeType.Tee: "TEE", NodeType.Concat: "CONCAT",
    NodeType.Super: "SUPER",
}

# Nodes a pipeline never fuses ACROSS (DryadLinqQueryNode.cs:523-533).
_NO_PIPELINE = {NodeType.InputTable, NodeType.Concat, NodeType.Tee, NodeType.DoWhile, NodeType.Dummy,
                NodeType.Fork, NodeType.Zip, NodeType.Reverse, NodeType.SlidingWindow}
_DISTRIBUTORS = {NodeType.HashPartition, NodeType.RangePartition, NodeType.Offsets, NodeType.RangeSampler}
# Stateful operators hold a whole partition (or a table) before emitting.
_STATEFUL = {NodeType.OrderBy, NodeType.GroupBy, NodeType.Join, NodeType.GroupJoin, NodeType.Distinct,
             NodeType.SetOperation, NodeType.BasicAggregate, NodeType.Aggregate, NodeType.Apply,
             NodeType.PartitionOp, NodeType.Contains}

_ids = itertools.count(0)


class DLinqQueryNode:
    def __init__(self, ntype: NodeType, children: List[Tuple["DLinqQueryNode", ConnOp]], partition_count: int,
                 info: Optional[DataSetInfo] = None, schema=None, name: str = "", **params):
        self.uid = -1                      # assigned children-first by QueryGen.assign_unique_ids
        self._tmp = next(_ids)
        self.type = ntype
        self.name = name or ntype.value
        self.children: List[Tuple[DLinqQueryNode, ConnOp]] = list(children)
        self.parents: List[DLinqQueryNode] = []
        for c, _ in self.children:
            c.parents.append(self)
        self.partition_count = int(partition_count)
        self.info = info or DataSetInfo(RandomPartition(self.partition_count))
        self.schema = schema
        self.params: Dict[str, Any] = params
        self.n_out_ports = 1
        self.channel_type = ChannelType.DiskFile
        self.dynamic_manager = DynamicManagerType.NoManager
        self.aggregation_levels = 0
        self.super: Optional[DLinqQueryNode] = None
        self.members: List[DLinqQueryNode] = []     # for Super nodes, in execution order
        self.is_stateful = ntype in _STATEFUL

    # ---- structure -------------------------------------------------------------------------
    @property
    def is_forked(self) -> bool:
        return len(self.parents) > 1

    def replace_child(self, old: "DLinqQueryNode", new: "DLinqQueryNode") -> None:
        for i, (c, op) in enumerate(self.children):
            if c is old:
                self.children[i] = (new, op)
                if self in old.parents:
                    old.parents.remove(self)
                new.parents.append(self)

    # ---- pipelining (CanBePipelined / CanNotBePipelinedWith) ------------------------------
    def can_pipeline_into_child(self) -> bool:
        """True when this node may join the super-node of its (single) child."""
        if len(self.children) != 1:
            return False
        child, op = self.children[0]
        if op != ConnOp.Pointwise or child.partition_count != self.partition_count:
            return False
        if child.is_forked or child.n_out_ports > 1:
            return False
        tail = child.members[-1] if child.type == NodeType.Super else child
        pipe = child.members if child.type == NodeType.Super else [child]
        if self.type in _NO_PIPELINE or any(m.type in _NO_PIPELINE for m in pipe):
            return False
        if tail.type in _DISTRIBUTORS or tail.type == NodeType.OutputTable:
            return False
        if self.type in (NodeType.Merge, NodeType.Super, NodeType.ForkBranch):
            return False
        if self.type == NodeType.Apply and self.params.get("writes_stream"):
            return False
        if self.is_stateful and any(m.is_stateful for m in pipe):
            return False            # two stateful operators never share a pipeline
        return True

    # ---- description -----------------------------------------------------------------------
    def describe(self) -> str:
        p = self.params
        bits = []
        for k in ("fn", "key", "outer_key", "inner_key", "result"):
            v = p.get(k)
            if v is not None and hasattr(v, "describe"):
                bits.append(f"{k}={v.describe()}")
        for k in ("mode", "kind", "op", "stage", "algorithm", "n", "window", "descending", "uri"):
            if k in p and p[k] is not None and p[k] is not False:
                bits.append(f"{k}={p[k]}")
        if "decomp" in p and p["decomp"] is not None:
            bits.append("decomp=[" + p["decomp"].describe() + "]")
        return f"{self.name}({', '.join(bits)})"

    def __repr__(self):
        return f"<{self.type.value}#{self.uid if self.uid >= 0 else 't%d' % self._tmp} x{self.partition_count}>"


def make_super(members: List[DLinqQueryNode]) -> DLinqQueryNode:
    """Fuse a pipeline (execution order: first member is the head that reads the channels)."""
    head, tail = members[0], members[-1]
    sup = DLinqQueryNode.__new__(DLinqQueryNode)
    sup.uid, sup._tmp = -1, next(_ids)
    sup.type = NodeType.Super
    sup.name = "Super__" + "_".join(m.type.value for m in members)
    sup.children = list(head.children)
    sup.parents = list(tail.parents)
    sup.partition_count = tail.partition_count
    sup.info, sup.schema = tail.info, tail.schema
    sup.params = {}
    sup.n_out_ports = tail.n_out_ports
    sup.channel_type = tail.channel_type
    sup.dynamic_manager = tail.dynamic_manager
    sup.aggregation_levels = tail.aggregation_levels
    sup.super = None
    sup.members = list(members)
    sup.is_stateful = any(m.is_stateful for m in members)
    for m in members:
        m.super = sup
    for c, _ in sup.children:
        c.parents = [sup if x is head else x for x in c.parents]
    for par in sup.parents:
        par.children = [(sup if c is tail else c, op) for c, op in par.children]
    return sup
