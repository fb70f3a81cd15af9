"""Query-plan file model and parser (reference: DryadLinqGraphManager/Query.cs:33-161 — Query,
Vertex, VertexInfo, Predecessor, DynamicManager — and QueryPlanParser.ParseQueryXml,
DryadLinqGraphManager/QueryParser.cs:129-347).  The XML written by `QueryGen.to_xml()` is the
contract between the client planner and the graph manager; this module reads it back so a plan
file is enough to rebuild the stage graph (`to_graph_manager`)."""
from __future__ import annotations

import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import Dict, List, Optional


@dataclass
class Predecessor:
    unique_id: int
    connection_operator: str = "POINTWISE"        # POINTWISE | CROSSPRODUCT
    affinity_constraint: str = "UseDefault"


@dataclass
class DynamicManagerInfo:
    type: str = "None"
    aggregation_levels: int = 0
    sample_rate: float = 0.0


@dataclass
class StorageSet:
    type: str = ""
    uri: str = ""
    is_temporary: bool = False
    compression: str = "None"


@dataclass
class Vertex:
    unique_id: int
    type: str                                     # INPUTTABLE, OUTPUTTABLE, WHERE, JOIN, FORK, TEE, CONCAT, SUPER
    name: str
    partitions: int
    channel_type: str = "DISKFILE"
    connection_operator: str = "POINTWISE"
    dynamic_manager: DynamicManagerInfo = field(default_factory=DynamicManagerInfo)
    storage: Optional[StorageSet] = None
    entry: Optional[tuple] = None                 # (assembly, class, method)
    children: List[Predecessor] = field(default_factory=list)
    explain: str = ""
    machines: List[str] = field(default_factory=list)


@dataclass
class QueryPlan:
    version: str = ""
    cluster_name: str = ""
    intermediate_compression: str = "None"
    enable_speculative_duplication: bool = True
    exec_host_args: List[str] = field(default_factory=list)
    vertices: Dict[int, Vertex] = field(default_factory=dict)

    def topological(self) -> List[Vertex]:
        return [self.vertices[k] for k in sorted(self.vertices)]     # ids are assigned children-first


def _text(e: Optional[ET.Element], default: str = "") -> str:
    return (e.text or default).strip() if e is not None and e.text is not None else default


def parse_query_xml(xml_text: str) -> QueryPlan:
    root = ET.fromstring(xml_text)
    if root.tag != "Query":
        raise ValueError("not a query plan: root element is " + root.tag)
    q = QueryPlan(version=_text(root.find("DryadLinqVersion")), cluster_name=_text(root.find("ClusterName")),
                  intermediate_compression=_text(root.find("IntermediateDataCompression"), "None"),
                  enable_speculative_duplication=_text(root.find("EnableSpeculativeDuplication"), "true").lower() == "true")
    args = root.find("XmlExecHostArgs")
    if args is not None:
        q.exec_host_args = [_text(a) for a in args.findall("Argument")]
    plan = root.find("QueryPlan")
    if plan is None:
        raise ValueError("query plan has no <QueryPlan> element")
    for ve in plan.findall("Vertex"):
        v = Vertex(unique_id=int(_text(ve.find("UniqueId"))), type=_text(ve.find("Type")), name=_text(ve.find("Name")),
                   partitions=int(_text(ve.find("Partitions"), "1")), channel_type=_text(ve.find("ChannelType"), "DISKFILE"),
                   connection_operator=_text(ve.find("ConnectionOperator"), "POINTWISE"), explain=_text(ve.find("Explain")))
        dm = ve.find("DynamicManager")
        if dm is not None:
            v.dynamic_manager = DynamicManagerInfo(_text(dm.find("Type"), "None"), int(_text(dm.find("AggregationLevels"), "0")),
                                                   float(_text(dm.find("SampleRate"), "0")))
        ss = ve.find("StorageSet")
        if ss is not None:
            uri = _text(ss.find("SourceURI")) or _text(ss.find("SinkURI"))
            v.storage = StorageSet(_text(ss.find("Type")), uri, _text(ss.find("IsTemporary"), "false").lower() == "true",
                                   _text(ss.find("OutputCompressionScheme"), "None"))
        en = ve.find("Entry")
        if en is not None:
            v.entry = (_text(en.find("AssemblyName")), _text(en.find("ClassName")), _text(en.find("MethodName")))
        ch = ve.find("Children")
        if ch is not None:
            for c in ch.findall("Child"):
                v.children.append(Predecessor(int(_text(c.find("UniqueId"))), _text(c.find("ConnectionOperator"), "POINTWISE"),
                                              _text(c.find("AffinityConstraint"), "UseDefault")))
        ms = ve.find("Machines")
        if ms is not None:
            v.machines = [_text(m) for m in ms.findall("Machine")]
        if v.unique_id in q.vertices:
            raise ValueError(f"duplicate vertex id {v.unique_id}")
        q.vertices[v.unique_id] = v
    for v in q.vertices.values():
        for p in v.children:
            if p.unique_id not in q.vertices:
                raise ValueError(f"vertex {v.unique_id} references unknown child {p.unique_id}")
    return q


def to_graph_manager(plan: QueryPlan, max_failures: int = 6, app=None):
    """Build the native graph manager's stage graph from a parsed plan file (GraphBuilder role,
    DryadLinqGraphManager/GraphBuilder.cs:564-748).  Edge channel types follow the child's
    <ChannelType> (DISKFILE -> file edge, MEMORYFIFO / TCPPIPE -> streamed edge = start clique,
    GraphBuilder.cs:508-562).  `app` (sched.app.DryadLinqApp) receives the plan-vertex -> stage
    identity map."""
    from ..sched.gm import CHANNEL_FIFO, CHANNEL_FILE, CROSS_PRODUCT, POINTWISE, GraphManager
    gm = GraphManager(max_failures=max_failures, speculative=plan.enable_speculative_duplication)
    sid = {}
    for v in plan.topological():
        sid[v.unique_id] = gm.add_stage(f"{v.name}#{v.unique_id}", v.partitions)
        if app is not None:
            app.record_identity(v.unique_id, sid[v.unique_id], v.name)
    for v in plan.topological():
        for p in v.children:
            child = plan.vertices[p.unique_id]
            kind = CROSS_PRODUCT if p.connection_operator == "CROSSPRODUCT" else POINTWISE
            if kind == POINTWISE and not (child.partitions == v.partitions or child.partitions == 1 or v.partitions == 1
                                          or (v.partitions and child.partitions % v.partitions == 0)):
                kind = CROSS_PRODUCT
            streamed = kind == POINTWISE and child.partitions == v.partitions and \
                child.channel_type.upper() in ("MEMORYFIFO", "TCPPIPE")
            gm.connect(sid[p.unique_id], sid[v.unique_id], kind, CHANNEL_FIFO if streamed else CHANNEL_FILE)
    return gm, sid
