"""Plan file round trip, golden plan shapes, object store, small utilities."""
import pytest

import dryad_b200 as D
from dryad_b200.plan import xmlplan
from dryad_b200.plan.object_store import ObjectStore
from dryad_b200.plan.querygen import QueryGen
from dryad_b200.utils.collections import BitVector, ReadOnlyAfterSubmitDict, ReadOnlyAfterSubmitList


def _plan(q):
    return QueryGen(q.context, [q.materialize_lazy()]).build()


def test_xml_plan_round_trip_and_graph_manager_from_file():
    ctx = D.DryadLinqContext(3)
    q = (ctx.from_enumerable(["a b", "c"], 3).select_many(lambda l: l.split(" ")).group_by(lambda w: w)
         .select(lambda g: (g.key, g.count())))
    gen = _plan(q)
    plan = xmlplan.parse_query_xml(gen.to_xml())
    assert plan.version == ctx.version and plan.enable_speculative_duplication
    assert [v.unique_id for v in plan.topological()] == [n.uid for n in gen.nodes]
    by_id = {n.uid: n for n in gen.nodes}
    for v in plan.topological():
        n = by_id[v.unique_id]
        assert v.partitions == n.partition_count
        assert [p.unique_id for p in v.children] == [c.uid for c, _ in n.children]
    assert any(p.connection_operator == "CROSSPRODUCT" for v in plan.vertices.values() for p in v.children)
    assert plan.vertices[0].type == "INPUTTABLE" and any(v.type == "SUPER" for v in plan.vertices.values())
    gm, sid = xmlplan.to_graph_manager(plan)
    gm.start()
    ran = 0
    while True:
        nxt = gm.next_ready(0.0)
        if nxt is None:
            break
        gm.completed(nxt[0], nxt[1], 0.01)
        ran += 1
    assert gm.state == "Succeeded" and ran == sum(v.partitions for v in plan.vertices.values())
    with pytest.raises(ValueError):
        xmlplan.parse_query_xml("<NotAQuery/>")


def _shape(q):
    out = []
    for n in _plan(q).nodes:
        out.append(n.type.value if not n.members else "Super[" + ",".join(m.type.value for m in n.members) + "]")
    return out


def test_golden_plan_shapes():
    ctx = D.DryadLinqContext(4)
    src = lambda: ctx.from_enumerable(range(100), 4)
    # decomposable GroupBy: partial -> HashPartition | Merge -> final -> finalizer (QueryGen.cs:2094-2357)
    assert _shape(src().group_by(lambda x: x % 5).select(lambda g: (g.key, g.sum()))) == [
        "InputTable", "Super[GroupBy,HashPartition]", "Super[Merge,GroupBy,Select,OutputTable]"]
    # OrderBy: sample -> one vertex -> separators; RangePartition x Merge; Sort (QueryGen.cs:2362-2521)
    assert _shape(src().order_by(lambda x: x)) == [
        "InputTable", "Sampler", "Super[Merge,RangeSampler]", "RangePartition", "Super[Merge,OrderBy,OutputTable]"]
    # scalar aggregate: stage 1 per partition, stage 2 on one partition (QueryGen.cs:1954-2005)
    assert _shape(src().sum_as_query()) == ["InputTable", "BasicAggregate", "Super[Merge,BasicAggregate,OutputTable]"]
    # Take(n): per-partition take, ordered merge to one partition, final take (QueryGen.cs:2546-2624)
    assert _shape(src().take(3)) == ["InputTable", "PartitionOp", "Super[Merge,PartitionOp,OutputTable]"]
    # non-homomorphic Apply merges to one partition first (QueryGen.cs:2925-2955)
    assert _shape(src().apply(lambda xs: xs)) == ["InputTable", "Super[Merge,Apply,OutputTable]"]
    # Where/Select pipeline fuses into one super vertex with the output
    assert _shape(src().where(lambda x: x > 1).select(lambda x: x + 1)) == ["InputTable", "Super[Where,Select,OutputTable]"]
    # Distinct: hash partition by element, then Distinct
    assert _shape(src().distinct()) == ["InputTable", "HashPartition", "Super[Merge,Distinct,OutputTable]"]


def test_object_store_and_utils(tmp_path):
    s = ObjectStore()
    cmp_keys = [3, 7, 11]
    i = s.put(cmp_keys)
    assert s.put(cmp_keys) == i and s.get(i) is cmp_keys
    s.save(str(tmp_path / "objs.bin"))
    with pytest.raises(RuntimeError):
        s.put("late")
    assert ObjectStore.load(str(tmp_path / "objs.bin")).get(i) == cmp_keys
    d, l = ReadOnlyAfterSubmitDict(a=1), ReadOnlyAfterSubmitList([1])
    d["b"] = 2
    l.append(2)
    d.freeze()
    l.freeze()
    with pytest.raises(D.DryadLinqException):
        d["c"] = 3
    with pytest.raises(D.DryadLinqException):
        l.append(3)
    bv = BitVector(4)
    bv.set(2)
    bv.set(70)
    assert bv[2] and bv[70] and not bv[3] and bv.count() == 2 and list(bv) == [2, 70]


def test_graph_manager_app_options_identity_map_and_streamed_edges(tmp_path):
    """DryadLINQApp option table (-mai/-mafi/-at/-nca), the <plan>.map identity file, and plan
    edges whose child channel is MEMORYFIFO becoming start cliques in the graph manager."""
    import pytest
    from dryad_b200.plan import static_config
    from dryad_b200.plan.xmlplan import parse_query_xml, to_graph_manager
    from dryad_b200.sched.app import DryadLinqApp, parse_data_size
    import dryad_b200 as D
    assert parse_data_size("12KB") == 12 << 10 and parse_data_size("20.5mb") == int(20.5 * (1 << 20)) and parse_data_size("77") == 77
    with pytest.raises(ValueError):
        parse_data_size("1.5")
    app = DryadLinqApp.parse(["-mai", "64", "--maxaggregatefilterinputs", "8", "-at", "256MB", "-nca", "-amd64", "plan.xml"])
    assert (app.max_aggregate_inputs, app.max_aggregate_filter_inputs, app.aggregate_threshold, app.cluster_affinity) == (64, 8, 256 << 20, False)
    assert app.positional == ["plan.xml"] and "-mai" in DryadLinqApp.usage()
    with pytest.raises(ValueError):
        DryadLinqApp.parse(["-bogus"])
    # -nca: grouping ignores topology; the input cap applies
    groups = app.aggregate_groups([10] * 20, machine=list(range(20)), pod=[i % 4 for i in range(20)], filtered=True)
    assert max(groups) + 1 == 3 and all(groups.count(g) <= 8 for g in set(groups))
    old = static_config.USE_MEMORY_FIFO
    static_config.USE_MEMORY_FIFO = True
    try:
        ctx = D.DryadLinqContext(3)
        a = ctx.from_enumerable(list(range(60)), 3).select(lambda x: x + 1).hash_partition(lambda x: x, 3)
        b = ctx.from_enumerable(list(range(30)), 3).select(lambda x: x * 2).hash_partition(lambda x: x, 3)
        xml = a.join(b, lambda x: x, lambda y: y, lambda x, y: x + y).to_xml()
    finally:
        static_config.USE_MEMORY_FIFO = old
    plan = parse_query_xml(xml)
    (tmp_path / "plan.xml").write_text(xml)
    app.open_identity_map(str(tmp_path / "plan.xml"))
    gm, sid = to_graph_manager(plan, app=app)
    ident = DryadLinqApp.read_identity_map(tmp_path / "plan.xml.map")
    assert {k: v[0] for k, v in ident.items()} == sid
    gm.start()
    assert any(len(gm.clique(v)) > 1 for v in range(gm.live_vertices))


def test_fork_value_type_system_and_query_doc():
    import xml.etree.ElementTree as ET
    import dryad_b200 as D
    from dryad_b200.plan import type_system as T
    from dryad_b200.plan.query_doc import QueryDoc
    e, v = D.ForkValue(), D.ForkValue(5)
    assert not e.has_value and v.has_value and v.value == 5 and hash(e) == 0 and e == D.ForkValue() and v != e
    e.value = 5
    assert e == v and e.has_value
    assert T.is_query_operator("GroupBy") and T.is_query_operator("select_many") and not T.is_query_operator("frobnicate")
    assert T.is_identity_function(None) and T.is_identity_function(lambda x: x) and not T.is_identity_function(lambda x: x + 1)
    assert T.is_anonymous_type({"a": 1}) and T.schema_of_record((1, 2.0)) == {"item1": "i", "item2": "f"}
    assert T.kind_of_value(True) == "b" and T.kind_of_value(1 << 70) == "o"
    assert any(p.endswith("type_system.py") for p in T.loaded_non_system_modules())
    doc = QueryDoc("Query").element("ClusterName", "local").open("Vertex", {"id": 3}).cdata("Explain", "x => x < 3 && y ]]> 0").close("Vertex")
    root = ET.fromstring(doc.to_string())
    assert root.find("ClusterName").text == "local" and root.find("Vertex/Explain").text == "x => x < 3 && y ]]> 0"


def test_golden_plans_of_dataset_info_elision_rules():
    """The DataSetInfo subtleties: which shuffles and sorts the planner skips
    because the data is already suitably partitioned / ordered, the broadcast-join rule, Fork / Tee,
    Assume*, Concat, and Where pushed below OrderBy (SimpleRewriter)."""
    ctx = D.DryadLinqContext(4)
    ctx.broadcast_join_rows = 0                      # size rule off unless a case turns it on
    src = lambda: ctx.from_enumerable([{"k": i % 7, "v": i} for i in range(100)], 4)   # noqa: E731
    K = lambda r: r.k                                # noqa: E731
    J = lambda a, b: {"k": a.k, "s": a.v + b.v}      # noqa: E731
    shuffled = ["InputTable", "HashPartition", "Merge"]
    # Join: neither side partitioned on the key -> both shuffled (QueryGen.cs:1483-1608)
    assert _shape(src().join(src(), K, K, J)) == shuffled + shuffled + ["Super[Join,OutputTable]"]
    # both sides already hash-partitioned on the key into the same count: NO further shuffle
    assert _shape(src().hash_partition(K, 4).join(src().hash_partition(K, 4), K, K, J)) == \
        shuffled + shuffled + ["Super[Join,OutputTable]"]
    # one side partitioned: only the OTHER side is shuffled, into the same partition count
    one = _plan(src().hash_partition(K, 4).join(src(), K, K, J))
    assert [n.type.value for n in one.nodes].count("HashPartition") == 2 and one.nodes[-1].partition_count == 4
    # set operations hash both sides on the element; Distinct after a matching HashPartition adds nothing
    setop = ["InputTable", "Super[Select,HashPartition]", "Merge"]
    assert _shape(src().select(lambda r: r.k).union(src().select(lambda r: r.v))) == setop + setop + ["Super[SetOperation,OutputTable]"]
    assert _shape(src().select(lambda r: r.k).hash_partition(lambda x: x, 4).distinct()) == [
        "InputTable", "Super[Select,HashPartition]", "Super[Merge,Distinct,OutputTable]"]
    # GroupBy over data already hash-partitioned on the key: one full aggregation, no partial stage, no shuffle
    assert _shape(src().hash_partition(K, 4).group_by(K).select(lambda g: {"k": g.key, "c": g.count()})) == [
        "InputTable", "HashPartition", "Super[Merge,GroupBy,Select,OutputTable]"]
    assert _shape(src().assume_hash_partition(K).group_by(K).select(lambda g: {"k": g.key, "c": g.count()})) == [
        "InputTable", "Dummy", "Super[GroupBy,Select,OutputTable]"]
    # OrderBy over data already range-partitioned on the key: only the per-partition sort is added;
    # a second OrderBy on the same key is dropped altogether
    ranged = ["InputTable", "Sampler", "Super[Merge,RangeSampler]", "RangePartition", "Super[Merge,OrderBy,OutputTable]"]
    assert _shape(src().range_partition(K, 4).order_by(K)) == ranged
    assert _shape(src().order_by(K).order_by(K)) == ranged
    # AssumeOrderBy: Take keeps the order (ordered merge to one partition), no sort is planned
    assert _shape(src().assume_order_by(K).take(5)) == ["InputTable", "Dummy", "PartitionOp", "Super[Merge,PartitionOp,OutputTable]"]
    # Where moves below OrderBy (and the shared input of sampler and distributor becomes a Tee)
    assert _shape(src().order_by(K).where(lambda r: r.v > 3)) == [
        "InputTable", "Where", "Tee", "Sampler", "Super[Merge,RangeSampler]", "RangePartition", "Super[Merge,OrderBy,OutputTable]"]
    # Concat keeps the partitions of both inputs side by side; a GroupBy after it re-partitions
    assert _shape(src().concat(src())) == ["InputTable", "InputTable", "Concat", "OutputTable"]
    assert _plan(src().concat(src())).nodes[2].partition_count == 8
    assert _shape(src().concat(src()).group_by(K).select(lambda g: {"k": g.key, "c": g.count()})) == [
        "InputTable", "InputTable", "Concat", "Super[GroupBy,HashPartition]", "Super[Merge,GroupBy,Select,OutputTable]"]
    # broadcast-join rule: a small inner source is replicated (Tee x CrossProduct Merge), the big side stays put
    ctx2 = D.DryadLinqContext(4)
    big = ctx2.from_enumerable([{"k": i % 7, "v": i} for i in range(100)], 4)
    small = ctx2.from_enumerable([{"k": i, "v": i * 2} for i in range(7)], 1)
    bplan = _plan(big.join(small, K, K, J))
    assert _shape(big.join(small, K, K, J)) == ["InputTable", "InputTable", "Tee", "Merge", "Super[Join,OutputTable]"]
    join = bplan.nodes[-1].members[0]
    assert join.params["broadcast"] and bplan.nodes[2].dynamic_manager.value == "Broadcast"
    # Fork: one Fork vertex, one ForkBranch per consumer; a dataset used by two queries of one job is a Tee
    f = src().fork(lambda r: (r.v if r.v % 2 == 0 else None, r.k if r.v % 2 else None), 2)
    multi = QueryGen(ctx, [f.first.select(lambda x: x + 1).materialize_lazy(), f.second.materialize_lazy()]).build()
    assert [n.type.value if not n.members else "Super[" + ",".join(m.type.value for m in n.members) + "]" for n in multi.nodes] == [
        "InputTable", "Fork", "Super[ForkBranch,Select,OutputTable]", "Super[ForkBranch,OutputTable]"]
    t = src().select(lambda r: r.v)
    tee = QueryGen(ctx, [t.where(lambda x: x > 5).materialize_lazy(), t.select(lambda x: x * 2).materialize_lazy()]).build()
    assert [n.type.value if not n.members else "Super[" + ",".join(m.type.value for m in n.members) + "]" for n in tee.nodes] == [
        "InputTable", "Select", "Tee", "Super[Where,OutputTable]", "Super[Select,OutputTable]"]


def test_plan_carries_its_closure_objects_and_static_config_switches(tmp_path):
    """ObjectStore in the plan writer (comparers / range keys by index), QueryDoc-built XML, and the
    GroupBy planner switch that turns local pre-aggregation off."""
    from dryad_b200.plan import static_config as SC
    ctx = D.DryadLinqContext(2, work_dir=str(tmp_path))
    q = ctx.from_enumerable(range(50), 2).range_partition(lambda x: x, 2, range_keys=[25])
    gen = _plan(q)
    xml = gen.to_xml()
    assert "<ObjectStoreCount>1</ObjectStoreCount>" in xml and '<Ref name="range_keys">0</Ref>' in xml
    assert gen.object_store.get(0) == [25]
    assert gen.save_object_store(str(tmp_path / "os.bin")) and ObjectStore.load(str(tmp_path / "os.bin")).get(0) == [25]
    assert xmlplan.parse_query_xml(xml).version == ctx.version
    q.to_list()                                              # the job directory carries the store too
    import os
    assert os.path.exists(os.path.join(str(tmp_path), "LocalJobs", "1", "DryadLinqObjectStore__0.bin"))
    assert os.path.exists(os.path.join(str(tmp_path), "DryadLinqClient.log")) or ctx.client_log.level < 4
    old = SC.GROUP_BY_LOCAL_AGGREGATION_IS_PARTIAL
    try:
        SC.GROUP_BY_LOCAL_AGGREGATION_IS_PARTIAL = False
        c2 = D.DryadLinqContext(3)
        shape = _shape(c2.from_enumerable(range(30), 3).group_by(lambda x: x % 4).select(lambda g: (g.key, g.count())))
        assert shape == ["InputTable", "HashPartition", "Super[Merge,GroupBy,Select,OutputTable]"]
        got = sorted(c2.from_enumerable(range(30), 3).group_by(lambda x: x % 4).select(lambda g: (g.key, g.count())).to_list())
        assert got == [(0, 8), (1, 8), (2, 7), (3, 7)]
    finally:
        SC.GROUP_BY_LOCAL_AGGREGATION_IS_PARTIAL = old
    # settings collections become read-only once a job was submitted
    with pytest.raises(D.DryadLinqException):
        ctx.job_environment_variables["X"] = "1"
