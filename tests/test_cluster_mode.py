"""Cluster mode: a graph-manager process + vertex-host processes (reference roles:
DryadLinqGraphManager/Program.cs, DryadVertex/VertexHost/vertexHost/main.cpp, dvertexpncontrol.cpp,
ProcessService, LocalScheduler).  Everything here runs on CPU."""
import json
import os

import pytest

from dryad_b200.api.context import DryadLinqContext, PlatformKind
from dryad_b200.api.errors import DryadLinqException


def _ctx(tmp_path, n=3, **kw):
    ctx = DryadLinqContext(n, platform=PlatformKind.LOCAL_CLUSTER, work_dir=str(tmp_path))
    ctx.job_max_nodes = kw.pop("hosts", 2)
    for k, v in kw.items():
        setattr(ctx, k, v)
    return ctx


def test_cluster_queries_match_oracle(tmp_path):
    ctx = _ctx(tmp_path)
    oracle = DryadLinqContext(3, local_debug=True)
    data = [(i * 7919) % 1000 for i in range(3000)]
    other = [(k, f"name{k}") for k in range(0, 1000, 3)]

    def build(c):
        a = c.from_enumerable(data, 3)
        b = c.from_enumerable(other, 2)
        g = a.group_by(lambda x: x % 11).select(lambda g: (g.key, g.count(), sum(g)))
        j = a.join(b, lambda x: x, lambda t: t[0], lambda x, t: (x, t[1]))
        o = a.where(lambda x: x % 2 == 0).order_by(lambda x: x)
        return g, j, o

    g1, j1, o1 = build(ctx)
    g0, j0, o0 = build(oracle)
    ctx.submit_and_wait(g1, j1, o1)                        # three queries, ONE job
    assert sorted(g1) == sorted(g0)
    assert sorted(j1) == sorted(j0)
    assert list(o1) == list(o0)
    st = ctx.last_cluster_status
    assert st["state"] == "Success" and st["hosts"] == 2 and st["pump_messages"] > 0
    # every vertex ran in a vertex-host process, and locality was tracked per vertex
    hosts = {v["host"] for v in st["vertices"]}
    assert hosts and all(h.startswith("w") for h in hosts)
    assert any(v["cache_hits"] > 0 for v in st["vertices"])
    job_dir = os.path.join(str(tmp_path), "ClusterJobs", "1")
    assert os.path.exists(os.path.join(job_dir, "calypso.log"))
    log = open(os.path.join(job_dir, "calypso.log")).read()
    assert "schedule," in log and "state=queued" in log and "state=matched" in log        # IProcessWatcher transitions
    assert any(f.endswith(".chan") for f in os.listdir(os.path.join(job_dir, "channels")))


def test_cluster_store_roundtrip_and_text(tmp_path):
    ctx = _ctx(tmp_path)
    uri = f"partfile://{tmp_path}/squares.pt"
    ctx.submit_and_wait(ctx.from_enumerable(range(100), 3).select(lambda x: (x, x * x)).to_store(uri))
    assert not [f for f in os.listdir(tmp_path) if f.endswith(".tmp")]
    back = sorted(ctx.from_store(uri).where(lambda t: t[0] < 5))
    assert back == [(i, i * i) for i in range(5)]
    text = tmp_path / "words.txt"
    text.write_text("a b a\nc b a\n" * 50)
    wc = dict(ctx.from_store(f"file://{text}").select_many(lambda l: l.line.split(" "))
              .group_by(lambda w: w).select(lambda g: (g.key, g.count())))
    assert wc == {"a": 150, "b": 100, "c": 50}


def test_cluster_host_death_and_vertex_retry(tmp_path):
    marker = str(tmp_path / "killed")

    def inject(node, part, version):
        # the first attempt of partition 1 of the grouping stage takes its whole host process down;
        # the first attempt of partition 0 fails with an ordinary exception
        if "GroupBy" in node.name and version == 0:
            if part == 1 and not os.path.exists(marker):
                open(marker, "w").close()
                os._exit(7)
            if part == 0:
                raise RuntimeError("injected vertex failure")

    ctx = _ctx(tmp_path, fault_injector=inject)
    out = sorted(ctx.from_enumerable(range(600), 3).group_by(lambda x: x % 5).select(lambda g: (g.key, g.count())))
    assert out == [(k, 120) for k in range(5)]
    st = ctx.last_cluster_status
    assert st["host_restarts"] >= 1 and st["counters"]["failures"] >= 2
    log = open(os.path.join(str(tmp_path), "ClusterJobs", "1", "calypso.log")).read()
    assert "exited" in log and "injected vertex failure" in log
    from dryad_b200.tools.jobreport import build_job, render
    job = build_job(os.path.join(str(tmp_path), "ClusterJobs", "1"))
    assert any("vertex-host process" in d for d in job["diagnosis"])
    assert any(p["exitcode"] == 7 for p in job["processes"]) and "Vertex hosts:" in render(job)
    # the browser views as one self-contained page: stage graph + schedule lanes per computer, as inline SVG
    import re
    import xml.dom.minidom
    from dryad_b200.tools.jobreport import render_html
    page = render_html(job)
    svgs = re.findall(r"<svg.*?</svg>", page, re.S)
    assert len(svgs) == 2 and all(xml.dom.minidom.parseString(x) for x in svgs)
    assert page.count("#c0392b") >= 1                         # failed attempts are drawn in red on the schedule


def test_cluster_unreadable_channel_reruns_producer(tmp_path):
    from dryad_b200.runtime.executor import InputReadError

    def inject(node, part, version):
        if "GroupBy" in node.name and part == 0 and version == 0:
            prod = node.children[0][0]
            raise InputReadError(prod.uid, 1, "injected: channel file lost")

    ctx = _ctx(tmp_path, fault_injector=inject)
    out = sorted(ctx.from_enumerable(range(300), 3).group_by(lambda x: x % 3).select(lambda g: (g.key, g.count())))
    assert out == [(0, 100), (1, 100), (2, 100)]
    st = ctx.last_cluster_status
    assert st["counters"]["invalidations"] >= 1
    reruns = [v for v in st["vertices"] if v["version"] > 0]
    assert any("HashPartition" in v["stage_name"] or "InputTable" in v["stage_name"] for v in reruns)     # the producer ran again


def test_cluster_job_failure_publishes_nothing(tmp_path):
    def inject(node, part, version):
        if "Select" in node.name and part == 0:
            raise ValueError("always broken")

    ctx = _ctx(tmp_path, fault_injector=inject, max_failures_per_vertex=2)
    uri = f"partfile://{tmp_path}/never.pt"
    with pytest.raises(DryadLinqException) as e:
        ctx.submit_and_wait(ctx.from_enumerable(range(30), 3).select(lambda x: x + 1).to_store(uri))
    assert "always broken" in str(e.value)
    assert not os.path.exists(tmp_path / "never.pt")
    assert not [f for f in os.listdir(tmp_path) if f.startswith("never")]
    st = json.load(open(os.path.join(str(tmp_path), "ClusterJobs", "1", "status.json")))
    assert st["state"] == "Failure"


def test_cluster_submission_cancel(tmp_path):
    import time
    from dryad_b200.api.context import JobStatus
    from dryad_b200.api.submission import ClusterJobSubmission
    from dryad_b200.runtime import jobpackage

    def slow(node, part, version):
        time.sleep(30)

    ctx = DryadLinqContext(2, work_dir=str(tmp_path))
    ctx.fault_injector = slow
    q = ctx.from_enumerable(range(10), 2).select(lambda x: x).to_store(f"partfile://{tmp_path}/slow.pt")
    job_dir = str(tmp_path / "job")
    jobpackage.write(job_dir, ctx, [q.materialize_lazy()])
    sub = ClusterJobSubmission(job_dir, workers=2)
    sub.submit_job()
    assert sub.get_status() == JobStatus.Running
    time.sleep(3.0)
    sub.terminate_job()
    assert sub.get_status() == JobStatus.Cancelled
    assert not os.path.exists(tmp_path / "slow.pt")


def test_cluster_runtime_aggregation_tree(tmp_path):
    """DYN_AGGREGATE: internal vertices are spliced into the RUNNING graph between the partial
    aggregates and their N:1 consumer (DrDynamicAggregateManager)."""
    import operator
    from dryad_b200.plan import static_config as SC
    ctx = _ctx(tmp_path, n=12, hosts=3, dynamic_opt_level=SC.DYN_BROADCAST | SC.DYN_AGGREGATE, aggregation_tree_fan_in=3)
    data = list(range(1, 1201))
    q = ctx.from_enumerable(data, 12)
    assert q.sum() == sum(data)
    st = ctx.last_cluster_status
    assert st["dynamic"]["aggregate"] >= 3                 # 12 partials, at most 3 inputs per vertex
    internal = [v for v in st["vertices"] if v["stage_name"].startswith("Agg__")]
    assert len(internal) == st["dynamic"]["aggregate"] and all(v["records_out"] == 1 for v in internal)
    final = [v for v in st["vertices"] if "BasicAggregate" in v["stage_name"] and v["stage"] < 1_000_000 and v["part"] == 0
             and v["records_in"] and len(v["bytes_in"]) <= 3 + 1][-1:]
    assert final                                           # the consumer read at most fan-in inputs (+ none folded)
    log = open(os.path.join(str(tmp_path), "ClusterJobs", "1", "calypso.log")).read()
    assert "dynamic,manager=aggregate" in log
    # average and max through the same trees
    assert q.average() == sum(data) / len(data)
    assert q.max() == 1200
    # an associative function that is NOT commutative: groups are contiguous runs, the order of the partials is kept
    words = [chr(ord("a") + i % 26) for i in range(120)]
    cat = ctx.from_enumerable(words, 12).aggregate(operator.add)
    assert cat == "".join(words)
    assert ctx.last_cluster_status["dynamic"]["aggregate"] >= 3


def test_cluster_runtime_broadcast_copies(tmp_path):
    """DYN_BROADCAST (on by default, like the reference): a broadcast to >= 5 consumers goes through
    floor(sqrt(n)) copy vertices spliced in when the producer completes (DrDynamicBroadcast)."""
    ctx = _ctx(tmp_path, n=6, hosts=3)
    big = ctx.from_enumerable(range(600), 6)
    small = ctx.from_enumerable([(k, k * k) for k in range(0, 600, 50)], 2)
    out = sorted(big.join(small, lambda x: x, lambda t: t[0], lambda x, t: (x, t[1])))   # small inner side: broadcast join
    assert out == [(k, k * k) for k in range(0, 600, 50)]
    st = ctx.last_cluster_status
    assert st["dynamic"]["copy"] == 2                      # floor(sqrt(6))
    copies = [v for v in st["vertices"] if v["stage_name"].startswith("CP__")]
    assert len(copies) == 2 and all(v["records_out"] == 12 for v in copies)


def test_cluster_runtime_distributor_resizes_consumers(tmp_path):
    """DYN_HASH_PARTITION: the number of consumers of a hash repartitioning is decided from the data
    volume once the upstream stage has finished; the distributor's vertices are held until then and the
    pointwise stages downstream follow the new width (DrDynamicDistributionManager, DrPipelineSplitManager)."""
    from dryad_b200.channel import store
    from dryad_b200.plan import static_config as SC
    ctx = _ctx(tmp_path, n=4, hosts=2, dynamic_opt_level=SC.DYN_BROADCAST | SC.DYN_HASH_PARTITION, dynamic_data_per_vertex=1500)
    uri = f"partfile://{tmp_path}/groups.pt"
    q = ctx.from_enumerable(range(1000), 4).group_by(lambda x: x % 13).select(lambda g: (g.key, g.count())).to_store(uri)
    ctx.submit_and_wait(q)
    st = ctx.last_cluster_status
    assert st["dynamic"]["resized"] >= 1
    info = store.stream_info(uri)
    assert info.partition_count != 4 and info.partition_count == st["tables"][list(st["tables"])[0]]["partitions"]
    assert sorted(ctx.from_store(uri)) == sorted((k, len([x for x in range(1000) if x % 13 == k])) for k in range(13))
    finals = [v for v in st["vertices"] if "Merge_GroupBy" in v["stage_name"]]
    assert len({v["part"] for v in finals}) == info.partition_count
    log = open(os.path.join(str(tmp_path), "ClusterJobs", "1", "calypso.log")).read()
    assert "dynamic,manager=distributor" in log
    # a co-partitioned join must NOT be resized on one side only: the manager leaves it alone
    a = ctx.from_enumerable(range(200), 4)
    b = ctx.from_enumerable([(k, -k) for k in range(0, 200, 2)], 4)
    ctx.broadcast_join_rows = 0
    j = sorted(a.join(b, lambda x: x, lambda t: t[0], lambda x, t: (x, t[1])))
    assert j == [(k, -k) for k in range(0, 200, 2)]


def test_cluster_speculative_duplicate_beats_a_straggler(tmp_path):
    """A vertex that runs far beyond its stage's model gets a duplicate on another host; the first
    version to finish wins (DrDefaultManager / DrStageStatistics; GraphManager/stagemanager/DrStageManager.cpp)."""
    import time

    def straggle(node, part, version):
        if "Select" in node.name and part == 3 and version == 0:
            time.sleep(25)

    ctx = _ctx(tmp_path, n=4, hosts=3, fault_injector=straggle, duplicate_default_seconds=1.0, duplicate_min_seconds=0.5)
    t0 = time.time()
    out = sorted(ctx.from_enumerable(range(40), 4).select(lambda x: x * 3))
    assert out == [x * 3 for x in range(40)]
    assert time.time() - t0 < 20                                  # did not wait for the sleeping attempt
    st = ctx.last_cluster_status
    assert st["counters"]["duplicates"] >= 1
    winners = [v for v in st["vertices"] if "Select" in v["stage_name"] and v["part"] == 3 and v["won"]]
    assert winners and winners[0]["version"] >= 1


def test_cluster_runtime_range_distributor(tmp_path):
    """DYN_RANGE_PARTITION: the range count of an OrderBy follows the data volume — the separator
    vertex computes (new width - 1) separators, the bucketizers get that many ports, the sort stage is
    resized (DrDynamicRangeDistributor)."""
    from dryad_b200.channel import store
    from dryad_b200.plan import static_config as SC
    ctx = _ctx(tmp_path, n=3, hosts=2, dynamic_opt_level=SC.DYN_BROADCAST | SC.DYN_RANGE_PARTITION, dynamic_data_per_vertex=2500)
    data = [(i * 7919) % 5000 for i in range(3000)]
    uri = f"partfile://{tmp_path}/sorted.pt"
    ctx.submit_and_wait(ctx.from_enumerable(data, 3).order_by(lambda x: x).to_store(uri))
    st = ctx.last_cluster_status
    assert st["dynamic"]["resized"] >= 1
    n = store.stream_info(uri).partition_count
    assert n > 3
    parts = store.open_table(uri).partitions
    assert [x for p in parts for x in p] == sorted(data)          # globally sorted across the NEW partitions
    assert sum(1 for p in parts if p) >= 3
    # operators that address their neighbours' partitions by planned position (Reverse, Zip, SlidingWindow,
    # Concat) pin the width of the stage below them: no resize there, and the result stays right
    assert list(ctx.from_enumerable(data, 3).order_by(lambda x: x).reverse()) == sorted(data, reverse=True)
    assert ctx.last_cluster_status["dynamic"]["resized"] == 0


def test_cluster_aggregation_tree_for_decomposed_group_by(tmp_path):
    """GroupBy with decomposable aggregates: consumer j of the hash edge gets internal vertices that
    fold port j of a group of producers with the GroupBy's own combiner (GroupByDynamicReduce)."""
    from dryad_b200.plan import static_config as SC
    ctx = _ctx(tmp_path, n=8, hosts=2, dynamic_opt_level=SC.DYN_BROADCAST | SC.DYN_AGGREGATE, aggregation_tree_fan_in=3)
    ctx.num_partitions = 2
    q = ctx.from_enumerable(range(2000), 8).hash_partition(lambda x: x, 8) \
           .group_by(lambda x: x % 13).select(lambda g: (g.key, g.count(), g.sum()))
    out = sorted(q)
    assert out == sorted((k, len(range(k, 2000, 13)), sum(range(k, 2000, 13))) for k in range(13))
    st = ctx.last_cluster_status
    assert st["dynamic"]["aggregate"] >= 2
    internal = [v for v in st["vertices"] if v["stage_name"].startswith("Agg__")]
    assert internal and all("GroupBy" in v["stage_name"] for v in internal)


def test_cluster_channels_gzip_and_custom_records(tmp_path):
    """Intermediate channels honour the compression setting; records the wire format cannot carry
    (a user class without a registered serializer) still travel between vertex hosts."""
    class P:                                               # local class: pickled by value with the job
        def __init__(self, k, v):
            self.k, self.v = k, v

    ctx = _ctx(tmp_path, n=3, hosts=2, intermediate_data_compression_scheme="Gzip")
    q = ctx.from_enumerable(range(3000), 3).select(lambda x: (x % 10, "v" * 40)).hash_partition(lambda t: t[0], 3) \
           .select(lambda t: t[0])
    assert sorted(q) == sorted(x % 10 for x in range(3000))
    chans = os.path.join(str(tmp_path), "ClusterJobs", "1", "channels")
    sizes = [os.path.getsize(os.path.join(chans, f)) for f in os.listdir(chans) if f.endswith(".chan")]
    assert sizes and max(sizes) < 3000 * 20                # 44-byte records, highly repetitive: gzip shrinks them
    q2 = ctx.from_enumerable(range(60), 3).select(lambda x: P(x % 4, x)).hash_partition(lambda p: p.k, 2) \
            .select(lambda p: (p.k, p.v))
    assert sorted(q2) == sorted((x % 4, x) for x in range(60))
    chans2 = os.path.join(str(tmp_path), "ClusterJobs", "2", "channels")
    assert any(f.endswith(".pkl") for f in os.listdir(chans2))
    # a user class of an importable module is serialized automatically, field by field, in the block-stream format
    import serialization_types as T
    q3 = ctx.from_enumerable([T.Point(i, i * 0.5) for i in range(40)], 3).hash_partition(lambda p: p.x, 2).select(lambda p: T.Point(p.x + 1, p.y))
    assert sorted(q3, key=lambda p: p.x) == [T.Point(i + 1, i * 0.5) for i in range(40)]
    chans3 = os.path.join(str(tmp_path), "ClusterJobs", "3", "channels")
    assert os.listdir(chans3) and not any(f.endswith(".pkl") for f in os.listdir(chans3))


def test_cluster_subgraph_vertex_streams_over_fifos(tmp_path):
    """UseMemoryFIFO: the inputs of a join stream through in-process FIFOs, so producer and consumer
    vertices form a start clique that runs as ONE sub-graph vertex on one host; a member's failure
    restarts the gang (DryadSubGraphVertex, channelfifo.cpp, DrClique.h, DrCohort.cpp:852-929)."""
    def inject(node, part, version):
        if "Join" in node.name and part == 1 and version == 0:
            raise RuntimeError("injected failure inside the gang")

    ctx = _ctx(tmp_path, n=3, hosts=2, use_memory_fifo=True, broadcast_join_rows=0, fault_injector=inject)
    a = ctx.from_enumerable(range(300), 3).select(lambda x: x * 2)
    b = ctx.from_enumerable([(k, str(k)) for k in range(0, 600, 3)], 3)
    out = sorted(a.join(b, lambda x: x, lambda t: t[0], lambda x, t: (x, t[1])))
    assert out == [(k, str(k)) for k in range(0, 600, 6)]
    st = ctx.last_cluster_status
    log = open(os.path.join(str(tmp_path), "ClusterJobs", "1", "calypso.log")).read()
    assert "gang=3" in log                                          # merge(outer)[i] + merge(inner)[i] + join[i] start together
    assert "injected failure inside the gang" in log
    # the gang of partition 1 ran twice, and its members did not leave channel files for the streamed edges
    joins = [v for v in st["vertices"] if "Join" in v["stage_name"] and v["part"] == 1 and v["won"]]
    assert joins and joins[0]["version"] >= 1
    merges = [v for v in st["vertices"] if v["stage_name"] == "Merge" and v["part"] == 1 and v["won"]]
    assert merges and all(v["version"] >= 1 for v in merges)
    merge_uids = {v["stage"] for v in st["vertices"] if v["stage_name"] == "Merge"}
    chans = os.listdir(os.path.join(str(tmp_path), "ClusterJobs", "1", "channels"))
    assert not [f for f in chans if int(f.split("_")[0]) in merge_uids]


def test_cluster_graph_manager_options_and_identity_map(tmp_path):
    """The graph-manager process has its own option table (DryadLinqApplication.cs:34-409): -mai bounds the
    inputs of an aggregation vertex; the identity map records plan vertex -> graph stage."""
    import subprocess, sys
    from dryad_b200.api.submission import ClusterJobSubmission
    from dryad_b200.plan import static_config as SC
    from dryad_b200.runtime import jobpackage
    from dryad_b200.sched.app import DryadLinqApp
    ctx = DryadLinqContext(8, work_dir=str(tmp_path))
    ctx.dynamic_opt_level = SC.DYN_BROADCAST | SC.DYN_AGGREGATE
    q = ctx.from_enumerable(range(800), 8).select(lambda x: x + 1).count_as_query().to_store(f"partfile://{tmp_path}/n.pt")
    job_dir = str(tmp_path / "job")
    jobpackage.write(job_dir, ctx, [q.materialize_lazy()])
    sub = ClusterJobSubmission(job_dir, workers=2)
    sub.add_job_option("mai", 2)
    sub.submit_job()
    assert sub.wait_for_completion(120).value == "Success", sub.error_text()
    st = sub.status_record()
    assert st["dynamic"]["aggregate"] >= 4                      # 8 partials, 2 inputs per internal vertex
    assert list(ctx.from_store(f"partfile://{tmp_path}/n.pt")) == [800]
    ident = DryadLinqApp.read_identity_map(os.path.join(job_dir, "DryadLinqProgram__0.xml.map"))
    assert ident and all(isinstance(v[0], int) for v in ident.values())
    bad = subprocess.run([sys.executable, "-m", "dryad_b200.sched.jm", job_dir, "-nosuchoption"], capture_output=True, text=True)
    assert bad.returncode == 2 and "maxaggregateinputs" in bad.stderr


def test_cluster_remote_channel_reads_over_http(tmp_path):
    """Channels that are not in the reading host's own cache are fetched from the process service's
    file server with ranged GETs (managedchannel HttpReader, ProcessService/HttpServer.cs ranged file GET)."""
    import time

    def slow_producers(node, part, version):
        if "HashPartition" in node.name:
            time.sleep(0.4)                      # long enough for the second host to take its share of the stage

    ctx = _ctx(tmp_path, n=4, hosts=2, remote_channel_reads=True, fault_injector=slow_producers)
    out = sorted(ctx.from_enumerable(range(4000), 4).select(lambda x: (x % 9, x)).group_by(lambda t: t[0])
                 .select(lambda g: (g.key, g.count())))
    assert out == sorted((k, len(range(k, 4000, 9))) for k in range(9))
    st = ctx.last_cluster_status
    assert sum(v["remote_reads"] for v in st["vertices"]) > 0
    assert any(v["cache_hits"] > 0 for v in st["vertices"])          # what a host produced itself never crosses HTTP


def test_cluster_operator_sweep_matches_oracle(tmp_path):
    """The operators whose vertices read neighbouring partitions or several ports (Zip, SlidingWindow,
    Reverse, Concat, Fork, Take/Skip, indexed Select, set operations, Distinct, Apply, GroupJoin, OrderBy
    descending, scalar aggregates) through vertex-host processes and file channels, against LocalDebug."""
    data = [(i * 37) % 101 for i in range(500)]
    other = [(i * 13) % 71 for i in range(300)]

    def build(c):
        a = c.from_enumerable(data, 4)
        b = c.from_enumerable(other, 3)
        qs = {
            "zip": a.zip(c.from_enumerable(list(range(500)), 3), lambda x, y: x * 1000 + y),
            "window": a.sliding_window(lambda w: sum(w), 3),
            "reverse": a.reverse(),
            "concat": a.concat(b),
            "take": a.take(17),
            "skip": a.skip(480),
            "indexed": a.select_indexed(lambda x, i: x + i),
            "union": a.union(b),
            "intersect": a.intersect(b),
            "except": a.except_(b),
            "distinct": a.distinct(),
            "apply": a.apply(lambda xs: [x for x in xs if x % 5 == 0]),
            "groupjoin": a.distinct().group_join(b, lambda x: x, lambda y: y, lambda x, ys: (x, len(list(ys)))),
            "orderdesc": a.order_by_descending(lambda x: x),
            "where_sel": a.where(lambda x: x % 3 == 1).select(lambda x: x * 2),
            "selmany": a.select_many(lambda x: [x, -x] if x % 10 == 0 else []),
        }
        return qs

    ctx = _ctx(tmp_path, n=4, hosts=3)
    oracle = DryadLinqContext(4, local_debug=True)
    mine, ref = build(ctx), build(oracle)
    ctx.submit_and_wait(*mine.values())                       # one job, sixteen outputs
    ordered = {"zip", "window", "reverse", "concat", "take", "skip", "indexed", "orderdesc", "where_sel", "selmany", "apply"}
    for name in mine:
        got, want = list(mine[name]), list(ref[name])
        if name in ordered:
            assert got == want, name
        else:
            assert sorted(got) == sorted(want), name
    # scalar aggregates: each is its own small job
    a = ctx.from_enumerable(data, 4)
    assert a.count() == 500 and a.sum() == sum(data) and a.min() == min(data) and a.max() == max(data)
    assert a.any(lambda x: x == 100) == (100 in data) and a.all(lambda x: x >= 0)
    assert a.first() == data[0] and a.last() == data[-1] and a.contains(data[7])
    assert abs(a.average() - sum(data) / len(data)) < 1e-9


def test_cluster_api_errors_keep_their_code(tmp_path):
    from dryad_b200.api.errors import ErrorCode
    ctx = _ctx(tmp_path, n=2, hosts=1, max_failures_per_vertex=1)
    with pytest.raises(DryadLinqException) as e:
        ctx.from_enumerable([], 2).select(lambda x: x).min()
    assert e.value.error_code == ErrorCode.MinNoElements
    with pytest.raises(DryadLinqException) as e:
        ctx.from_enumerable([1, 2, 3], 2).apply(lambda xs: [1 // 0 for _ in xs]).count()
    assert e.value.error_code == ErrorCode.FailureInUserApplyFunction


def test_cluster_submit_is_asynchronous_and_cancellable(tmp_path):
    """Submit returns a job handle at once; Wait blocks; CancelJob stops the graph manager and its hosts."""
    import time
    from dryad_b200.api.context import JobStatus

    def slow(node, part, version):
        time.sleep(20)

    ctx = _ctx(tmp_path, n=2, hosts=2)
    job = ctx.submit(ctx.from_enumerable(range(100), 2).select(lambda x: x + 1).to_store(f"partfile://{tmp_path}/fast.pt"))
    assert job.status in (JobStatus.Waiting, JobStatus.Running, JobStatus.Success)
    res = job.wait()
    assert job.status == JobStatus.Success and sorted(res[0]) == list(range(1, 101))
    ctx2 = _ctx(tmp_path / "b", n=2, hosts=2, fault_injector=slow)
    t0 = time.time()
    job2 = ctx2.submit(ctx2.from_enumerable(range(10), 2).select(lambda x: x).to_store(f"partfile://{tmp_path}/slow.pt"))
    assert time.time() - t0 < 5                          # did not wait for the 20 s vertices
    time.sleep(3.0)
    job2.cancel_job()
    assert job2.status == JobStatus.Cancelled
    with pytest.raises(DryadLinqException):
        job2.wait()
    assert not os.path.exists(tmp_path / "slow.pt")


def test_cluster_job_package_references_stored_tables(tmp_path):
    """A stored input travels as (URI, byte ranges): the package stays small and every vertex host opens
    only the partitions it runs (the reference hands input channel URIs to the vertices)."""
    text = tmp_path / "big.txt"
    line = "alpha beta gamma delta alpha\n"
    text.write_text(line * 40000)                                   # ~1.1 MB
    ctx = _ctx(tmp_path, n=4, hosts=2)
    lines = ctx.from_store(f"file://{text}")
    assert lines.expr.params["partition_count"] == 4                # cut into byte ranges at line boundaries
    wc = dict(lines.select_many(lambda l: l.line.split(" ")).group_by(lambda w: w).select(lambda g: (g.key, g.count())))
    assert wc == {"alpha": 80000, "beta": 40000, "gamma": 40000, "delta": 40000}
    pkg = os.path.join(str(tmp_path), "ClusterJobs", "1", "job.pkl")
    assert os.path.getsize(pkg) < 200_000                           # not the megabyte of text
    uri = f"partfile://{tmp_path}/nums.pt"
    ctx.submit_and_wait(ctx.from_enumerable(range(50000), 4).to_store(uri))
    assert ctx.from_store(uri).select(lambda x: x % 7).sum() == sum(x % 7 for x in range(50000))
    pkg3 = os.path.join(str(tmp_path), "ClusterJobs", "3", "job.pkl")
    assert os.path.getsize(pkg3) < 100_000


def test_cluster_failed_vertex_can_be_rerun_alone(tmp_path):
    """A finally-failed vertex of a cluster job leaves its start command; the job directory has the vertex
    code (package) and its inputs (channel files), so the vertex replays alone, in one process
    (LocalDebuggingAndProfiling.cs, DumpRestartCommand)."""
    import subprocess
    import sys
    from dryad_b200.tools import rerun
    flag = str(tmp_path / "fixed")

    def inject(node, part, version):
        if "GroupBy" in node.name and "Merge" in node.name and part == 1 and not os.path.exists(flag):
            raise ValueError("bug in the user's function")

    ctx = _ctx(tmp_path, n=3, hosts=2, fault_injector=inject, max_failures_per_vertex=2)
    with pytest.raises(DryadLinqException):
        ctx.from_enumerable(range(300), 3).group_by(lambda x: x % 6).select(lambda g: (g.key, g.count())).to_list()
    job_dir = os.path.join(str(tmp_path), "ClusterJobs", "1")
    found = rerun.list_restarts(job_dir)
    assert len(found) == 1 and found[0]["partition"] == 1 and found[0]["inputs"]
    tag = f"{found[0]['stage']}.1"
    res = rerun.replay_cluster(job_dir, tag)
    assert not res["ok"] and "bug in the user's function" in res["message"]          # reproduces the failure
    open(flag, "w").close()                                                          # "fix the bug"
    res = rerun.replay_cluster(job_dir, tag)
    assert res["ok"] and res["n_in"] > 0 and res["n_out"] == 2                       # keys 1 and ... of 6 land in partition 1 of 3
    p = subprocess.run([sys.executable, "-m", "dryad_b200.tools.rerun", job_dir, tag, "--run"], capture_output=True, text=True,
                       cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert p.returncode == 0 and "completed" in p.stdout, p.stdout + p.stderr


def test_context_from_a_cluster_descriptor(tmp_path):
    """DryadLinqContext(DryadLinqCluster) (DryadLinqContext.cs:672): platform, sizes and the place for
    job-internal data come from the descriptor."""
    import dryad_b200 as D
    cl = D.DryadLinqLocalCluster(hosts=2, work_dir=str(tmp_path), partitions=3)
    ctx = D.DryadLinqContext(cluster=cl)
    assert ctx.platform == D.PlatformKind.LOCAL_CLUSTER and ctx.num_partitions == 3 and ctx.job_max_nodes == 2
    assert cl.MakeInternalClusterUri("staging", "x.pt") == f"partfile://{tmp_path}/internal/staging/x.pt"
    assert sorted(ctx.from_enumerable(range(30), 3).select(lambda x: x + 1)) == list(range(1, 31))
    assert ctx.last_cluster_status["hosts"] == 2
    assert ctx.ClientVersion() == ctx.version
    ctx.RegisterAzureAccount("acct", "k3y")
    assert ctx.AzureAccountKey("acct") == "k3y" and ctx.AzureAccountKey("other") is None
    b = D.DryadLinqB200Cluster(gpus=8)
    assert b.kind == D.PlatformKind.B200


def test_cluster_hosts_that_cannot_start_fail_the_job_with_their_log(tmp_path):
    """A vertex host that dies again and again (here: a broken environment makes it exit at start) ends the
    job with the tail of its log, instead of restarting forever."""
    ctx = _ctx(tmp_path, n=2, hosts=1)
    ctx.job_environment_variables["PYTHONPATH"] = ""            # the hosts still find the framework: the manager adds its own root

    def inject(node, part, version):
        os._exit(9)                                              # every attempt takes the host down

    ctx.fault_injector = inject
    ctx.max_failures_per_vertex = 50
    with pytest.raises(DryadLinqException) as e:
        ctx.from_enumerable(range(10), 2).select(lambda x: x).to_list()
    assert "died" in str(e.value) and "exit code 9" in str(e.value)


def test_cluster_subgraph_member_failing_before_it_writes_does_not_hang(tmp_path):
    """Every producer of a gang fails before it touched its FIFO: the consumer must see the abort (the FIFOs
    exist before any member runs) and the gang restarts."""
    import time

    def inject(node, part, version):
        if node.name == "Merge" and version == 0:
            raise RuntimeError("producer fails first")

    ctx = _ctx(tmp_path, n=3, hosts=2, use_memory_fifo=True, broadcast_join_rows=0, fault_injector=inject)
    t0 = time.time()
    a = ctx.from_enumerable(range(60), 3)
    b = ctx.from_enumerable(range(0, 120, 2), 3)
    assert sorted(a.except_(b)) == [x for x in range(60) if x % 2 == 1]
    assert time.time() - t0 < 60
    assert ctx.last_cluster_status["counters"]["failures"] >= 3


def test_cluster_vertex_timeout_kills_and_reruns_a_hung_vertex(tmp_path):
    import time

    def hang(node, part, version):
        if "Select" in node.name and part == 1 and version == 0:
            time.sleep(120)

    ctx = _ctx(tmp_path, n=2, hosts=2, fault_injector=hang, vertex_timeout_seconds=2, enable_speculative_duplication=False)
    t0 = time.time()
    assert sorted(ctx.from_enumerable(range(20), 2).select(lambda x: x * 2)) == [x * 2 for x in range(20)]
    assert time.time() - t0 < 60
    st = ctx.last_cluster_status
    assert st["host_restarts"] >= 1 and st["counters"]["failures"] >= 1
    log = open(os.path.join(str(tmp_path), "ClusterJobs", "1", "calypso.log")).read()
    assert "state=killed" in log


def test_cluster_intermediate_data_is_removed_when_not_retained(tmp_path):
    ctx = _ctx(tmp_path, n=3, hosts=2, retain_intermediate=False)
    assert sorted(ctx.from_enumerable(range(90), 3).hash_partition(lambda x: x, 3).select(lambda x: x + 1)) == list(range(1, 91))
    job_dir = os.path.join(str(tmp_path), "ClusterJobs", "1")
    assert not os.path.exists(os.path.join(job_dir, "channels")) and os.path.exists(os.path.join(job_dir, "calypso.log"))
