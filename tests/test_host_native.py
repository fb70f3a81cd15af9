"""Native host runtime (libdryad_host.so): graph manager state machine, straggler model, dynamic
managers, message pump, mailbox, delay scheduler, FIFO, cache/spill, fingerprint, record streams."""
import os
import threading
import time

import pytest

from dryad_b200.channel import records as R
from dryad_b200.channel.fifo import BlockCache, Fifo, fingerprint
from dryad_b200.runtime.services import LocalScheduler, Mailbox, MessagePump
from dryad_b200.sched import gm as G


def _drain(gm, work=lambda v, ver: None, now=0.0):
    order = []
    while True:
        nxt = gm.next_ready(now)
        if nxt is None:
            break
        v, ver = nxt
        work(v, ver)
        order.append((v, ver))
        gm.completed(v, ver, 0.01, 100.0)
    return order


def test_gm_runs_dag_in_dependency_order_and_finishes():
    gm = G.GraphManager()
    a = gm.add_stage("in", 3)
    b = gm.add_stage("part", 3)
    c = gm.add_stage("merge", 2)
    d = gm.add_stage("out", 1)
    gm.connect(a, b, G.POINTWISE)
    gm.connect(b, c, G.CROSS_PRODUCT)
    gm.connect(c, d, G.POINTWISE)
    gm.start()
    order = _drain(gm)
    assert [v for v, _ in order] == list(range(9)) and all(ver == 0 for _, ver in order)
    assert gm.state == "Succeeded" and gm.counters()["completed"] == 9


def test_gm_retries_then_aborts_after_max_failures():
    gm = G.GraphManager(max_failures=3)
    s = gm.add_stage("s", 1)
    gm.start()
    for attempt in range(3):
        v, ver = gm.next_ready(0.0)
        assert ver == attempt                     # every attempt is a new version
        rc = gm.failed(v, ver, "boom")
        assert rc == (1 if attempt == 2 else 0)
    assert gm.state == "Failed" and "failed 3 times" in gm.error()


def test_gm_read_error_invalidates_and_reruns_the_producer():
    gm = G.GraphManager()
    p = gm.add_stage("producer", 1)
    c = gm.add_stage("consumer", 1)
    gm.connect(p, c)
    gm.start()
    v, ver = gm.next_ready(0.0)
    gm.completed(v, ver, 0.1)
    cv, cver = gm.next_ready(0.0)
    assert gm.vertex(cv)["stage"] == c
    assert gm.failed(cv, cver, "cannot read input", bad_input=0) == 2
    v2, ver2 = gm.next_ready(0.0)                 # the producer runs again as version 1
    assert v2 == v and ver2 == 1 and gm.counters()["invalidations"] == 1
    gm.completed(v2, ver2, 0.1)
    cv2, cver2 = gm.next_ready(0.0)
    assert cv2 == cv and cver2 == 1
    gm.completed(cv2, cver2, 0.1)
    assert gm.state == "Succeeded"


def test_gm_speculative_duplicate_first_finisher_wins():
    gm = G.GraphManager()
    gm.set_outlier_params(default_s=5.0, min_s=0.1)
    s = gm.add_stage("s", 2)
    gm.start()
    v0, ver0 = gm.next_ready(0.0)
    v1, ver1 = gm.next_ready(0.0)
    gm.completed(v1, ver1, 1.0)
    assert gm.check_duplicates(now=3.0) == 0      # below the threshold
    assert gm.check_duplicates(now=6.0) == 1      # straggler: duplicate requested
    dv, dver = gm.next_ready(6.0)
    assert dv == v0 and dver == 1
    assert gm.completed(dv, dver, 0.5) is True    # the duplicate wins...
    assert gm.completed(v0, ver0, 9.0) is False   # ...and the original loses
    assert gm.state == "Succeeded" and gm.counters()["duplicates"] == 1


def test_stage_statistics_fit_and_outlier_threshold():
    gm = G.GraphManager()
    gm.set_outlier_params(default_s=600.0, min_s=0.5)
    s = gm.add_stage("s", 40)
    gm.start()
    for i in range(30):
        v, ver = gm.next_ready(0.0)
        b = 1e6 * (1 + i % 5)
        t = 2.0 + 3e-6 * b + (0.01 if i % 2 else -0.01)
        if i == 7:
            t = 500.0                              # one gross outlier must not wreck the fit
        gm.completed(v, ver, t, b)
    m = gm.stage_model(s, 3e6)
    assert m["have_model"]
    assert abs(m["startup"] - 2.0) < 0.5 and abs(m["per_byte"] - 3e-6) < 1e-6
    assert m["outlier_threshold"] < 600.0


def test_dynamic_managers():
    groups = G.aggregate_groups([100] * 10, machine=[0, 0, 1, 1, 2, 2, 3, 3, 4, 4], pod=[0] * 5 + [1] * 5,
                                max_inputs=4, max_bytes=10_000)
    assert len(set(groups)) == 4                  # <=4 inputs per group, never across pods
    assert groups[0] != groups[9]
    big = G.aggregate_groups([10, 10**12, 10], [0, 0, 0], [0, 0, 0])
    assert big[1] not in (big[0], big[2])         # oversize inputs are left alone
    assert G.consumer_count(10 << 30, 1 << 30) == 10 and G.consumer_count(5, 1 << 30) == 1
    copies, assign = G.broadcast_tree(16)
    assert copies == 4 and sorted(set(assign)) == [0, 1, 2, 3]
    assert G.broadcast_tree(4)[0] == 0


def test_message_pump_delivers_under_listener_lock_and_delays():
    seen, lock_violations, active = [], [], {}
    mu = threading.Lock()

    def handler(listener, message):
        with mu:
            if active.get(listener):
                lock_violations.append(listener)
            active[listener] = True
        time.sleep(0.001)
        with mu:
            seen.append((listener, message, time.monotonic()))
            active[listener] = False

    pump = MessagePump(handler, threads=4)
    t0 = time.monotonic()
    for i in range(50):
        pump.post(i % 3, i)
    pump.post(7, 999, delay=0.2)
    assert pump.wait_delivered(51, timeout=5.0) == 51
    pump.close()
    assert not lock_violations and len(seen) == 51
    late = [t for l, m, t in seen if m == 999][0]
    assert late - t0 >= 0.19


def test_mailbox_versions_and_long_poll():
    mb = Mailbox()
    assert mb.block_on_status("cmd", 0, timeout=0.05) is None
    v1 = mb.set("cmd", "start", b"payload-1")
    got = mb.block_on_status("cmd", 0, timeout=0.1)
    assert got == (v1, "start", b"payload-1")
    assert mb.block_on_status("cmd", v1, timeout=0.05) is None      # nothing newer
    threading.Timer(0.1, lambda: mb.set("cmd", "terminate", b"x")).start()
    got = mb.block_on_status("cmd", v1, timeout=2.0)
    assert got is not None and got[0] == v1 + 1 and got[1] == "terminate"
    mb.close()


def test_delay_scheduler_locality_then_rack_then_cluster():
    ls = LocalScheduler(rack_of=[0, 0, 1, 1], rack_delay=1.0, cluster_delay=2.0)
    ls.submit(1, [0])
    ls.submit(2, [3], hard=True)
    assert ls.computer_idle(2) is None                  # no affinity, nothing aged yet
    assert ls.computer_idle(1, age_bias=1.5) == 1       # same rack as computer 0 after the rack delay
    assert ls.computer_idle(0, age_bias=10.0) is None   # hard constraint never relaxes
    assert ls.computer_idle(3) == 2
    ls.submit(3, [0])
    assert ls.computer_idle(2, age_bias=1.5) is None    # other rack: needs the cluster delay
    assert ls.computer_idle(2, age_bias=2.5) == 3
    assert ls.waiting == 0
    ls.close()


def test_fifo_blocking_depth_and_end_of_stream():
    f = Fifo(depth=2)
    assert f.write(b"a") and f.write(b"b")
    assert f.write(b"c", timeout=0.05) is False         # full: blocking writer times out
    got = []
    t = threading.Thread(target=lambda: [got.append(f.read()) for _ in range(4)])
    t.start()
    assert f.write(b"c", timeout=2.0)
    f.close_writer()
    t.join()
    assert got == [b"a", b"b", b"c", None]
    f.close()


def test_cache_spills_and_reloads(tmp_path):
    c = BlockCache(str(tmp_path), low_water=3000, high_water=5000)
    for i in range(8):
        c.put(f"blk{i}", bytes([i]) * 1000)
    st = c.stats()
    assert st["spills"] >= 3 and st["resident"] <= 5000
    for i in range(8):
        assert c.get(f"blk{i}") == bytes([i]) * 1000
    assert c.stats()["loads"] >= 3 and c.get("nope") is None
    c.close()


def test_fingerprint_properties():
    a = fingerprint(b"hello world")
    assert a == fingerprint(b"hello world") and a != fingerprint(b"hello worle") and a != fingerprint(b"hello worl")
    assert fingerprint(b"") == 0 and fingerprint(b"abc", seed=5) != fingerprint(b"abc")


def test_record_codec_primitives_and_roundtrip(tmp_path):
    import ctypes as C
    from dryad_b200 import _native as N
    from dryad_b200.api.types import Grouping, LineRecord, Pair, Record
    assert R.compact(0) == b"\x00" and R.compact(0x7F) == b"\x7f"
    assert R.compact(0x80) == b"\x80\x00\x00\x80" and R.compact(0x12345678) == b"\x92\x34\x56\x78"
    assert R.enc_string("hé") == b"\x02\x03h\xc3\xa9"
    lib = N.host()
    p = str(tmp_path / "prim.bin")
    w = lib.dh_writer_open(p.encode(), 0, 0)
    lib.dh_write_compact(w, 0x80)
    lib.dh_write_i64(w, -2)
    lib.dh_write_string(w, "hé".encode(), 3, 2)
    lib.dh_writer_end_record(w)
    lib.dh_writer_close(w, None, None, None)
    raw = open(p, "rb").read()[16:]                       # skip the block header
    assert raw == R.compact(0x80) + (-2).to_bytes(8, "little", signed=True) + R.enc_string("hé")
    objs = [None, True, -5, 2**70, 1.5, "x", b"\x00\x01", (1, "a"), [1, [2]], Record(a=1, b="z"), Pair("k", 3),
            Grouping("g", [1, 2]), LineRecord("line"), {"k": (1, 2)}]
    for o in objs:
        buf = memoryview(R.encode_record(o))
        back, pos = R.decode(buf, 0)
        assert back == o and pos == len(buf) and type(back) is type(o)


def test_block_stream_never_splits_records_and_gzip(tmp_path):
    from dryad_b200.channel import store
    recs = [("r%d" % i, "x" * (i % 700)) for i in range(3000)]     # > one 256 KiB block
    for comp in (None, "Gzip"):
        uri = f"partfile://{tmp_path}/t_{comp}.pt"
        store.write_table(uri, [recs[:1000], recs[1000:]], None, comp)
        t = store.open_table(uri)
        assert [len(p) for p in t.partitions] == [1000, 2000]
        assert t.partitions[0] + t.partitions[1] == recs
    raw = os.path.getsize(f"{tmp_path}/t_None.00000001")
    gz = os.path.getsize(f"{tmp_path}/t_Gzip.00000001")
    assert gz < raw / 5


def test_vertex_command_wire_format_round_trip():
    from dryad_b200.runtime.command import CMD_START, STATE_COMPLETED, VertexCommand, affinity_merge
    c = VertexCommand(CMD_START, vertex_id=17, version=2, args=["--flag", "Select__3"],
                      inputs=[("file:///tmp/3_0_1.tmp", 1234), ("fifo://32/4_1_0", -1)], outputs=[("hbm://17_0_2", 1 << 30)])
    wire = c.serialize()
    back = VertexCommand.parse(wire)
    assert back == c
    st = VertexCommand(0, 17, 2, state=STATE_COMPLETED, exit_code=0, message="ok", channel_status=[(1234, 0), (99, 5)])
    assert VertexCommand.parse(st.serialize()) == st
    assert VertexCommand.parse(wire[:-3]) is None and VertexCommand.parse(b"garbage") is None
    assert affinity_merge([(0, 900), (1, 50), (2, 50)], rack_of=[0, 0, 1]) == ("computer", 0)
    assert affinity_merge([(0, 400), (1, 350), (2, 250)], rack_of=[0, 0, 1]) == ("rack", 0)
    assert affinity_merge([(0, 300), (2, 300), (4, 400)], rack_of=[0, 0, 1, 1, 2, 2])[0] == "cluster"


def test_channel_factory_file_fifo_null(tmp_path):
    from dryad_b200.channel.factory import ChannelReader, ChannelWriter, channel_name, fifo_uri
    recs = [("k%d" % i, i, [i, i + 1]) for i in range(5000)]
    uri = "file://" + str(tmp_path / channel_name(3, 0, 1))
    w = ChannelWriter(uri)
    for r in recs:
        w.write(r)
    w.close()
    assert list(ChannelReader(uri)) == recs and w.records == 5000
    fu = fifo_uri(9, 0, 0)
    got = []
    t = threading.Thread(target=lambda: got.extend(ChannelReader(fu)))
    t.start()
    fw = ChannelWriter(fu)
    for r in recs:
        fw.write(r)
    fw.close()
    t.join()
    assert got == recs
    nw = ChannelWriter("null://x")
    nw.write(1)
    nw.close()
    assert list(ChannelReader("null://x")) == []


def _ready_all(g, now=0.0):
    out = []
    while True:
        x = g.next_ready(now)
        if not x:
            return out
        out.append(x)


def test_start_cliques_gang_start_and_gang_restart():
    """Vertices joined by FIFO edges start together and restart together
    (GraphManager/vertex/DrClique.h:28-64, DrCohort.cpp:852-929)."""
    from dryad_b200.sched.gm import CHANNEL_FIFO, GraphManager
    g = GraphManager()
    src, a, b, sink = g.add_stage("src", 2), g.add_stage("read", 2), g.add_stage("map", 2), g.add_stage("sink", 1)
    g.connect(src, a, 0)                       # file edge: external input of the clique
    g.connect(a, b, 0, CHANNEL_FIFO)           # streamed: read[i] and map[i] form a clique
    g.connect(b, sink, 0)
    g.start()
    assert sorted(g.clique(2)) == [2, 4] and g.clique(6) == [6]
    first = _ready_all(g)
    assert sorted(v for v, _ in first) == [0, 1]            # clique members wait for their external inputs
    g.completed(0, 0, 1.0, 10)
    ready = _ready_all(g)
    assert sorted(v for v, _ in ready) == [2, 4]            # both members of clique 0 at once, not clique 1
    g.completed(1, 0, 1.0, 10)
    assert sorted(v for v, _ in _ready_all(g)) == [3, 5]
    # the consumer of clique 0 fails: the whole gang (2 and 4) runs again, clique 1 is untouched
    assert g.failed(4, 0, "boom") == 0
    assert g.gang_restarts == 1
    again = _ready_all(g)
    assert sorted(again) == [(2, 1), (4, 1)]
    for v, ver in again + [(3, 0), (5, 0)]:
        g.completed(v, ver, 1.0, 10)
    (last,) = _ready_all(g)
    g.completed(*last, 1.0, 1)
    assert g.state == "Succeeded"


def test_dynamic_resize_follows_pointwise_stages_and_range_distributor_sizing():
    """DrDynamicRangeDistributor sizes the consumers from the sample volume; DrPipelineSplitManager
    propagates the new width to the pointwise stages downstream."""
    from dryad_b200.sched.gm import CROSS_PRODUCT, GraphManager, range_consumer_count
    assert range_consumer_count(5 << 20, 0.001, 1 << 30) == 5          # 5 MiB of samples at 0.1 % = 5 GiB
    assert range_consumer_count(0, 0.001, 1 << 30) == 1
    g = GraphManager()
    samp, dist, sort, out = (g.add_stage(n, 4) for n in ("sample", "distribute", "sort", "output"))
    g.connect(samp, dist, 0)
    g.connect(dist, sort, CROSS_PRODUCT)
    g.connect(sort, out, 0)
    g.start()
    for v, ver in _ready_all(g):
        g.completed(v, ver, 1.0, 100)
    assert g.resize_stage(sort, 6) == 2                                 # sort and the pointwise output follow
    assert len(g.stage_vertices(sort)) == 6 and len(g.stage_vertices(out)) == 6 and g.live_vertices == 4 + 4 + 6 + 6
    for v, ver in _ready_all(g):                                            # the distributors
        g.completed(v, ver, 1.0, 100)
    sorts = _ready_all(g)
    assert sorted(v for v, _ in sorts) == sorted(g.stage_vertices(sort))
    for v, ver in sorts:
        g.completed(v, ver, 1.0, 100)
    for v, ver in _ready_all(g):
        g.completed(v, ver, 1.0, 100)
    assert g.state == "Succeeded"
    # a started stage cannot be resized
    g2 = GraphManager()
    s0, s1 = g2.add_stage("a", 2), g2.add_stage("b", 2)
    g2.connect(s0, s1, 0)
    g2.start()
    _ready_all(g2)
    import pytest
    with pytest.raises(ValueError):
        g2.resize_stage(s0, 3)


def test_builtin_and_custom_record_serializers_round_trip():
    """Guid / DateTime / Decimal / ForkValue built-ins and a [CustomDryadLinqSerializer] type
    (LinqToDryad/DryadLinqSerialization.cs:41-789, Attributes.cs:37-275)."""
    import datetime
    import decimal
    import uuid
    import dryad_b200 as D
    from dryad_b200.channel import records as R

    class Money:
        def __init__(self, cents):
            self.cents = cents

        def __eq__(self, o):
            return isinstance(o, Money) and o.cents == self.cents

    R.register_serializer(Money, "test.Money", lambda m: m.cents.to_bytes(8, "little", signed=True),
                          lambda b: Money(int.from_bytes(b, "little", signed=True)))
    rec = D.Record(id=uuid.UUID("12345678-1234-5678-1234-567812345678"), t=datetime.datetime(2014, 3, 1, 12, 30, 15, 250000),
                   d=decimal.Decimal("12.3400"), fv=D.ForkValue((1, "x")), none=D.ForkValue(), m=Money(-995))
    buf = R.encode_record(rec)
    back, pos = R.decode(memoryview(buf), 0)
    assert pos == len(buf) and back == rec and back.m == Money(-995) and not back.none.has_value
    # .NET Guid layout: first three groups little-endian
    assert buf[buf.index(bytes((R.T_GUID,))) + 1:][:4] == bytes.fromhex("78563412")


def test_vertex_env_buffer_sizing_threads_and_error_report(tmp_path, monkeypatch):
    from dryad_b200.runtime.vertex_ops import VertexEnv
    import dryad_b200 as D
    ctx = D.DryadLinqContext(1)
    env = VertexEnv(ctx, 7, 2, 1, n_outputs=4)
    assert env.write_buffer_size(available_physical=1 << 40) == 8 << 20
    assert env.write_buffer_size(available_physical=64 << 20) == (64 << 20) // 4 // 4
    assert env.write_buffer_size(available_physical=100) == 16 << 10
    assert VertexEnv(ctx, 7, 2, use_large_buffer=True).write_buffer_size(available_physical=1 << 40) == 256 << 20
    monkeypatch.setenv("DRYAD_THREADS_PER_WORKER", "3")
    assert env.threads_per_worker == 3
    try:
        raise ValueError("bad record")
    except ValueError as ex:
        path = env.report_vertex_error(ex, str(tmp_path))
    text = open(path).read()
    assert path.endswith("VertexException.txt") and "vertex 7 partition 2 version 1" in text and "bad record" in text


def test_gm_splice_rewrites_a_running_graph():
    """Run-time rewrite (DrDynamicAggregateManager / DrDynamicBroadcast): a new vertex takes over edges
    of vertices that have not started; retries and upstream invalidation keep working through it."""
    from dryad_b200.sched.gm import GraphManager
    gm = GraphManager(max_failures=4, speculative=False)
    a = gm.add_stage("partial", 4)
    b = gm.add_stage("final", 1)
    dyn = gm.add_stage("Agg__final", 0)
    gm.connect(a, b)
    gm.start()
    started = {}
    while True:
        nxt = gm.next_ready(0.0)
        if nxt is None:
            break
        started[nxt[0]] = nxt[1]
    assert sorted(started) == [0, 1, 2, 3]
    gm.completed(0, 0, 0.1)
    gm.completed(1, 0, 0.1)
    x = gm.splice(dyn, [0, 1], [4])                       # both inputs complete: the internal vertex is runnable at once
    assert gm.vertex(x)["stage"] == dyn and gm.vertex(x)["part"] == 0
    assert gm.vertex_inputs(4) == [2, 3, x] and gm.vertex_inputs(x) == [0, 1]
    assert gm.next_ready(0.0) == (x, 0)
    gm.completed(2, 0, 0.1)
    gm.completed(3, 0, 0.1)
    assert gm.next_ready(0.0) is None and gm.vertex(4)["state"] == "Waiting"      # the consumer waits for the new vertex
    with pytest.raises(ValueError):
        gm.splice(dyn, [0], [4])                           # 0 is no longer an input of the consumer
    gm.completed(x, 0, 0.1)
    assert gm.next_ready(0.0) == (4, 0)
    # the consumer cannot read the internal vertex's output: it is revoked and runs again from ITS inputs
    assert gm.failed(4, 0, "cannot read", bad_input=2) == 2
    assert gm.next_ready(0.0) == (x, 1)
    gm.completed(x, 1, 0.1)
    v, ver = gm.next_ready(0.0)
    assert v == 4
    with pytest.raises(ValueError):
        gm.splice(dyn, [2], [4])                           # a started consumer cannot be rewired
    gm.completed(4, ver, 0.1)
    assert gm.state == "Succeeded" and gm.counters()["completed"] == 6
    # copy vertex between one producer and several consumers
    g2 = GraphManager(speculative=False)
    p = g2.add_stage("tee", 1)
    c = g2.add_stage("join", 6)
    cp = g2.add_stage("CP__tee", 0)
    g2.connect(p, c)
    g2.start()
    assert g2.next_ready(0.0) == (0, 0)
    g2.completed(0, 0, 0.1)
    x1 = g2.splice(cp, [0], [1, 2, 3])
    x2 = g2.splice(cp, [0], [4, 5, 6])
    ready = []
    while (n := g2.next_ready(0.0)) is not None:
        ready.append(n[0])
    assert sorted(ready) == [x1, x2]
    g2.completed(x1, 0, 0.1)
    ready = []
    while (n := g2.next_ready(0.0)) is not None:
        ready.append(n[0])
    assert sorted(ready) == [1, 2, 3]


def test_native_record_codec_is_byte_exact_with_the_python_codec():
    """csrc/py/pyrecords.cpp against channel/records.py: same bytes out, same objects back, for the types it
    handles itself and for the ones it hands back (big ints, Grouping, ForkTuple, GUID, DateTime, Decimal,
    subclasses), at any nesting depth."""
    import datetime
    import decimal
    import random
    import uuid
    from dryad_b200.api.types import ForkTuple, ForkValue, Grouping, LineRecord, Pair, Record
    from dryad_b200.channel import records as R
    if R._native is None:
        pytest.skip("native record codec not built")
    rng = random.Random(11)

    def rand(depth=0):
        k = rng.randrange(20 if depth < 3 else 9)
        if k == 0: return None
        if k == 1: return rng.random() < 0.5
        if k == 2: return rng.choice([0, 1, -1, 127, 128, -(1 << 63), (1 << 63) - 1, rng.randrange(-10**12, 10**12)])
        if k == 3: return rng.choice([0.0, -0.0, 1.5, float("inf"), 1e-300, rng.random()])
        if k == 4: return rng.choice(["", "a", "héllo wörld", "x" * 200, "日本語", "\\u0000z"])
        if k == 5: return rng.choice([b"", b"\\x00\\xff", bytes(range(256))])
        if k == 6: return LineRecord("line " + str(rng.randrange(100)))
        if k == 7: return 1 << rng.choice([64, 70, 200])                       # big integer: Python fallback
        if k == 8: return rng.choice([uuid.UUID(int=rng.getrandbits(128)), datetime.datetime(2024, 2, 29, 12, 30, 5, 123456),
                                      decimal.Decimal("12.3400"), ForkValue(3), ForkValue()])
        if k in (9, 10): return tuple(rand(depth + 1) for _ in range(rng.randrange(4)))
        if k in (11, 12): return [rand(depth + 1) for _ in range(rng.randrange(4))]
        if k == 13: return {rng.choice(["a", 1, 2.5, ("t", 1)]): rand(depth + 1) for _ in range(rng.randrange(3))}
        if k in (14, 15): return Record({f"f{i}": rand(depth + 1) for i in range(rng.randrange(1, 4))})
        if k == 16: return Pair(rand(depth + 1), rand(depth + 1))
        if k == 17:
            g = Grouping(rand(depth + 1))
            g.extend(rand(depth + 1) for _ in range(rng.randrange(3)))
            return g
        if k == 18: return ForkTuple(rand(depth + 1), None)
        return Record({"nested": Record({"deep": [1, (2, "x")]})})

    recs = [rand() for _ in range(3000)]
    for r in recs:
        assert R.encode_record(r) == R._encode_record_py(r), r
    data, offs, n = R.encode_many(recs)
    assert n == len(recs) and data == b"".join(R._encode_record_py(r) for r in recs)
    back = R.decode_many(data)
    mv, pos, want = memoryview(data), 0, []
    while pos < len(mv):
        o, pos = R.decode(mv, pos)
        want.append(o)
    assert len(back) == len(want) == len(recs)
    for a, b in zip(back, want):
        assert type(a) is type(b) and repr(a) == repr(b)
    # truncated and malformed streams raise instead of reading out of bounds
    for cut in (1, 5, len(data) // 2, len(data) - 1):
        try:
            R.decode_many(data[:cut])
        except Exception:  # noqa: BLE001
            pass
    with pytest.raises(Exception):
        R.decode_many(bytes([7, 0xFF, 0xFF, 0xFF, 0xFF]))                          # a tuple that claims 2^31 items
