"""Cluster abstraction, node daemon over HTTP, local multi-process job submission."""
import os
import sys
import urllib.request

from dryad_b200.api.context import JobStatus
from dryad_b200.api.submission import LocalJobSubmission
from dryad_b200.runtime.process_service import H_STATUS, H_VERSION, ProcessService
from dryad_b200.sched.cluster import Affinity, IProcessWatcher, LocalCluster, Process

HERE = os.path.dirname(os.path.abspath(__file__))


def test_local_cluster_matches_processes_to_idle_gpus():
    events = []

    class W(IProcessWatcher):
        def on_queued(self, p): events.append(("queued", p.id))
        def on_matched(self, p, c): events.append(("matched", p.id, c.name))

    cl = LocalCluster(4, "/tmp")
    assert [c.name for c in cl.get_computers()] == ["gpu0", "gpu1", "gpu2", "gpu3"]
    cl.schedule_process(Process(1, "worker --vertex 1", Affinity([2], 100), W()))
    cl.schedule_process(Process(2, "worker --vertex 2", Affinity([3], 100, hard=True), W()))
    assert cl.computer_idle(0) is None                     # nothing prefers gpu0 yet
    assert cl.computer_idle(2).id == 1
    assert cl.computer_idle(0, age_bias=5.0) is None       # hard constraint on gpu3 never relaxes
    assert cl.computer_idle(3).id == 2
    assert ("matched", 1, "gpu2") in events and cl.local_path(cl.computers[1], "3_0_1.tmp") == "file:///tmp/3_0_1.tmp?c=None"


def test_process_service_mailbox_long_poll_and_ranged_file_get(tmp_path):
    svc = ProcessService(str(tmp_path))
    try:
        (tmp_path / "chan.tmp").write_bytes(bytes(range(200)))
        req = urllib.request.Request(f"{svc.uri}/process/p1/setvalue?key=cmd&status=start", data=b"launch-descriptor", method="PUT")
        with urllib.request.urlopen(req) as r:
            v1 = int(r.headers[H_VERSION])
        with urllib.request.urlopen(f"{svc.uri}/process/p1/status?key=cmd&version=0&timeout=1") as r:
            assert r.status == 200 and int(r.headers[H_VERSION]) == v1 and r.headers[H_STATUS] == "start"
            assert r.read() == b"launch-descriptor"
        with urllib.request.urlopen(f"{svc.uri}/process/p1/status?key=cmd&version={v1}&timeout=0.1") as r:
            assert r.status == 204                         # nothing newer: long poll timed out
        req = urllib.request.Request(f"{svc.uri}/file/chan.tmp", headers={"Range": "bytes=10-19"})
        with urllib.request.urlopen(req) as r:
            assert r.status == 206 and r.read() == bytes(range(10, 20))
    finally:
        svc.close()


def test_local_job_submission_runs_spmd_workers(tmp_path):
    job = LocalJobSubmission(os.path.join(HERE, "mgpu", "spmd_worker.py"), ["--mode", "cpu"], num_processes=2,
                             dryad_home=str(tmp_path))
    job.add_job_option("friendlyname", "spmd smoke")
    assert job.get_status() == JobStatus.NotSubmitted
    job.submit_job()
    assert job.wait_for_completion(timeout=300) == JobStatus.Success, job.error_text()
    assert os.path.isdir(os.path.join(job.job_dir, "log")) and "checks passed" in job.error_text()
    bad = LocalJobSubmission(os.path.join(HERE, "mgpu", "does_not_exist.py"), num_processes=1, dryad_home=str(tmp_path))
    bad.submit_job()
    assert bad.wait_for_completion(timeout=120) == JobStatus.Failure


def test_vertices_are_placed_by_the_delay_scheduler_and_records_keep_locality(tmp_path):
    """ClusterInterface + LocalScheduler on the path of a job: every vertex is matched with a computer
    through affinity votes of its inputs; the vertex record keeps where it ran and how many records
    were local (DrVertexRecord per-channel statistics)."""
    import os
    import dryad_b200 as D
    ctx = D.DryadLinqContext(3, work_dir=str(tmp_path))
    got = ctx.from_enumerable(list(range(90)), 3).select(lambda x: x + 1).where(lambda x: x % 2 == 0).to_list()
    assert sorted(got) == [x + 1 for x in range(90) if (x + 1) % 2 == 0]
    log = open(os.path.join(str(tmp_path), "LocalJobs", "1", "calypso.log")).read()
    placed = [l for l in log.splitlines() if ",vertex," in l and "state=completed" in l]
    assert placed and all("computer=" in l and "locality=" in l for l in placed)
    # pointwise stages follow their input: everything they read is local to the computer they ran on
    follow = [l for l in placed if "locality=computer" in l]
    assert follow and all("remote=0" in l for l in follow)


def test_multi_port_vertex_reader_and_writer(tmp_path):
    """WriteItem(item, port) / per-port record writers / permuted multi-port reads over file and FIFO channels."""
    from dryad_b200.channel.factory import VertexReader, VertexWriter, port_order
    uris = [f"file://{tmp_path}/p{i}.bin" for i in range(3)] + ["fifo://32/mp_test_3"]
    w = VertexWriter(uris)
    for x in range(40):
        w.write_item((x, f"v{x}"), x % 4)
    w.get_record_writer(0).write((999, "extra"))
    assert w.records == [11, 10, 10, 10]
    w.close()
    ordered = list(VertexReader(uris, keep_port_order=True))
    assert [r[0] for r in ordered[:11]] == [0, 4, 8, 12, 16, 20, 24, 28, 32, 36, 999]
    assert sorted(r[0] for r in ordered) == sorted(list(range(40)) + [999])
    w2 = VertexWriter(uris[:3])
    w2.write_item_sequence([1, 2, 3], port=2)
    w2.close()
    assert port_order(8, False, 5) == port_order(8, False, 5) != list(range(8))     # seeded: retries read in the same order
    assert port_order(8, True, 5) == list(range(8))
    shuffled = VertexReader(uris[:3], keep_port_order=False, seed=5)
    assert sorted(shuffled.order) == [0, 1, 2] and list(shuffled) == [1, 2, 3]


def test_commands_and_mailbox_values_have_no_size_limit():
    """A start command with thousands of input channels (or a long substitution map) travels whole."""
    from dryad_b200.runtime.command import CMD_START, VertexCommand
    from dryad_b200.runtime.services import Mailbox
    m = Mailbox()
    big = VertexCommand(CMD_START, 7, 3, args=["name", "0", "5", "subst=" + "x" * 100000],
                        inputs=[(f"chan://{i}/{i}/0", i) for i in range(5000)])
    data = big.serialize()
    assert len(data) > (1 << 16)
    m.set("cmd", "start", data)
    v, st, got = m.block_on_status("cmd", 0, 1.0)
    assert got == data and st == "start"
    back = VertexCommand.parse(got)
    assert back.args[3] == big.args[3] and len(back.inputs) == 5000 and back.inputs[4999] == ("chan://4999/4999/0", 4999)
    m.close()


def test_process_service_daemon_main(tmp_path):
    """The node daemon as its own process: a client sets and long-polls mailbox values and fetches a file
    range over HTTP (ProcessService/Program.cs, HttpServer.cs)."""
    import signal
    import subprocess
    import sys
    from dryad_b200.runtime.process_service import MailboxClient
    (tmp_path / "data.bin").write_bytes(bytes(range(200)))
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.Popen([sys.executable, "-m", "dryad_b200.runtime.process_service", "--root", str(tmp_path)],
                         stdout=subprocess.PIPE, text=True, cwd=root)
    try:
        line = p.stdout.readline().strip()
        assert line.startswith("DRYAD_PROCESS_SERVER_URI=http://127.0.0.1:")
        c = MailboxClient(line.split("=", 1)[1], "test")
        assert c.get("proc1", "k", 0, 0.1) is None
        c.set("proc1", "k", "running", b"payload")
        v, st, data = c.get("proc1", "k", 0, 1.0)
        assert (st, data) == ("running", b"payload") and v >= 1
        assert c.fetch_file("data.bin", 10, 19) == bytes(range(10, 20))
        c.close()
    finally:
        p.send_signal(signal.SIGTERM)
        assert p.wait(10) == 0
