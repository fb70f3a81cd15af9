"""Failure, re-execution and observability (the reference has no automated tests for
these paths; fault injection is added here)."""
import os

import pytest

import dryad_b200 as D
from dryad_b200 import DryadLinqException, ErrorCode
from dryad_b200.plan.qnodes import NodeType
from dryad_b200.runtime.executor import InputReadError
from dryad_b200.tools import jobreport

DATA = list(range(100))


def _ctx(tmp_path, parts=3):
    return D.DryadLinqContext(parts, work_dir=str(tmp_path))


def test_transient_vertex_failures_are_retried_with_new_versions(tmp_path):
    ctx = _ctx(tmp_path)
    seen = []

    def inject(node, part, version):
        seen.append((node.uid, part, version))
        if node.type != NodeType.InputTable and part == 1 and version < 2:
            raise RuntimeError("injected transient fault")
    ctx.fault_injector = inject
    got = ctx.from_enumerable(DATA, 3).select(lambda x: x * 2).to_list()
    assert sorted(got) == [x * 2 for x in DATA]
    versions = [v for (_, p, v) in seen if p == 1]
    assert max(versions) == 2                      # two failed attempts, third version succeeded
    job = jobreport.build_job(os.path.join(str(tmp_path), "LocalJobs", "1"))
    assert job["success"] and job["failures"] == 2
    assert any("transient" in d for d in job["diagnosis"])


def test_job_aborts_after_max_failures_and_reports_vertex_error(tmp_path):
    ctx = _ctx(tmp_path)

    def bad(x):
        raise ValueError("user code is broken")
    with pytest.raises(DryadLinqException) as ei:
        ctx.from_enumerable(DATA, 3).select(bad).to_list()
    assert ei.value.error_code == ErrorCode.DidNotCompleteSuccessfully
    assert "failed 6 times" in str(ei.value) and "user code is broken" in str(ei.value)
    job = jobreport.build_job(os.path.join(str(tmp_path), "LocalJobs", "1"))
    assert job["success"] is False
    assert any("deterministic failure" in d for d in job["diagnosis"])
    text = jobreport.render(job)
    assert "FAILURE" in text and "Diagnosis:" in text
    assert os.path.exists(os.path.join(str(tmp_path), "LocalJobs", "1", "log", "error.txt"))


def test_read_error_reexecutes_the_producer(tmp_path):
    ctx = _ctx(tmp_path, parts=2)
    runs = {"producer": 0}
    state = {"raised": False}

    def produce(x):
        return x + 1

    def inject(node, part, version):
        members = node.members if node.type == NodeType.Super else [node]
        if any(m.type == NodeType.HashPartition for m in members) and part == 0:
            runs["producer"] += 1
        if node.type == NodeType.Super and members[0].type == NodeType.Merge and part == 0 and not state["raised"]:
            state["raised"] = True
            producer = node.children[0][0]
            raise InputReadError(producer.uid, 0, "cannot read channel 0: checksum mismatch")
    ctx.fault_injector = inject
    got = ctx.from_enumerable(DATA, 2).select(produce).hash_partition(lambda x: x, 2).to_list()
    assert sorted(got) == [x + 1 for x in DATA]
    assert runs["producer"] == 2                   # the revoked producer partition ran twice
    job = jobreport.build_job(os.path.join(str(tmp_path), "LocalJobs", "1"))
    assert job["invalidations"] == 1 and job["success"]
    assert any("READING its input" in d for d in job["diagnosis"])


def test_failed_job_leaves_no_partial_output(tmp_path):
    ctx = _ctx(tmp_path)
    uri = f"partfile://{tmp_path}/never.pt"

    def inject(node, part, version):
        if node.type != NodeType.InputTable and part == 2:
            raise RuntimeError("always fails")
    ctx.fault_injector = inject
    with pytest.raises(DryadLinqException):
        ctx.from_enumerable(DATA, 3).select(lambda x: x).to_store(uri).submit_and_wait()
    assert not any(f.startswith("never") for f in os.listdir(tmp_path))


def test_event_log_plan_files_and_report(tmp_path):
    ctx = _ctx(tmp_path)
    q = (ctx.from_enumerable(["a b", "b c", "c"], 2).select_many(lambda l: l.split(" "))
         .group_by(lambda w: w).select(lambda g: (g.key, g.count())))
    assert sorted(q.to_list()) == [("a", 1), ("b", 2), ("c", 2)]
    jd = os.path.join(str(tmp_path), "LocalJobs", "1")
    assert os.path.exists(os.path.join(jd, "calypso.log"))
    xml = open(os.path.join(jd, "DryadLinqProgram__0.xml")).read()
    assert "<QueryPlan>" in xml and "<ConnectionOperator>CROSSPRODUCT</ConnectionOperator>" in xml
    job = jobreport.build_job(jd)
    names = [s["name"] for s in job["stages"]]
    assert any("HashPartition" in n for n in names) and job["success"]
    assert jobreport.main([jd]) == 0


def test_memory_fifo_channels_form_start_cliques_and_results_are_unchanged(monkeypatch):
    """StaticConfig.UseMemoryFIFO streams the inputs of a Join: those edges are FIFO in the plan
    and the graph manager starts producer and consumer as one clique; the query result is the same."""
    import dryad_b200 as D
    from dryad_b200.plan import static_config
    left = [{"k": i % 50, "a": i} for i in range(400)]
    right = [{"k": i, "b": i * 3} for i in range(50)]

    def run(ctx):
        l = ctx.from_enumerable(left, 4).select(lambda r: {"k": r.k, "a": r.a + 1}).hash_partition(lambda r: r.k, 4)
        r = ctx.from_enumerable(right, 4).select(lambda r: {"k": r.k, "b": r.b}).hash_partition(lambda r: r.k, 4)
        return l.join(r, lambda x: x.k, lambda y: y.k, lambda x, y: (x.k, x.a + y.b))

    want = sorted(run(D.DryadLinqContext(1, local_debug=True)).to_list())
    monkeypatch.setattr(static_config, "USE_MEMORY_FIFO", True)
    ctx = D.DryadLinqContext(4)
    q = run(ctx)
    assert "MEMORYFIFO" in q.to_xml()
    assert sorted(q.to_list()) == want
    gm = ctx.last_executor.gm
    assert any(len(gm.clique(v)) > 1 for v in range(gm.live_vertices))
