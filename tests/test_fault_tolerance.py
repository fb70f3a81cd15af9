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


def test_failed_vertex_leaves_a_restart_descriptor_and_can_be_rerun_alone(tmp_path):
    """DumpRestartCommand + local re-run of one failed vertex (LocalDebuggingAndProfiling role)."""
    from dryad_b200.tools import rerun
    ctx = _ctx(tmp_path, parts=2)
    state = {"broken": True}

    def f(x):
        if state["broken"] and x == 37:
            raise ValueError("bad record 37")
        return x * 10
    q = ctx.from_enumerable(DATA, 2).select(lambda x: x + 0).where(lambda x: x >= 0).select(f)
    with pytest.raises(DryadLinqException):
        q.to_list()
    job_dir = os.path.join(str(tmp_path), "LocalJobs", "1")
    found = rerun.list_restarts(job_dir)
    assert len(found) == 1 and "bad record 37" in found[0]["error"] and found[0]["inputs"]
    tag = f"{found[0]['stage']}.{found[0]['partition']}"
    assert rerun.main([job_dir]) == 0
    with pytest.raises(ValueError):                      # same inputs, same code: same failure, now under a debugger
        rerun.replay(q, job_dir, tag)
    state["broken"] = False                              # "fix the bug", replay only that vertex
    ports = rerun.replay(q, job_dir, tag, profile=True)
    half = DATA[:50] if found[0]["partition"] == 0 else DATA[50:]
    assert list(ports[0]) == [x * 10 for x in half]
    # the diagnosis names the failure family
    job = jobreport.build_job(job_dir)
    assert any("deterministic failure" in d for d in job["diagnosis"])


def test_diagnosis_rule_families(tmp_path):
    """One synthetic event log per failure family of the diagnosis engine (Diagnosis.cs rule set)."""
    def job(events, plan="<Query><DryadLinqVersion>x</DryadLinqVersion><QueryPlan></QueryPlan></Query>"):
        d = tmp_path / f"job{len(list(tmp_path.iterdir()))}"
        (d / "log").mkdir(parents=True)
        lines = ["t0,start,job=j,stages=2", "t0,topology,stage=1,stagename=Select,type=Select,partitions=4,children="]
        for i, (vertex, version, state, kv) in enumerate(events):
            body = ",".join(f"{k}={v}" for k, v in kv.items())
            lines.append(f"t{i + 1},vertex,vertex={vertex},version={version},stagename=Select,stage=1,partInStage={vertex.split('.')[1]},state={state}" + ("," + body if body else ""))
        lines.append("tz,stop,success=False,error=boom")
        (d / "calypso.log").write_text("\n".join(lines) + "\n")
        if plan is not None:
            (d / "DryadLinqProgram__0.xml").write_text(plan)
        return jobreport.build_job(str(d))["diagnosis"]
    so = job([("1.0", 0, "failed", {"error": "RecursionError: maximum recursion depth exceeded"})])
    assert any("stack overflow" in x for x in so)
    ser = job([("1.0", 0, "failed", {"error": "PicklingError: cannot pickle lambda"})])
    assert any("serialization error" in x for x in ser)
    rd = job([("1.0", 0, "failed", {"error": "InputReadError: cannot read output of stage 0[0]"}), ("1.0", 1, "completed", {})])
    assert any("failed while READING" in x for x in rd)
    det_in = job([("1.0", v, "failed", {"error": "InputReadError: cannot read output of stage 0[0]"}) for v in range(3)])
    assert any("deterministic input failure" in x for x in det_in)
    corr = job([("1.0", 0, "failed", {"error": "ReadFileError: x"}), ("1.1", 0, "failed", {"error": "ReadFileError: y"})])
    assert any("correlated read failures" in x for x in corr)
    mach = job([("1.0", 0, "failed", {"error": "RuntimeError: a", "computer": 2}), ("1.1", 0, "failed", {"error": "RuntimeError: b", "computer": 2}),
                ("1.2", 0, "completed", {"computer": 0}), ("1.3", 0, "completed", {"computer": 1})])
    assert any("correlated machine failures" in x and "computer 2" in x for x in mach)
    bad_plan = job([("1.0", 0, "failed", {"error": "RuntimeError: a"})], plan="<Query><Broken")
    assert any("XML query plan cannot be parsed" in x for x in bad_plan)
    lost = job([("1.0", 0, "failed", {"error": "rank 1 did not report stage 1 v0 within 60 s (rank lost)"})])
    assert any("stopped answering" in x for x in lost)
