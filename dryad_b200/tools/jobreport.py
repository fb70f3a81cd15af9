"""Headless job report: plan, per-stage statistics, schedule and automated failure diagnosis from a
job directory (calypso.log + plan files).

Reference: JobBrowser/JOM (job object model parsing `calypso.log`, JOM/jobinfo.cs, JOM/dryadlog.cs:32-345),
the schedule (Gantt) view (JobBrowser/JobBrowser/jobschedule.cs) and the diagnosis engine
(JobBrowser/JobBrowser/Diagnosis.cs: VertexFailureDiagnosis :311, JobFailureDiagnosis :594 with
rules for serialization errors :545, failed-when-reading :573, plan parse errors :673, correlated
machine failures :749, deterministic input failure :807).  The WinForms GUI itself is out of scope
(no windowing here); this produces the same information as text / JSON.

    python -m dryad_b200.tools.jobreport <job_dir> [--json | --html [out.html]]
"""
from __future__ import annotations

import json
import os
import sys
from collections import defaultdict
from typing import Any, Dict, List


def parse_calypso(path: str) -> List[Dict[str, Any]]:
    events = []
    with open(path) as f:
        for line in f:
            parts = line.rstrip("\n").split(",")
            if len(parts) < 2:
                continue
            ev = {"ts": parts[0], "event": parts[1]}
            for kv in parts[2:]:
                if "=" in kv:
                    k, v = kv.split("=", 1)
                    ev[k] = v
            events.append(ev)
    return events


def build_job(job_dir: str) -> Dict[str, Any]:
    events = parse_calypso(os.path.join(job_dir, "calypso.log"))
    stages: Dict[str, Dict[str, Any]] = {}
    attempts: Dict[tuple, Dict[str, Any]] = {}
    job: Dict[str, Any] = {"dir": job_dir, "success": None, "error": None, "seconds": None}
    processes: Dict[str, Dict[str, Any]] = {}       # cluster mode: vertex-host processes
    rewrites: List[Dict[str, Any]] = []             # run-time graph rewrites (dynamic managers)
    for e in events:
        if e["event"] == "process":
            p = processes.setdefault(e.get("host", "?"), {"host": e.get("host", "?"), "pid": None, "created": None, "started": None,
                                                       "exited": None, "exitcode": None, "vertices": 0})
            if e.get("pid"):
                p["pid"] = int(e["pid"])
            if e.get("state") in ("created", "started", "exited"):
                p[e["state"]] = e["ts"]
            if e.get("exitcode") not in (None, "None"):
                p["exitcode"] = int(e["exitcode"])
            continue
        if e["event"] == "dynamic":
            rewrites.append({k: v for k, v in e.items() if k != "event"})
            continue
        if e["event"] == "topology":
            stages[e["stage"]] = {"id": int(e["stage"]), "name": e.get("stagename", ""), "type": e.get("type", ""),
                                  "partitions": int(e.get("partitions", 0)), "children": e.get("children", "").split(),
                                  "completed": 0, "failed": 0, "seconds": [], "records_in": 0, "records_out": 0}
        elif e["event"] == "vertex":
            key = (e["vertex"], e["version"])
            a = attempts.setdefault(key, {"vertex": e["vertex"], "version": int(e["version"]), "stage": e.get("stage"),
                                          "part": int(e.get("partInStage", 0)), "state": None, "start": None,
                                          "end": None, "error": None})
            a["state"] = e["state"]
            if "computer" in e:
                a["computer"] = e["computer"]
            if e["state"] == "created":
                continue
            if e["state"] == "starting":
                a["start"] = e["ts"]
            else:
                a["end"] = e["ts"]
            st = stages.get(e.get("stage"))
            if e["state"] == "completed" and st is not None:
                st["completed"] += 1
                st["seconds"].append(float(e.get("seconds", 0)))
                st["records_in"] += int(e.get("records_in", 0))
                st["records_out"] += int(e.get("records_out", 0))
            elif e["state"] == "failed":
                a["error"] = e.get("error")
                if st is not None:
                    st["failed"] += 1
        elif e["event"] == "stop":
            job["success"] = e.get("success") == "True"
            job["error"] = e.get("error")
            job["seconds"] = float(e["seconds"]) if "seconds" in e else None
            for k in ("completed", "failures", "duplicates", "invalidations"):
                if k in e:
                    job[k] = int(e[k])
    job["processes"] = sorted(processes.values(), key=lambda p: p["host"])
    job["rewrites"] = rewrites
    job["stages"] = sorted(stages.values(), key=lambda s: s["id"])
    job["attempts"] = sorted(attempts.values(), key=lambda a: (a["start"] or "", a["vertex"]))
    # static plan as written by the client: a job whose plan file cannot be parsed never had a graph
    plan_file = os.path.join(job_dir, "DryadLinqProgram__0.xml")
    job["plan_error"] = None
    if os.path.exists(plan_file):
        try:
            from ..plan import xmlplan
            xmlplan.parse_query_xml(open(plan_file).read())
        except Exception as ex:  # noqa: BLE001
            job["plan_error"] = f"{type(ex).__name__}: {ex}"
    job["diagnosis"] = diagnose(job)
    return job


def diagnose(job: Dict[str, Any]) -> List[str]:
    """Post-mortem rules, one per failure family of the reference's diagnosis engine
    (JobBrowser/JobBrowser/Diagnosis.cs: stack overflow :441, serialization error :545, failed when
    reading :573, XML plan parse error :673, correlated machine failures :749, deterministic input
    failure :807, correlated read failures :858), plus the families this engine adds (checked
    arithmetic, memory, lost rank, stage-wide operator failure)."""
    out: List[str] = []
    if job.get("plan_error"):
        out.append(f"the XML query plan cannot be parsed ({job['plan_error']}): the graph manager never built a graph — "
                   "client and graph manager disagree on the plan format")
    failed = [a for a in job["attempts"] if a["state"] == "failed"]
    if not failed and job.get("success") and not any(p.get("exitcode") not in (None, 0) for p in job.get("processes", [])):
        return out + ["job succeeded; no failed vertex attempts"]
    by_vertex = defaultdict(list)
    for a in failed:
        by_vertex[a["vertex"]].append(a)
    read_fail = lambda m: "cannot read" in m.lower() or "ReadFileError" in m or "StreamDoesNotExist" in m   # noqa: E731
    readers = []
    for v, atts in by_vertex.items():
        errs = {(a["error"] or "").split(":")[0] for a in atts}
        msg = (atts[-1]["error"] or "")
        same = len(atts) >= 2 and len(errs) == 1
        if same and not read_fail(msg):
            out.append(f"vertex {v} failed {len(atts)} times with the same error ({msg[:120]}): deterministic failure — "
                       "most likely a bug in the user function or bad input records, not the cluster")
        if "RecursionError" in msg or "stack overflow" in msg.lower() or "maximum recursion depth" in msg:
            out.append(f"vertex {v}: stack overflow — unbounded recursion in the user function (or a very deep expression)")
        if "TypeNotSerializable" in msg or "FailedToDeserialize" in msg or "pickle" in msg.lower():
            out.append(f"vertex {v}: serialization error — a record type cannot be (de)serialised")
        if read_fail(msg):
            readers.append(v)
            if same:
                out.append(f"vertex {v} failed {len(atts)} times READING the same input: deterministic input failure — the "
                           "stored data itself is bad (corrupt partition file or a record the reader cannot decode); "
                           "re-running the producer will not help")
            else:
                out.append(f"vertex {v}: failed while READING its input — the producer's output was lost or is corrupt; "
                           "the graph manager re-executes the producer")
        if "overflow" in msg.lower() and "stack" not in msg.lower():
            out.append(f"vertex {v}: checked arithmetic overflow in an aggregate")
        if "MemoryError" in msg or "out of memory" in msg.lower():
            out.append(f"vertex {v}: ran out of memory — increase the partition count")
        if "rank lost" in msg or "did not report" in msg:
            out.append(f"stage {v.split('.')[0]}: a rank stopped answering the graph manager (process died or hung); the job was aborted on every rank")
    if len(readers) >= 2:
        out.append(f"correlated read failures: {len(readers)} vertices ({', '.join(sorted(readers)[:6])}) failed reading their inputs — "
                   "look at the storage / the producing stage they share, not at the readers")
    machines = defaultdict(set)
    for a in failed:
        if a.get("computer") is not None:
            machines[a["computer"]].add(a["vertex"])
    all_machines = {a.get("computer") for a in job["attempts"] if a.get("computer") is not None}
    for m, vs in machines.items():
        if len(vs) >= 2 and len(all_machines) > 1 and sum(len(x) for x in machines.values()) == len(vs):
            out.append(f"correlated machine failures: all {len(vs)} failed vertices ran on computer {m} while the other "
                       f"{len(all_machines) - 1} computers had none — suspect that GPU / worker, not the program")
    died = [p for p in job.get("processes", []) if p.get("exitcode") not in (None, 0)]
    if died:
        slots = defaultdict(int)
        for p in died:
            slots[p["host"].split(".")[0]] += 1
        worst = max(slots.items(), key=lambda kv: kv[1])
        out.append(f"{len(died)} vertex-host process(es) died (exit codes {sorted({p['exitcode'] for p in died})}); the graph manager "
                   f"restarted them and re-ran their vertices" + (f" — slot {worst[0]} died {worst[1]} times: suspect that worker" if worst[1] >= 2 else ""))
    stages_failed = {a["stage"] for a in failed}
    if len(by_vertex) >= 3 and len(stages_failed) == 1:
        out.append(f"{len(by_vertex)} different vertices of stage {next(iter(stages_failed))} failed: the stage's "
                   "operator fails independently of the partition")
    if job.get("success") is False and job.get("error"):
        out.append("job error: " + job["error"][:300])
    if not out:
        out.append("failed attempts were transient: every failed vertex succeeded on a later version")
    return out


def render(job: Dict[str, Any]) -> str:
    lines = [f"Job {job['dir']}", f"  result: {'SUCCESS' if job['success'] else 'FAILURE' if job['success'] is False else 'unknown'}"
             + (f" in {job['seconds']:.3f}s" if job.get("seconds") is not None else "")]
    for k in ("completed", "failures", "duplicates", "invalidations"):
        if k in job:
            lines.append(f"  {k}: {job[k]}")
    lines.append("Stages:")
    lines.append(f"  {'id':>3} {'name':<38} {'parts':>5} {'done':>5} {'fail':>4} {'rec in':>10} {'rec out':>10} {'max s':>8} {'sum s':>8}")
    for s in job["stages"]:
        secs = s["seconds"] or [0.0]
        lines.append(f"  {s['id']:>3} {s['name'][:38]:<38} {s['partitions']:>5} {s['completed']:>5} {s['failed']:>4} "
                     f"{s['records_in']:>10} {s['records_out']:>10} {max(secs):>8.4f} {sum(secs):>8.4f}")
    lines.append("Schedule (attempt order):")
    for a in job["attempts"][:200]:
        lines.append(f"  {a['start'] or '-':<28} v{a['vertex']:<8} ver {a['version']} -> {a['state']}")
    if job.get("processes"):
        lines.append("Vertex hosts:")
        for p in job["processes"]:
            lines.append(f"  {p['host']:<8} pid {p['pid'] or '-':<8} created {p['created'] or '-':<28} exit {p['exitcode'] if p['exitcode'] is not None else '-'}")
    if job.get("rewrites"):
        lines.append("Run-time graph rewrites:")
        for r in job["rewrites"]:
            lines.append("  " + " ".join(f"{k}={v}" for k, v in r.items() if k != "ts"))
    lines.append("Diagnosis:")
    lines += ["  - " + d for d in job["diagnosis"]]
    plan = os.path.join(job["dir"], "QueryGraph__0.txt")
    if os.path.exists(plan):
        lines.append("Plan:")
        lines += ["  " + l for l in open(plan).read().splitlines()]
    return "\n".join(lines)


def _ts(text: str) -> float:
    import datetime
    return datetime.datetime.strptime(text, "%Y-%m-%dT%H:%M:%S.%fZ").replace(tzinfo=datetime.timezone.utc).timestamp()


def render_html(job: Dict[str, Any]) -> str:
    """Self-contained HTML page with two inline SVG drawings — the stage graph with per-stage progress and
    the schedule (one lane per computer, one bar per vertex attempt) — followed by the text report.
    The information content of the reference's job browser views (JobBrowser/JobBrowser: static / dynamic
    plan drawing, `jobschedule.cs` Gantt view, diagnosis pane), without a window system."""
    import html
    stages = job["stages"]
    by_id = {str(s["id"]): s for s in stages}
    level: Dict[str, int] = {}

    def depth(sid: str) -> int:
        if sid not in level:
            level[sid] = 0
            kids = [c for c in by_id[sid]["children"] if c in by_id]
            level[sid] = 1 + max((depth(c) for c in kids), default=-1)
        return level[sid]

    for sid in by_id:
        depth(sid)
    rows: Dict[int, List[str]] = defaultdict(list)
    for sid, lv in level.items():
        rows[lv].append(sid)
    bw, bh, gx, gy = 210, 44, 30, 36
    pos = {}
    for lv, ids in rows.items():
        for k, sid in enumerate(sorted(ids, key=int)):
            pos[sid] = (20 + k * (bw + gx), 20 + lv * (bh + gy))
    gw = max((x for x, _ in pos.values()), default=0) + bw + 20
    gh = max((y for _, y in pos.values()), default=0) + bh + 20
    g = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{gw}" height="{gh}" font-family="monospace" font-size="11">']
    for sid, s in by_id.items():
        x, y = pos[sid]
        for c in s["children"]:
            if c in pos:
                cx, cy = pos[c]
                g.append(f'<line x1="{cx + bw / 2}" y1="{cy + bh}" x2="{x + bw / 2}" y2="{y}" stroke="#888"/>')
    for sid, s in by_id.items():
        x, y = pos[sid]
        done = s["completed"] >= s["partitions"] > 0
        fill = "#f6c7c7" if s["failed"] and not done else ("#cfe8cf" if done else "#eeeeee")
        g.append(f'<rect x="{x}" y="{y}" width="{bw}" height="{bh}" rx="5" fill="{fill}" stroke="#444"/>')
        g.append(f'<text x="{x + 6}" y="{y + 16}">#{sid} {html.escape(s["name"][:28])}</text>')
        g.append(f'<text x="{x + 6}" y="{y + 32}">x{s["partitions"]}  done {s["completed"]}  failed {s["failed"]}</text>')
    g.append("</svg>")
    # schedule
    atts = [a for a in job["attempts"] if a.get("start") and a.get("end")]
    sched = ["<p>(no timed attempts)</p>"]
    if atts:
        t0 = min(_ts(a["start"]) for a in atts)
        t1 = max(_ts(a["end"]) for a in atts)
        span = max(t1 - t0, 1e-6)
        lanes = sorted({str(a.get("computer", "-")) for a in atts}, key=lambda v: (len(v), v))
        lw, lh, left = 900, 22, 90
        sched = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{lw + left + 20}" height="{len(lanes) * (lh + 6) + 40}" '
                 f'font-family="monospace" font-size="10">']
        for k, lane in enumerate(lanes):
            y = 10 + k * (lh + 6)
            sched.append(f'<text x="4" y="{y + 15}">computer {html.escape(lane)}</text>')
            sched.append(f'<rect x="{left}" y="{y}" width="{lw}" height="{lh}" fill="#fafafa" stroke="#ddd"/>')
        for a in atts:
            k = lanes.index(str(a.get("computer", "-")))
            x = left + (_ts(a["start"]) - t0) / span * lw
            w = max(1.5, (_ts(a["end"]) - _ts(a["start"])) / span * lw)
            color = {"completed": "#4c9f4c", "failed": "#c0392b", "completed_duplicate": "#d4a017"}.get(a["state"], "#999")
            tip = f'v{a["vertex"]} ver {a["version"]} {a["state"]}' + (f' — {a["error"][:80]}' if a.get("error") else "")
            sched.append(f'<rect x="{x:.1f}" y="{10 + k * (lh + 6) + 2}" width="{w:.1f}" height="{lh - 4}" fill="{color}">'
                         f'<title>{html.escape(tip)}</title></rect>')
        sched.append(f'<text x="{left}" y="{len(lanes) * (lh + 6) + 26}">0 s</text>')
        sched.append(f'<text x="{left + lw - 60}" y="{len(lanes) * (lh + 6) + 26}">{span:.3f} s</text>')
        sched.append("</svg>")
    return ("<!doctype html><html><head><meta charset='utf-8'><title>" + html.escape(job["dir"]) + "</title></head><body>"
            "<h2>Stage graph</h2>" + "".join(g) + "<h2>Schedule</h2>" + "".join(sched) +
            "<h2>Report</h2><pre>" + html.escape(render(job)) + "</pre></body></html>")


def main(argv=None):
    argv = argv or sys.argv[1:]
    if not argv:
        print(__doc__)
        return 2
    job = build_job(argv[0])
    if "--html" in argv:
        i = argv.index("--html")
        out = argv[i + 1] if i + 1 < len(argv) and not argv[i + 1].startswith("--") else os.path.join(argv[0], "report.html")
        with open(out, "w") as f:
            f.write(render_html(job))
        print(out)
        return 0
    print(json.dumps(job, indent=1) if "--json" in argv else render(job))
    return 0


if __name__ == "__main__":
    sys.exit(main())
