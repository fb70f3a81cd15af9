"""Re-run ONE failed vertex of a finished job, alone, for debugging or profiling.

Reference: JobBrowser/JobBrowser/LocalDebuggingAndProfiling.cs (re-create the failed vertex's
working directory on the developer's machine and start the vertex host on it, optionally under a
profiler) and DumpRestartCommand (DryadVertex/VertexHost/system/dprocess/src/dvertexpncontrol.h:72).

A failed vertex leaves `<job dir>/restart/vertex-<stage>-<partition>.cmd` (its start command on the
TLV wire format) and its inputs as partition files (runtime/executor.py `_dump_restart`).  The
vertex CODE is the user's query, so a replay needs the query again:

  in a program:   rerun.replay(query, job_dir, "5.1")           # returns the output ports, or raises
                  rerun.replay(query, job_dir, "5.1", profile=True)   # ... under cProfile, prints the top
  from a shell:   python -m dryad_b200.tools.rerun <job_dir> [<stage>.<partition>]
                  lists the restart descriptors of the job (stage, partition, version, inputs, error)

A CLUSTER-MODE job directory holds the job package (the vertex code) and every channel as a file, so its
failed vertices can be replayed without the client program:

  from a shell:   python -m dryad_b200.tools.rerun <job_dir> <stage>.<partition> --run [--profile] [--pdb]
                  runs the vertex alone in this process with the inputs it had (exit code 0 = it completed now)
  in a program:   rerun.replay_cluster(job_dir, "4.1") -> result record of the attempt
"""
from __future__ import annotations

import glob
import os
import sys
from typing import List, Optional


def list_restarts(job_dir: str) -> List[dict]:
    from ..runtime.command import VertexCommand
    out = []
    for path in sorted(glob.glob(os.path.join(job_dir, "restart", "vertex-*.cmd"))):
        with open(path, "rb") as f:
            cmd = VertexCommand.parse(f.read())
        if cmd is None:
            continue
        out.append({"file": path, "stage": cmd.vertex_id, "stage_name": cmd.args[0] if cmd.args else "",
                    "partition": int(cmd.args[1]) if len(cmd.args) > 1 else 0, "version": cmd.version,
                    "inputs": cmd.inputs, "error": cmd.message})
    return out


def replay(query, job_dir: str, vertex: str, profile: bool = False):
    """Plan `query` again (stage ids are deterministic), pick the stage of `vertex` = "<stage>.<partition>",
    feed it the inputs saved by the failed job and run it alone."""
    from ..plan.querygen import QueryGen
    from ..runtime.executor import JobExecutor
    uid, part = (int(x) for x in vertex.split("."))
    restart = os.path.join(job_dir, "restart", f"vertex-{uid}-{part}.cmd")
    if not os.path.exists(restart):
        raise FileNotFoundError(f"no restart descriptor for vertex {vertex} in {job_dir}")
    ctx = query.context
    ex = JobExecutor(ctx, QueryGen(ctx, [query.materialize_lazy()]))
    if not profile:
        return ex.run_single_vertex(uid, part, restart_file=restart)
    import cProfile
    import pstats
    pr = cProfile.Profile()
    pr.enable()
    try:
        return ex.run_single_vertex(uid, part, restart_file=restart)
    finally:
        pr.disable()
        pstats.Stats(pr).sort_stats("cumulative").print_stats(15)


def replay_cluster(job_dir: str, vertex: str, profile: bool = False, post_mortem: bool = False) -> dict:
    """Run one vertex of a cluster-mode job alone, in this process, from its saved start command: same job
    package, same input channel versions.  Outputs are written as a version far above the job's own, so
    nothing the job produced is overwritten."""
    from ..runtime.command import VertexCommand
    from ..runtime.vertex_host import VertexHost, _Run
    uid, part = (int(x) for x in vertex.split("."))
    path = os.path.join(job_dir, "restart", f"vertex-{uid}-{part}.cmd")
    if not os.path.exists(path):
        raise FileNotFoundError(f"no restart descriptor for vertex {vertex} in {job_dir}")
    with open(path, "rb") as f:
        cmd = VertexCommand.parse(f.read())
    host = VertexHost("http://127.0.0.1:9", "rerun", job_dir)          # no manager: the mailbox is never contacted
    host.remote_reads = False
    try:
        extra = cmd.args[3:]
        subst = {}
        for a in extra:
            if a.startswith("subst="):
                import json
                subst = json.loads(a[6:])
            elif a.startswith("widths="):
                import json
                host._apply_widths(json.loads(a[7:]))
        internal = extra[1:4] if extra and extra[0] == "internal" else None
        member = {"uid": cmd.vertex_id, "part": int(cmd.args[1]), "gmv": int(cmd.args[2]), "version": cmd.version + 1000,
                  "inputs": [u.split("://", 1)[1].split("/") for u, _n in cmd.inputs]}
        run = _Run(host, member, subst, internal)
        if post_mortem:
            host.raise_errors = True                   # the vertex's exception reaches the caller (and its debugger)
        if profile:
            import cProfile
            import pstats
            pr = cProfile.Profile()
            pr.enable()
            try:
                return host.run_member(run)
            finally:
                pr.disable()
                pstats.Stats(pr).sort_stats("cumulative").print_stats(15)
        return host.run_member(run)
    finally:
        host.close()


def main(argv: Optional[List[str]] = None) -> int:
    argv = sys.argv[1:] if argv is None else argv
    if not argv:
        print(__doc__)
        return 2
    if "--run" in argv:
        flags = [a for a in argv if a.startswith("--")]
        pos = [a for a in argv if not a.startswith("--")]
        if len(pos) < 2:
            print("usage: rerun <job_dir> <stage>.<partition> --run [--profile] [--pdb]")
            return 2
        try:
            res = replay_cluster(pos[0], pos[1], profile="--profile" in flags, post_mortem="--pdb" in flags)
        except Exception:  # noqa: BLE001  (--pdb: stop where the vertex code raised)
            import pdb
            import traceback
            traceback.print_exc()
            pdb.post_mortem()
            return 1
        print(f"vertex {pos[1]}: {'completed' if res['ok'] else 'FAILED'} in {res['seconds']:.3f}s, "
              f"{res['n_in']} records in, {res['n_out']} out")
        if not res["ok"]:
            print("  " + res["message"][:2000])
        return 0 if res["ok"] else 1
    want = argv[1] if len(argv) > 1 else None
    found = list_restarts(argv[0])
    for r in found:
        tag = f"{r['stage']}.{r['partition']}"
        if want and tag != want:
            continue
        print(f"vertex {tag}  {r['stage_name']}  version {r['version']}")
        for u, n in r["inputs"]:
            print(f"  input  {u}  ({n} records)")
        print(f"  error  {r['error'][:300]}")
    if not found:
        print("no restart descriptors (the job had no finally-failed vertex)")
    return 0


if __name__ == "__main__":
    sys.exit(main())
