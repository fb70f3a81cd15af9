"""The job package: what the client ships so that OTHER processes can plan and run the same job.

The reference compiles the query into a vertex DLL + `DryadLinqProgram__.xml` and stages both,
with the serialised client objects, as job resources (LinqToDryad/DryadLinqQueryGen.cs:316-399
GenerateDryadProgram, DryadLinqObjectStore.cs, LinqToDryad/PeloponneseJobSubmission.cs:36-179).  Here
the "vertex code" is the logical query itself: the root expressions with their lambdas (pickled by
value, so a lambda defined inside a function travels), the context properties that influence
planning and execution, and the client's module search path.  Every process that loads the package
runs the same deterministic planner and gets the same physical plan — stage ids, partition counts,
channel types — which is what lets the graph manager address vertices by (stage id, partition).

    <job dir>/job.pkl               this package
    <job dir>/DryadLinqProgram__.xml   the plan as XML (same artefact the in-process executor writes)
"""
from __future__ import annotations

import os
import pickle
import sys
from typing import Any, Dict

PACKAGE, PLAN_XML = "job.pkl", "DryadLinqProgram__.xml"

_CTX_FIELDS = ("num_partitions", "storage_scheme", "intermediate_data_compression_scheme",
               "output_data_compression_scheme", "compile_for_vertex_debugging", "job_friendly_name",
               "job_min_nodes", "job_max_nodes", "threads_per_worker", "enable_speculative_duplication",
               "select_order_preserving", "max_failures_per_vertex", "fail_fast_on_deterministic_errors",
               "retain_intermediate", "sample_rate", "group_capacity_hint", "broadcast_join_rows", "fault_injector",
               "job_runtime_limit", "dynamic_opt_level", "aggregation_tree_fan_in", "broadcast_tree_min_consumers",
               "dynamic_data_per_vertex", "duplicate_default_seconds", "duplicate_min_seconds", "use_memory_fifo", "remote_channel_reads", "vertex_timeout_seconds")


def write(job_dir: str, context, roots) -> str:
    """`roots`: Query objects rooted at to_store.  Returns the package path."""
    import cloudpickle
    os.makedirs(job_dir, exist_ok=True)
    settings: Dict[str, Any] = {k: getattr(context, k) for k in _CTX_FIELDS if hasattr(context, k)}
    settings["runtime_logging_level"] = context.runtime_logging_level.value
    settings["job_environment_variables"] = dict(context.job_environment_variables)
    pkg = {"version": context.version, "settings": settings, "roots": [q.expr for q in roots]}
    path = os.path.join(job_dir, PACKAGE)
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        # two pickles: the module search path first, because functions pickled by reference need the
        # client's modules importable BEFORE the second pickle is loaded
        pickle.dump({"sys_path": [p for p in sys.path if p], "cwd": os.getcwd()}, f)
        from ..api.context import Table
        Table.by_reference.on = True                  # stored tables travel as (URI, ranges), not as data
        try:
            cloudpickle.dump(pkg, f, protocol=pickle.HIGHEST_PROTOCOL)
        finally:
            Table.by_reference.on = False
    os.replace(tmp, path)
    return path


def load(job_dir: str):
    """-> (context, plan): a LOCAL context configured like the client's and the built physical plan."""
    with open(os.path.join(job_dir, PACKAGE), "rb") as f:
        env = pickle.load(f)
        for p in [env.get("cwd")] + list(env.get("sys_path", [])):
            if p and p not in sys.path:
                sys.path.append(p)
        pkg = pickle.load(f)
    from ..api.context import DryadLinqContext
    from ..api.query import Query
    from ..plan.querygen import QueryGen
    s = pkg["settings"]
    ctx = DryadLinqContext(s.get("num_partitions"), s.get("storage_scheme", "partfile"), work_dir=job_dir)
    for k, v in s.items():
        if k == "runtime_logging_level":
            from ..api.context import QueryLoggingLevel
            ctx.runtime_logging_level = QueryLoggingLevel(v)
        elif k == "job_environment_variables":
            for kk, vv in v.items():
                ctx.job_environment_variables[kk] = vv
                os.environ[kk] = str(vv)
        else:
            setattr(ctx, k, v)
    ctx.schedule_vertices = False          # placement is the graph-manager process's business
    top = 0
    for e in pkg["roots"]:
        for n in e.walk():
            top = max(top, n.id)
            if n.op == "source" and "table" in n.params:
                ctx._tables[n.id] = n.params["table"]
    # expressions the planner creates here (rewrites) must not reuse the ids of the client's expressions
    import itertools
    from ..plan import logical
    logical._ids = itertools.count(max(top, next(logical._ids)) + 1)
    plan = QueryGen(ctx, [Query(ctx, e) for e in pkg["roots"]]).build()
    return ctx, plan
