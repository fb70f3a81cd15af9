"""Job executor: runs a physical plan stage by stage, vertex by vertex.

Role of the GraphManager run loop + vertex hosts (reference call stack SURVEY §3.2/§3.3:
GraphManager/vertex/DrGraph.cpp:86-116, DrVertex.cpp:431-885, DrCohort.cpp:720-745): every
(stage, partition) is a vertex with VERSIONED attempts; a failed attempt is re-executed from its
retained inputs; the job aborts after `max_failures_per_vertex` failures of the same vertex
(GraphManager/graph/DrGraphParameters.cpp:52); outputs are committed only when the whole job
succeeded (DrGraph.cpp:204-254).  The state machine itself lives in the native scheduler
(`csrc/host/sched.cpp`, bound in `dryad_b200.sched`); this module wires data to it.
"""
from __future__ import annotations

import time
import traceback
from typing import Any, Callable, Dict, List, Optional, Tuple

from ..api.errors import DryadLinqException, ErrorCode
from ..plan.qnodes import ConnOp, DLinqQueryNode, NodeType
from ..plan.querygen import QueryGen
from . import vertex_ops
from .vertex_ops import VertexEnv


class VertexFailure(Exception):
    def __init__(self, node, part, version, cause):
        super().__init__(f"vertex {node.type.value}#{node.uid}[{part}] v{version} failed: {cause}")
        self.node, self.part, self.version, self.cause = node, part, version, cause


class InputReadError(Exception):
    """A vertex could not read the output of (producer stage uid, partition): the graph manager
    revokes that output and re-executes the producer (DrGraph::ReportFailure, DrGraph.cpp:392-415)."""

    def __init__(self, producer_uid: int, producer_part: int, message: str = ""):
        super().__init__(message or f"cannot read output of stage {producer_uid}[{producer_part}]")
        self.producer_uid, self.producer_part = producer_uid, producer_part


class JobExecutor:
    def __init__(self, context, plan: QueryGen, reporter=None):
        self.context = context
        self.plan = plan.build()
        self.reporter = reporter or context._make_reporter()
        self.out: Dict[int, List[Optional[List[List]]]] = {}     # node uid -> per partition -> ports
        self.stats: List[dict] = []
        self.backend = context._backend()
        from ..utils import trace
        self._placement: Dict[Tuple[int, int], int] = {}          # (stage uid, partition) -> computer its output lives on
        self.cluster = None
        self.timeline = trace.GpuTimeline() if (trace.enabled(context) and getattr(self.backend, "name", "") == "gpu") else None
        self.gpu_timeline: List[dict] = []

    # ---- input assembly (GraphBuilder.ConnectPointwise / ConnectCrossProduct) --------------
    def _inputs(self, node: DLinqQueryNode, i: int) -> List[List]:
        ins: List[List] = []
        t = node.type if node.type != NodeType.Super else node.members[0].type
        if t == NodeType.Zip:
            lp = [p[0] for p in self.out[node.children[0][0].uid]]
            rp = [p[0] for p in self.out[node.children[1][0].uid]]
            if all(hasattr(b, "cols") and not b.is_rows for b in lp + rp):
                # columnar re-alignment: slices of the right side's partitions, no records
                from .columnar import ColumnBatch
                start = sum(b.n for b in lp[:i])
                need = lp[i].n
                segs, pos = [], 0
                for b in rp:
                    lo, hi = max(start, pos), min(start + need, pos + b.n)
                    if lo < hi:
                        segs.append(b.slice(lo - pos, hi - pos))
                    pos += b.n
                right = ColumnBatch.concat(segs) if segs else rp[0].slice(0, 0)
                return [[lp[i]], [right]]
            tr = self.backend.to_records
            left = [[tr(p[0])] for p in self.out[node.children[0][0].uid]]
            right = [[tr(p[0])] for p in self.out[node.children[1][0].uid]]
            start = sum(len(left[s][0]) for s in range(i))
            need = len(left[i][0])
            flat: List = []
            pos = 0
            for s in range(node.children[1][0].partition_count):   # re-align right to left's boundaries
                seg = right[s][0]
                lo, hi = max(start, pos), min(start + need, pos + len(seg))
                if lo < hi:
                    flat.extend(seg[lo - pos:hi - pos])
                pos += len(seg)
            return [[left[i][0]], [flat]]
        for k, (child, op) in enumerate(node.children):
            co = self.out[child.uid]
            cpc = child.partition_count
            if t == NodeType.ForkBranch:
                idx = (node.members[0] if node.type == NodeType.Super else node).params["index"]
                ins.append([co[i][idx]])
            elif op == ConnOp.CrossProduct:
                ins.append([co[s][i] for s in range(cpc)])
            elif t == NodeType.Concat:
                ins.append([])      # filled below
            elif t == NodeType.Reverse and k == 0:
                ins.append([co[cpc - 1 - i][0]])
            elif cpc == node.partition_count:
                ins.append([co[i][0]])
            elif cpc == 1:
                ins.append([co[0][0]])                               # 1:N broadcast
            elif node.partition_count == 1:
                ins.append([co[s][0] for s in range(cpc)])           # N:1 gather, port order
            else:
                g = (node.members[0] if node.type == NodeType.Super else node).params.get("group")
                if g:
                    ins.append([co[s][0] for s in range(g * i, min(cpc, g * i + g))])
                else:
                    raise DryadLinqException(ErrorCode.Internal,
                                             f"cannot connect {cpc} -> {node.partition_count} partitions pointwise")
        if t == NodeType.Concat:
            acc = 0
            for child, _ in node.children:
                if i < acc + child.partition_count:
                    return [[self.out[child.uid][i - acc][0]]]
                acc += child.partition_count
        if t == NodeType.SlidingWindow:
            w = (node.members[0] if node.type == NodeType.Super else node).params["window"]
            cb = [p[0] for p in self.out[node.children[0][0].uid]]
            if all(hasattr(b, "cols") and not b.is_rows for b in cb):
                # columnar halo: the last w-1 rows that precede partition i, as slices
                from .columnar import ColumnBatch
                parts_h, have, s = [], 0, i - 1
                while s >= 0 and have < w - 1:
                    take = min(cb[s].n, (w - 1) - have)
                    if take:
                        parts_h.insert(0, cb[s].slice(cb[s].n - take, cb[s].n))
                        have += take
                    s -= 1
                ins.append([ColumnBatch.concat(parts_h) if parts_h else cb[i].slice(0, 0)])
                return ins
            co = [[self.backend.to_records(p[0])] for p in self.out[node.children[0][0].uid]]
            halo: List = []
            s = i - 1
            while s >= 0 and len(halo) < w - 1:
                seg = co[s][0]
                take = (w - 1) - len(halo)
                halo = list(seg[-take:]) + halo if take > 0 and seg else halo
                s -= 1
            ins.append([halo if i > 0 else []])
        return ins

    # ---- restart descriptors (DumpRestartCommand, dvertexpncontrol.h:72; LocalDebuggingAndProfiling.cs) ----
    def _dump_restart(self, node: DLinqQueryNode, part: int, version: int, ins, error: str) -> Optional[str]:
        """A failed vertex leaves everything needed to run it ALONE again: its start command on the
        native TLV wire format (stage, partition, version, input channel URIs) and its inputs as
        partition files, under <job dir>/restart/.  `tools/rerun.py` replays it."""
        import os
        from ..channel import store
        from .command import CMD_START, STATE_FAILED, VertexCommand
        job_dir = getattr(self.reporter, "job_dir", None)
        if not job_dir:
            return None
        d = os.path.join(job_dir, "restart")
        os.makedirs(d, exist_ok=True)
        stem = os.path.join(d, f"vertex-{node.uid}-{part}")
        uris = []
        try:
            for k, payloads in enumerate(ins):
                uri = f"partfile://{stem}-in{k}.pt"
                store.write_table(uri, [self.backend.to_records(p) for p in payloads])
                uris.append((uri, sum(self.backend.count(p) for p in payloads)))
        except Exception:  # noqa: BLE001  (records that the wire format cannot carry: keep the command only)
            uris = []
        cmd = VertexCommand(CMD_START, node.uid, version, args=[node.name, str(part), str(len(ins))], inputs=uris,
                            state=STATE_FAILED, exit_code=1, message=error[:4000])
        with open(stem + ".cmd", "wb") as f:
            f.write(cmd.serialize())
        return stem + ".cmd"

    def run_single_vertex(self, uid: int, part: int, inputs=None, restart_file: Optional[str] = None):
        """Run ONE vertex of this plan outside the graph manager — from explicit inputs or from a
        restart descriptor written by a failed job — and return its output ports."""
        from ..channel import store
        from .command import VertexCommand
        node = next(n for n in self.plan.nodes if n.uid == uid)
        version = 0
        if inputs is None:
            with open(restart_file, "rb") as f:
                cmd = VertexCommand.parse(f.read())
            version = cmd.version + 1
            inputs = [[list(p) for p in store.open_table(u).partitions] for u, _ in cmd.inputs]
        env = VertexEnv(self.context, node.uid, part, version)
        return self.backend.run_vertex(node, part, inputs, env)

    # ---- placement (ClusterInterface + LocalScheduler roles) ---------------------------------------
    def _place(self, v: int, node: DLinqQueryNode, i: int, ins) -> Dict[str, Any]:
        """Choose the computer (GPU / worker slot) for one vertex: every input partition votes for the
        computer that holds it with its record count (DrResources affinity), the votes are merged into
        a computer-, rack- or cluster-level preference (`affinity_merge`), and the native delay
        scheduler matches the vertex with an idle computer (LocalScheduler/LocalScheduler.cs:132-268).
        Returns what the vertex record keeps: where it ran, at which locality, and per input channel
        how many records came from which computer (GraphManager/vertex/DrVertexRecord.cpp)."""
        from ..sched.cluster import Affinity, LocalCluster, Process
        from .command import affinity_merge
        ctx = self.context
        if self.cluster is None:
            n = max(1, min(int(getattr(ctx, "job_max_nodes", 0) or 0) or max(x.partition_count for x in self.plan.nodes), 64))
            self.cluster = LocalCluster(n, getattr(ctx, "work_dir", "/tmp"), rack_delay=0.0, cluster_delay=0.0)
        ncomp = len(self.cluster.computers)
        votes: Dict[int, int] = {}
        channels = []
        for (uid, part), payloads in zip(self._gm_inputs.get(v, []), [p for inp in ins for p in inp]):
            where = self._placement.get((uid, part))
            recs = self.backend.count(payloads)
            channels.append({"from": f"{uid}.{part}", "records": recs, "computer": where})
            if where is not None and recs:
                votes[where] = votes.get(where, 0) + recs
        if votes:
            level, choice = affinity_merge(sorted(votes.items()), [0] * ncomp)
            prefer = [choice] if level == "computer" else []
        else:
            level, prefer = "cluster", [i % ncomp]                    # sources: spread by partition index
        proc = Process(v, f"{node.name}#{node.uid}[{i}]", Affinity(prefer, sum(votes.values()), False))
        self.cluster.schedule_process(proc)
        got = None
        for c in (prefer + [x for x in range(ncomp) if x not in prefer]):
            got = self.cluster.computer_idle(c, age_bias=1e9)
            if got is not None:
                break
        comp = got.computer.index if got is not None and got.computer is not None else (prefer[0] if prefer else 0)
        self._placement[(node.uid, i)] = comp
        local = sum(ch["records"] for ch in channels if ch["computer"] == comp)
        total = sum(ch["records"] for ch in channels)
        return {"computer": comp, "locality": level if votes else "source", "channels": channels,
                "records_local": local, "records_remote": total - local}

    # ---- graph manager wiring -------------------------------------------------------------------
    @staticmethod
    def edge_is_cross_product(node: DLinqQueryNode, child: DLinqQueryNode, op) -> bool:
        """How the edge child -> node is expanded in the graph manager: cross product when the plan
        says so, and conservatively for the operators whose vertices read neighbouring partitions."""
        special = (NodeType.Concat, NodeType.Zip, NodeType.Reverse, NodeType.SlidingWindow)
        head = node.members[0].type if node.type == NodeType.Super else node.type
        cpc, pc = child.partition_count, node.partition_count
        conservative = head in special or not (cpc == pc or cpc == 1 or pc == 1 or (pc and cpc % pc == 0))
        return op == ConnOp.CrossProduct or conservative

    def _build_gm(self):
        """Mirror the physical plan into the native graph manager: one GM stage per plan node, one
        GM vertex per partition, edges expanded by the same connection rules the data follows."""
        from ..sched.gm import CHANNEL_FIFO, CHANNEL_FILE, CROSS_PRODUCT, POINTWISE, GraphManager
        ctx = self.context
        gm = GraphManager(max_failures=ctx.max_failures_per_vertex, speculative=ctx.enable_speculative_duplication)
        # straggler thresholds (DrGraphParameters.cpp:50-69: default outlier threshold 10 min, never below 10 s)
        gm.set_outlier_params(float(getattr(ctx, "duplicate_default_seconds", 600.0)),
                              float(getattr(ctx, "duplicate_min_seconds", 10.0)))
        sid = {}
        self._gm_inputs: Dict[int, List[Tuple[int, int]]] = {}       # gm vertex -> [(producer uid, part)]
        special = (NodeType.Concat, NodeType.Zip, NodeType.Reverse, NodeType.SlidingWindow)
        for node in self.plan.nodes:
            sid[node.uid] = gm.add_stage(f"{node.name}#{node.uid}", node.partition_count)
        for node in self.plan.nodes:
            head = node.members[0].type if node.type == NodeType.Super else node.type
            first = gm.stage_first_vertex[sid[node.uid]]
            for child, op in node.children:
                cpc, pc = child.partition_count, node.partition_count
                kind = CROSS_PRODUCT if self.edge_is_cross_product(node, child, op) else POINTWISE
                # a MemoryFIFO edge streams: producer and consumer vertices form one start clique
                fifo = kind == POINTWISE and cpc == pc and getattr(child.channel_type, "value", "") == "MemoryFIFO"
                gm.connect(sid[child.uid], sid[node.uid], kind, CHANNEL_FIFO if fifo else CHANNEL_FILE)
                for d in range(pc):                      # same expansion order as gm_connect
                    lst = self._gm_inputs.setdefault(first + d, [])
                    if kind == CROSS_PRODUCT:
                        srcs = range(cpc)
                    elif cpc == pc:
                        srcs = [d]
                    elif cpc == 1:
                        srcs = [0]
                    elif pc == 1:
                        srcs = range(cpc)
                    else:
                        g = cpc // pc
                        srcs = range(d * g, d * g + g)
                    lst.extend((child.uid, s) for s in srcs)
        self._gm_stage_node = {sid[n.uid]: n for n in self.plan.nodes}
        return gm

    # ---- run -----------------------------------------------------------------------------------
    def run(self) -> Dict[int, Any]:
        """Drive the job: the native graph manager decides WHAT runs next (dependencies, versions,
        retries, upstream invalidation, abort); this loop moves the data and runs the vertex code."""
        ctx = self.context
        rep = self.reporter
        rep.job_start(self.plan)
        gm = self._build_gm()
        self.gm = gm
        for node in self.plan.nodes:
            self.out[node.uid] = [None] * node.partition_count
        t_job = time.perf_counter()
        gm.start()
        try:
            while True:
                nxt = gm.next_ready(time.perf_counter() - t_job)
                if nxt is None:
                    break
                v, version = nxt
                info = gm.vertex(v)
                node, i = self._gm_stage_node[info["stage"]], info["part"]
                rep.vertex_event(node, i, version, "starting")
                t0 = time.perf_counter()
                ins = []
                try:
                    ins = self._inputs(node, i)
                    placed = self._place(v, node, i, ins) if ctx.schedule_vertices else None
                    if ctx.fault_injector is not None:
                        ctx.fault_injector(node, i, version)
                    env = VertexEnv(ctx, node.uid, i, version)
                    if self.timeline is not None:
                        with self.timeline.span(f"{node.name}#{node.uid}[{i}]", stage=node.uid, part=i, version=version):
                            ports = self.backend.run_vertex(node, i, ins, env)
                    else:
                        ports = self.backend.run_vertex(node, i, ins, env)
                    dt = time.perf_counter() - t0
                    self.out[node.uid][i] = ports
                    n_in = sum(self.backend.count(p) for inp in ins for p in inp)
                    n_out = sum(self.backend.count(p) for p in ports)
                    rec = {"stage": node.uid, "stage_name": node.name, "part": i, "version": version,
                           "seconds": dt, "records_in": n_in, "records_out": n_out}
                    if placed is not None:
                        rec.update(placed)
                    self.stats.append(rec)
                    gm.completed(v, version, dt, float(n_in))
                    rep.vertex_event(node, i, version, "completed", seconds=dt, records_in=n_in, records_out=n_out,
                                     **({"computer": placed["computer"], "locality": placed["locality"],
                                         "local": placed["records_local"], "remote": placed["records_remote"]} if placed else {}))
                    self._release_finished_children(node)
                except Exception as ex:  # noqa: BLE001  (vertex code is user code)
                    text = "".join(traceback.format_exception_only(type(ex), ex)).strip()
                    rep.vertex_event(node, i, version, "failed", error=text,
                                     **({"computer": self._placement.get((node.uid, i))} if (node.uid, i) in self._placement else {}))
                    bad = -1
                    if isinstance(ex, InputReadError):
                        want = (ex.producer_uid, ex.producer_part)
                        edges = self._gm_inputs.get(v, [])
                        bad = edges.index(want) if want in edges else -1
                    rc = gm.failed(v, version, text, bad_input=bad)
                    if rc == 1 or ctx.compile_for_vertex_debugging:
                        try:
                            self._dump_restart(node, i, version, ins, text)
                        except Exception:  # noqa: BLE001  (diagnostics must never mask the failure)
                            pass
                    deterministic = isinstance(ex, DryadLinqException) and not getattr(ex, "transient", False)
                    if rc == 2:                      # producer revoked: it will be re-executed
                        self.out[ex.producer_uid][ex.producer_part] = None
                        continue
                    if rc == 1 or (deterministic and ctx.fail_fast_on_deterministic_errors):
                        rep.job_stop(False, text)
                        if isinstance(ex, DryadLinqException):
                            raise
                        raise DryadLinqException(ErrorCode.DidNotCompleteSuccessfully,
                                                 f"{gm.error() or text}", ex) from ex
            if gm.state != "Succeeded":
                raise DryadLinqException(ErrorCode.DidNotCompleteSuccessfully,
                                         f"job ended in state {gm.state}: {gm.error()}")
            results = self._commit_outputs()
            self._flush_timeline()
            rep.job_stop(True, "", seconds=time.perf_counter() - t_job, **gm.counters())
            return results
        except BaseException:
            self._discard_outputs()
            raise

    def _flush_timeline(self):
        """Resolve the device timeline (one synchronize) and store it next to the event log."""
        if self.timeline is None:
            return
        import os
        spans = self.timeline.resolve()
        self.gpu_timeline = spans
        job_dir = getattr(self.reporter, "job_dir", None)
        if job_dir and spans:
            self.timeline.write(os.path.join(job_dir, "gpu_timeline.jsonl"), spans)

    def _release_finished_children(self, node: DLinqQueryNode):
        """Drop a producer's buffers once every consumer stage has completed (bounded HBM / RAM);
        with `retain_intermediate` they are kept, like the reference's intermediate files, so a
        later failure could re-run a consumer without re-running the producer."""
        if self.context.retain_intermediate:
            return
        from ..utils.collections import BitVector
        done = BitVector(max((n.uid for n in self.plan.nodes), default=0) + 1)       # finished stages, by uid
        for n in self.plan.nodes:
            if n.uid in self.out and all(x is not None for x in self.out[n.uid]):
                done.set(n.uid)
        for child, _ in node.children:
            if child.parents and all(done[p.uid] for p in child.parents):
                self.out[child.uid] = [None] * child.partition_count

    # ---- atomic output commit (DrGraph::FinalizeGraph) ----------------------------------------
    def _commit_outputs(self) -> Dict[int, Any]:
        results: Dict[int, Any] = {}
        for root in self.plan.roots:
            node = root if root.type == NodeType.OutputTable else root.members[-1]
            parts = [self.out[root.uid][i][0] for i in range(root.partition_count)]   # batches stay columnar
            table = self.context._commit_table(node.params.get("uri"), parts, root.schema, root.info,
                                               node.params.get("compression"), node.params.get("temporary", False))
            results[node.params["qexpr_id"]] = table
        return results

    def _discard_outputs(self):
        self.out.clear()
