"""Graph-manager process of a cluster-mode job.

    python -m dryad_b200.sched.jm <job dir> [--workers N]

Role of the reference's job manager executable (DryadLinqGraphManager/Program.cs:30-131 +
LinqToDryadJM.cs:79-259: read the plan, build the graph, run it, exit code = job outcome) with the
pieces it drives: the native graph (csrc/host/sched.cpp: versions, retries, upstream invalidation,
duplicates), the cluster interface + delay scheduler that match vertices with idle computers
(ClusterInterface/Interfaces.cs, LocalScheduler/LocalScheduler.cs:132-268), the process service
whose mailbox carries start commands and status blocks (ProcessService/ProcessService.cs:674-746),
and the message pump that serialises every state change of the graph
(GraphManager/kernel/DrMessagePump.h:23-135).

Threads: one watcher per vertex-host process blocks on that host's status key and on the process
itself; whatever it sees becomes a MESSAGE posted to the pump.  The pump's handler is the only code
that touches the graph: it applies the status (completed / failed / host died), asks the graph for
newly runnable vertices, queues them with the scheduler under their input affinities, and lets idle
hosts claim work.  Channel data never passes through this process: producers write block-stream files
under <job dir>/channels, consumers read them (or their own cache); output tables are written by the
output vertices as uncommitted versions and published here, atomically, only when the whole job
succeeded (DrGraph::FinalizeGraph, GraphManager/filesystem/DrPartitionFile.cpp:561-597).
"""
from __future__ import annotations

import argparse
import json
import os
import signal
import subprocess
import sys
import threading
import time
from collections import deque
from typing import Any, Deque, Dict, List, Optional, Tuple

EV_DISPATCH, EV_STATUS, EV_EXIT, EV_TICK, EV_HELLO = range(5)
STATUS_FILE = "status.json"


class _Host:
    """One vertex-host process slot (a 'computer')."""

    def __init__(self, index: int):
        self.index = index
        self.incarnation = -1
        self.name = ""
        self.proc: Optional[subprocess.Popen] = None
        self.ready = False
        self.busy: Optional[Tuple[List[Tuple[int, int]], float]] = None      # ([(gm vertex, version), ...], start time)
        self.restarts = 0
        self.vertices_run = 0


class _Watcher:
    """IProcessWatcher of the scheduler (ClusterInterface/Interfaces.cs): queued / matched transitions of every
    vertex process go to the event log, so the schedule view shows how long a vertex waited for a computer."""

    def __init__(self, jm):
        self.jm = jm

    def on_queued(self, proc):
        self.jm.reporter._emit("schedule", process=proc.id, what=proc.command_line, state="queued",
                               prefer=" ".join(map(str, proc.affinity.computers)) or None, weight=proc.affinity.weight or None)

    def on_matched(self, proc, computer):
        self.jm.reporter._emit("schedule", process=proc.id, what=proc.command_line, state="matched", computer=computer.index)

    def on_created(self, proc): ...
    def on_started(self, proc): ...

    def on_exited(self, proc, state, exit_code, error):
        self.jm.reporter._emit("schedule", process=proc.id, state="cancelled", error=error)


class JobManager:
    def __init__(self, job_dir: str, workers: int = 0, max_host_restarts: int = 8, app=None):
        from ..runtime import jobpackage
        from ..runtime.executor import JobExecutor
        from ..runtime.process_service import ProcessService
        from ..runtime.services import MessagePump
        from .cluster import LocalCluster
        from .reporting import CalypsoReporter
        self.job_dir = os.path.abspath(job_dir)
        self.ctx, self.plan = jobpackage.load(self.job_dir)
        from .app import DryadLinqApp
        self.app = app or DryadLinqApp()                  # the graph manager's own option table (-mai, -at, -nca, ...)
        self.reporter = CalypsoReporter(self.job_dir, level=self.ctx.runtime_logging_level.value)
        self.ex = JobExecutor(self.ctx, self.plan, reporter=self.reporter)
        self.gm = self.ex._build_gm()
        self.nodes = self.ex._gm_stage_node                                   # gm stage id -> plan node
        self.node_by_uid = {n.uid: n for n in self.plan.nodes}
        self.vid: Dict[Tuple[int, int], int] = {}                             # (stage uid, part) -> gm vertex
        self.vkey: Dict[int, Tuple[int, int]] = {}                            # gm vertex -> (stage uid, part)
        for st, n in self.nodes.items():
            first = self.gm.stage_first_vertex[st]
            for p_ in range(n.partition_count):
                self.vid[(n.uid, p_)] = first + p_
                self.vkey[first + p_] = (n.uid, p_)
        width = max(n.partition_count for n in self.plan.nodes)
        want = workers or int(self.ctx.job_max_nodes or 0) or min(width, os.cpu_count() or 1)
        self.nhosts = max(1, max(int(self.ctx.job_min_nodes or 0), want))
        self.service = ProcessService(self.job_dir)
        self.cluster = LocalCluster(self.nhosts, self.job_dir, rack_delay=0.02, cluster_delay=0.05)
        self.hosts = [_Host(i) for i in range(self.nhosts)]
        self.max_host_restarts = max_host_restarts
        self.lock = threading.RLock()
        self.inbox: Dict[int, Deque[Tuple[int, bytes]]] = {i: deque() for i in range(self.nhosts)}
        self.pending: Dict[int, Tuple[int, int]] = {}                         # scheduler process id -> (gm vertex, version)
        self._next_pid = 1
        self.versions: Dict[Tuple[int, int], int] = {}                        # (stage uid, part) -> completed version
        self.where: Dict[Tuple[int, int], int] = {}                           # (stage uid, part) -> host that holds it in cache
        self.records: Dict[Tuple[int, int], int] = {}                         # (stage uid, part) -> output records
        self.out_bytes: Dict[Tuple[int, int], int] = {}                       # (stage uid, part) -> channel bytes written
        self.stage_of_uid = {n.uid: st for st, n in self.nodes.items()}
        self.part_files: Dict[Tuple[int, int], Tuple[str, str, int]] = {}     # root (uid, part) -> (tmp, final, size)
        self.sent: Dict[Tuple[int, int], bytes] = {}                          # (gm vertex, version) -> start command as sent
        self.done = threading.Event()
        self.ok = False
        self.error = ""
        self.stopping = False
        self.t0 = time.perf_counter()
        self.stats: List[dict] = []
        self.messages = 0
        self._watchers: List[threading.Thread] = []
        self._watcher = _Watcher(self)
        from .dynamic import DynamicManagers
        self.dyn = DynamicManagers(self)
        self.pump = MessagePump(self._on_message, threads=1)                  # ONE consumer: graph changes are serialised

    # ---- vertex hosts -----------------------------------------------------------------------------
    def _launch(self, h: _Host) -> None:
        h.incarnation += 1
        h.name = f"w{h.index}.{h.incarnation}"
        h.ready, h.busy = False, None
        log = open(os.path.join(self.job_dir, "log", f"host-{h.name}.txt"), "wb")
        env = dict(os.environ)
        env["DRYAD_JOB_DIR"] = self.job_dir
        root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        env["PYTHONPATH"] = os.pathsep.join([root] + [p for p in env.get("PYTHONPATH", "").split(os.pathsep) if p and p != root])
        if "DRYAD_THREADS_PER_WORKER" not in self.ctx.job_environment_variables:
            # the hosts share this box: split its cores between them (DRYAD_THREADS_PER_WORKER, VertexEnv.cs:72-86)
            env["DRYAD_THREADS_PER_WORKER"] = str(max(1, (os.cpu_count() or 1) // self.nhosts))
        # hosts live as long as the job: an idle host waits (another host may run a long vertex); it leaves when it
        # is told to, or when the manager's mailbox is gone
        cmd = [sys.executable, "-m", "dryad_b200.runtime.vertex_host", "--service", self.service.uri, "--id", h.name,
               "--job-dir", self.job_dir, "--idle-timeout", "604800"]
        h.proc = subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT, env=env, cwd=self.job_dir)
        with self.service._lock:
            self.service.procs[h.name] = h.proc
        self.reporter._emit("process", host=h.name, pid=h.proc.pid, state="created")
        t = threading.Thread(target=self._watch, args=(h, h.incarnation), daemon=True, name=f"watch-{h.name}")
        self._watchers.append(t)
        t.start()

    def _mbox(self, name: str):
        from ..runtime.services import Mailbox
        with self.service._lock:
            m = self.service.mailboxes.get(name)
            if m is None:
                m = self.service.mailboxes[name] = Mailbox()
            return m

    def _watch(self, h: _Host, incarnation: int) -> None:
        """Blocks on the host's keys; turns what it sees into pump messages.  Never touches the graph."""
        mb = self._mbox(f"w{h.index}.{incarnation}")
        seen, hello = 0, False
        proc = h.proc
        while not self.done.is_set():
            if not hello:
                got = mb.block_on_status("hello", 0, 0.05)
                if got:
                    hello = True
                    self.pump.post(EV_HELLO, h.index * 1000 + incarnation)
            got = mb.block_on_status("status", seen, 0.2 if hello else 0.05)
            if got:
                seen = got[0]
                self.inbox[h.index].append((incarnation, got[2]))
                self.pump.post(EV_STATUS, h.index * 1000 + incarnation)
                continue
            if proc.poll() is not None:
                self.pump.post(EV_EXIT, h.index * 1000 + incarnation)
                return

    # ---- the pump handler: the only place the graph changes -----------------------------------------
    def _on_message(self, listener: int, message: int) -> None:
        try:
            with self.lock:
                self.messages += 1
                if self.done.is_set():
                    return
                if listener in (EV_HELLO, EV_STATUS, EV_EXIT):
                    h = self.hosts[message // 1000]
                    data = self.inbox[h.index].popleft()[1] if listener == EV_STATUS else b""   # one entry per status message
                    if h.incarnation != message % 1000:
                        return                                                # news of a replaced incarnation
                    if listener == EV_HELLO:
                        h.ready = True
                        self.reporter._emit("process", host=h.name, state="started")
                    elif listener == EV_STATUS:
                        self._on_status(h, data)
                    else:
                        self._on_exit(h)
                elif listener == EV_TICK:
                    self.gm.check_duplicates(time.perf_counter() - self.t0)
                    limit_v = float(getattr(self.ctx, "vertex_timeout_seconds", 0) or 0)
                    if limit_v > 0:
                        # a vertex that runs beyond the limit is treated like a dead host: its process is killed (the
                        # watcher reports the exit, the attempt fails, the host is replaced, the vertex runs again)
                        for h_ in self.hosts:
                            if h_.busy is not None and h_.proc is not None and h_.proc.poll() is None \
                                    and time.perf_counter() - h_.busy[1] > limit_v:
                                self.reporter._emit("process", host=h_.name, state="killed", reason=f"vertex ran longer than {limit_v:g} s")
                                h_.proc.kill()
                    limit = self.ctx.job_runtime_limit
                    if limit and time.perf_counter() - self.t0 > float(limit):
                        self._finish(False, f"job exceeded its runtime limit of {limit} s")
                        return
                    self.pump.post(EV_TICK, 0, 0.25)
                self._dispatch()
                st = self.gm.state
                if st == "Succeeded":
                    self._commit()
                elif st == "Failed":
                    self._finish(False, self.error or self.gm.error() or "job failed")
        except Exception as ex:  # noqa: BLE001  (a bug here must end the job, not hang it)
            import traceback
            self._finish(False, "graph manager error: " + "".join(traceback.format_exception(type(ex), ex, ex.__traceback__))[-3000:])

    def _vertex(self, v: int):
        info = self.gm.vertex(v)
        return self.nodes[info["stage"]], info["part"]

    def _on_status(self, h: _Host, data: bytes) -> None:
        import json
        from ..runtime.command import STATE_COMPLETED, VertexCommand
        st = VertexCommand.parse(data)
        if st is None or h.busy is None:
            return
        attempts, _t_start = h.busy
        if st.args and st.args[0] == "subgraph":
            results = json.loads(st.args[1])
            if {(r["gmv"], r["version"]) for r in results} != set(attempts):
                return                                                        # stale
            h.busy = None
            # a streamed edge keeps nothing to re-read: one member's failure restarts the gang, so the failure is
            # applied first and the completions of the same attempt then lose (they are no longer running)
            failed = [r for r in results if not r["ok"]]
            own = [r for r in failed if not r["message"].startswith("dryad_b200.runtime.executor.InputReadError: aborted")]
            for r in (own or failed)[:1]:
                self._apply_result(h, r)
            for r in results:
                if r["ok"]:
                    self._apply_result(h, r)
            return
        v, version = attempts[0]
        if int(st.args[1]) != v or st.version != version:
            return                                                            # stale
        h.busy = None
        part_file = next(((pth.split("|", 1)[0], pth.split("|", 1)[1], sz) for pth, sz in st.outputs if "|" in pth), None)
        self._apply_result(h, {
            "gmv": v, "version": version, "ok": st.state == STATE_COMPLETED, "n_in": int(st.args[2]), "n_out": int(st.args[3]),
            "seconds": float(st.args[4]), "hits": int(st.args[5]), "remote_reads": int(st.args[6]) if len(st.args) > 6 else 0, "bytes_in": [b for b, _ in st.channel_status],
            "bad": next((k for k, (_b, err) in enumerate(st.channel_status) if err), -1), "message": st.message,
            "out_bytes": sum(sz for pth, sz in st.outputs if "|" not in pth), "out_sizes": [sz for pth, sz in st.outputs if "|" not in pth],
            "part_file": part_file, "deterministic": st.exit_code == 2})

    def _apply_result(self, h: _Host, r: dict) -> None:
        """One vertex attempt's outcome -> the graph, the bookkeeping, the event log."""
        v, version = int(r["gmv"]), int(r["version"])
        h.vertices_run += 1
        node, part = self._vertex(v)
        n_in, n_out, seconds, hits = r["n_in"], r["n_out"], r["seconds"], r["hits"]
        if r["ok"]:
            won = self.gm.completed(v, version, seconds, float(sum(r.get("bytes_in", []))))
            if won:
                self.versions[(node.uid, part)] = version
                self.where[(node.uid, part)] = h.index
                self.records[(node.uid, part)] = n_out
                self.out_bytes[(node.uid, part)] = r["out_bytes"]
                if r.get("part_file"):
                    tmp, final, size = r["part_file"]
                    self.part_files[(node.uid, part)] = (tmp, final, size)
                self.dyn.on_completed(v, h.index)                             # run-time rewrites, before anything is dispatched
            ins = self.ex._gm_inputs.get(v, [])
            local = sum(self.records.get(e, 0) for e in ins if self.where.get(e) == h.index)
            total = sum(self.records.get(e, 0) for e in ins)
            rec = {"stage": node.uid, "stage_name": node.name, "part": part, "version": version, "seconds": seconds,
                   "records_in": n_in, "records_out": n_out, "computer": h.index, "host": h.name, "won": bool(won),
                   "records_local": local, "records_remote": total - local, "cache_hits": hits,
                   "remote_reads": r.get("remote_reads", 0), "bytes_in": r.get("bytes_in", []), "bytes_out": r.get("out_sizes", [r["out_bytes"]])}
            self.stats.append(rec)
            self.reporter.vertex_event(node, part, version, "completed" if won else "completed_duplicate", seconds=seconds,
                                       records_in=n_in, records_out=n_out, computer=h.index, local=local,
                                       remote=total - local, cache_hits=hits)
        else:
            bad = r["bad"]
            self.reporter.vertex_event(node, part, version, "failed", error=r["message"], computer=h.index)
            rc = self.gm.failed(v, version, r["message"], bad_input=bad)
            if rc == 1 or self.ctx.compile_for_vertex_debugging:
                self._dump_restart(node, part, v, version, r["message"])
            if rc == 2 and bad >= 0:
                prod = self.ex._gm_inputs[v][bad]
                self.versions.pop(prod, None)
                self.where.pop(prod, None)
            elif rc == 1 or (r["deterministic"] and self.ctx.fail_fast_on_deterministic_errors):
                self.error = r["message"]
                if rc != 1:
                    self._finish(False, r["message"])

    def _dump_restart(self, node, part: int, v: int, version: int, error: str) -> None:
        """The start command of a finally-failed vertex, as it was sent (DumpRestartCommand,
        dvertexpncontrol.h:72): its inputs are channel files of this job directory, so
        `python -m dryad_b200.tools.rerun <job dir> <stage>.<partition>` can run the vertex alone."""
        wire = self.sent.get((v, version))
        if wire is None:
            return
        try:
            d = os.path.join(self.job_dir, "restart")
            os.makedirs(d, exist_ok=True)
            with open(os.path.join(d, f"vertex-{node.uid}-{part}.cmd"), "wb") as f:
                f.write(wire)
            with open(os.path.join(d, f"vertex-{node.uid}-{part}.error.txt"), "w") as f:
                f.write(error)
        except OSError:
            pass                                           # diagnostics must never mask the failure

    def _on_exit(self, h: _Host) -> None:
        code = h.proc.poll() if h.proc is not None else None
        self.reporter._emit("process", host=h.name, state="exited", exitcode=code)
        if self.stopping:
            return
        if h.busy is not None:
            attempts, _ = h.busy
            # members of a sub-graph vertex are one gang: failing one restarts them all
            for v, version in (attempts[:1] if len(attempts) > 1 else attempts):
                node, part = self._vertex(v)
                msg = f"vertex host {h.name} exited with code {code} while running {node.name}#{node.uid}[{part}] v{version}"
                self.reporter.vertex_event(node, part, version, "failed", error=msg, computer=h.index)
                if self.gm.failed(v, version, msg) == 1:
                    self.error = msg
            h.busy = None
        for key in [k for k, w in self.where.items() if w == h.index]:          # its cache died with it; the files remain
            del self.where[key]
        h.restarts += 1
        if h.restarts > self.max_host_restarts:
            tail = ""
            try:
                with open(os.path.join(self.job_dir, "log", f"host-{h.name}.txt"), errors="replace") as f:
                    tail = f.read()[-1500:]
            except OSError:
                pass
            self.error = f"vertex host slot {h.index} died {h.restarts} times (last exit code {code}); its log ends with:\n{tail}"
            self._finish(False, self.error)
            return
        self._launch(h)

    def enqueue(self, attempts) -> None:
        """Queue vertex attempts with the scheduler as ONE process — a single vertex, or the members of a
        start clique that run together as a sub-graph vertex — with the input affinities of all of them."""
        from ..runtime.command import affinity_merge
        from .cluster import Affinity, Process
        if isinstance(attempts, tuple):
            attempts = [attempts]
        members = {self.vkey.get(v) for v, _ in attempts}
        votes: Dict[int, int] = {}
        for v, _version in attempts:
            for e in self.ex._gm_inputs.get(v, []):
                w = self.where.get(e)
                if w is not None and e not in members:
                    votes[w] = votes.get(w, 0) + max(1, self.records.get(e, 0))
        prefer: List[int] = []
        if votes:
            level, choice = affinity_merge(sorted(votes.items()), [0] * self.nhosts)
            if level == "computer":
                prefer = [choice]
        pid = self._next_pid
        self._next_pid += 1
        self.pending[pid] = list(attempts)
        node, part = self._vertex(attempts[0][0])
        label = f"{node.name}#{node.uid}[{part}]" + (f" +{len(attempts) - 1}" if len(attempts) > 1 else "")
        self.cluster.schedule_process(Process(pid, label, Affinity(prefer, sum(votes.values()), False), self._watcher))
        for v, version in attempts:
            n_, p_ = self._vertex(v)
            self.reporter.vertex_event(n_, p_, version, "created", prefer=prefer[0] if prefer else None,
                                       gang=len(attempts) if len(attempts) > 1 else None)

    def rebuild_maps(self) -> None:
        """After a run-time resize the native graph is the truth: vertex ids and input lists are re-read."""
        self.stage_of_uid = {n.uid: st for st, n in self.nodes.items()}
        for st, n in self.nodes.items():
            for idx, vid in enumerate(self.gm.stage_vertices(st)):
                self.vid[(n.uid, idx)] = vid
                self.vkey[vid] = (n.uid, idx)
        for vid in list(self.vkey):
            if self.vid.get(self.vkey[vid]) == vid:
                self.ex._gm_inputs[vid] = [self.vkey[p] for p in self.gm.vertex_inputs(vid)]

    def _dispatch(self) -> None:
        from ..runtime.command import CMD_START, VertexCommand
        now = time.perf_counter() - self.t0
        gangs: Dict[int, List[Tuple[int, int]]] = {}
        while True:
            nxt = self.gm.next_ready(now)
            if nxt is None:
                break
            v, version = nxt
            if self.dyn.hold(v, version):
                continue
            clique = self.gm.clique(v)
            if len(clique) > 1:
                gangs.setdefault(min(clique), []).append((v, version))       # a start clique runs as ONE sub-graph vertex
            else:
                self.enqueue([(v, version)])
        for members in gangs.values():
            self.enqueue(sorted(members))
        waiting = self.cluster.scheduler.waiting
        if not waiting:
            return
        idle = [h for h in self.hosts if h.ready and h.busy is None and h.proc is not None and h.proc.poll() is None]
        for h in idle:
            p = self.cluster.computer_idle(h.index)
            if p is None:
                continue
            attempts = self.pending.pop(p.id)
            v, version = attempts[0]
            node, part = self._vertex(v)
            if len(attempts) == 1:
                inputs = [(f"chan://{uid}/{s}/{self.versions.get((uid, s), 0)}", self.records.get((uid, s), 0))
                          for uid, s in self.ex._gm_inputs.get(v, [])]
                args = [node.name, str(part), str(v)] + self.dyn.command_args(v)
            else:
                import json
                members, inputs = [], []
                for mv, mver in attempts:
                    mn, mp = self._vertex(mv)
                    mins = [[uid, s, self.versions.get((uid, s), 0)] for uid, s in self.ex._gm_inputs.get(mv, [])]
                    members.append({"uid": mn.uid, "part": mp, "gmv": mv, "version": mver, "inputs": mins})
                    inputs += [(f"chan://{u}/{s}/{ver}", self.records.get((u, s), 0)) for u, s, ver in mins]
                args = [node.name, str(part), str(v), "subgraph", json.dumps(members)] + \
                       [a for a in self.dyn.command_args(v) if a.startswith("widths=")]
            cmd = VertexCommand(CMD_START, node.uid, version, args=args, inputs=inputs,
                                outputs=[(f"chan://{node.uid}/{part}/{version}", 0)])
            h.busy = (list(attempts), time.perf_counter())
            wire = cmd.serialize()
            for mv, mver in attempts:
                self.sent[(mv, mver)] = wire
            self._mbox(h.name).set("cmd", "start", wire)
            for mv, mver in attempts:
                mn, mp = self._vertex(mv)
                self.reporter.vertex_event(mn, mp, mver, "starting", computer=h.index)
        if self.cluster.scheduler.waiting and any(h.ready and h.busy is None for h in self.hosts):
            self.pump.post(EV_DISPATCH, 0, 0.03)                              # delay scheduling: let preferences age

    # ---- end of job ---------------------------------------------------------------------------------
    def _commit(self) -> None:
        """Every vertex completed: publish the output tables (rename the winning part versions, then the
        metadata), write the job's status file."""
        from ..channel import store
        from ..runtime.vertex_host import read_channel, temp_output_uri
        tables: Dict[str, Any] = {}
        try:
            for root in self.plan.roots:
                onode = root if root.type.value == "OutputTable" else root.members[-1]
                uri = onode.params.get("uri") or temp_output_uri(self.job_dir, onode.params["qexpr_id"])
                comp = onode.params.get("compression") or self.ctx.output_data_compression_scheme
                if store.supports_part_commit(uri):
                    sizes = []
                    for i in range(root.partition_count):
                        tmp, final, size = self.part_files[(root.uid, i)]
                        store.commit_part(tmp, final, uri)
                        sizes.append(size)
                    store.finish_table(uri, sizes, root.schema, comp)
                else:
                    parts = [read_channel(self.job_dir, root.uid, i, 0, self.versions[(root.uid, i)])[0]
                             for i in range(root.partition_count)]
                    store.write_table(uri, parts, root.schema, comp)
                tables[str(onode.params["qexpr_id"])] = {"uri": uri, "temporary": bool(onode.params.get("temporary", False)),
                                                         "partitions": root.partition_count}
        except Exception as ex:  # noqa: BLE001
            self._finish(False, f"output commit failed: {type(ex).__name__}: {ex}")
            return
        self._remove_uncommitted()                       # part versions of duplicate attempts that lost
        self._finish(True, "", tables)

    def _finish(self, ok: bool, error: str, tables: Optional[dict] = None) -> None:
        if self.done.is_set():
            return
        self.stopping = True
        if not ok:
            for tmp, _final, _ in self.part_files.values():                  # failed jobs leave no partial outputs
                try:
                    os.remove(tmp)
                except OSError:
                    pass
        counters = self.gm.counters()
        self.reporter.job_stop(ok, error, seconds=time.perf_counter() - self.t0, **counters)
        status = {"state": "Success" if ok else "Failure", "error": error, "tables": tables or {}, "counters": counters,
                  "hosts": self.nhosts, "host_restarts": sum(h.restarts for h in self.hosts),
                  "pump_messages": self.messages, "dynamic": dict(self.dyn.spliced), "seconds": time.perf_counter() - self.t0, "vertices": self.stats}
        tmp = os.path.join(self.job_dir, STATUS_FILE + ".tmp")
        with open(tmp, "w") as f:
            json.dump(status, f)
        os.replace(tmp, os.path.join(self.job_dir, STATUS_FILE))
        self.ok = ok
        self.done.set()

    def _remove_uncommitted(self) -> None:
        """Failed and cancelled jobs leave no partial outputs: drop every uncommitted part version."""
        import glob
        from ..channel import store
        from ..runtime.vertex_host import temp_output_uri
        for root in self.plan.roots:
            onode = root if root.type.value == "OutputTable" else root.members[-1]
            uri = onode.params.get("uri") or temp_output_uri(self.job_dir, onode.params["qexpr_id"])
            try:
                _, path = store.parse_uri(uri)
            except Exception:  # noqa: BLE001
                continue
            for f in glob.glob(glob.escape(os.path.splitext(path)[0]) + ".*.tmp"):
                try:
                    os.remove(f)
                except OSError:
                    pass

    def run(self) -> int:
        os.makedirs(os.path.join(self.job_dir, "log"), exist_ok=True)
        self.reporter.job_start(self.plan)
        # which plan vertex became which graph-manager stage (DryadLinqApplication.cs identity map, `<plan>.map`)
        self.app.open_identity_map(os.path.join(self.job_dir, "DryadLinqProgram__0.xml"))
        for st, n in sorted(self.nodes.items()):
            self.app.record_identity(n.uid, st, n.name)
        for h in self.hosts:
            self._launch(h)
        self.gm.start()
        self.pump.post(EV_DISPATCH, 0)
        self.pump.post(EV_TICK, 0, 0.25)
        try:
            self.done.wait()
        finally:
            self.shutdown()
        return 0 if self.ok else 1

    def cancel(self, why: str = "cancelled") -> None:
        with self.lock:
            self._finish(False, why)

    def shutdown(self) -> None:
        from ..runtime.command import terminate_command
        self.stopping = True
        data = terminate_command().serialize()
        for h in self.hosts:
            if h.proc is not None and h.proc.poll() is None:
                self._mbox(h.name).set("cmd", "terminate", data)
        deadline = time.monotonic() + 5.0
        for h in self.hosts:
            if h.proc is None:
                continue
            try:
                h.proc.wait(max(0.05, deadline - time.monotonic()))
            except subprocess.TimeoutExpired:
                h.proc.kill()                                                 # the exact process we started
        self._remove_uncommitted()                                            # a host may have written a part version after the verdict
        if self.ok and not self.ctx.retain_intermediate:
            # intermediate channels and spill files are temporary data of the job (the reference's vertices write
            # them as temp files of the job directory); kept by default for inspection and single-vertex replay
            import shutil
            for sub in ("channels", "spill"):
                shutil.rmtree(os.path.join(self.job_dir, sub), ignore_errors=True)
        for t in self._watchers:                                              # nobody may post to a pump that is gone
            t.join(5.0)
        self.pump.close()
        self.service.close()
        self.reporter.close()


def main(argv: Optional[List[str]] = None) -> int:
    ap = argparse.ArgumentParser(prog="jm")
    ap.add_argument("job_dir")
    ap.add_argument("--workers", type=int, default=0)
    a, rest = ap.parse_known_args(argv)
    from .app import DryadLinqApp
    try:
        app = DryadLinqApp.parse(rest)                    # -mai / -mafi / -at / -nca and the legacy switches
    except ValueError as ex:
        print(f"{ex}\noptions:\n{DryadLinqApp.usage()}", file=sys.stderr)
        return 2
    jm = JobManager(a.job_dir, a.workers, app=app)
    signal.signal(signal.SIGTERM, lambda *_: jm.cancel("cancelled (SIGTERM)"))
    return jm.run()


if __name__ == "__main__":
    sys.exit(main())
