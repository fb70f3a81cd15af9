"""Vertex host: the worker process of a cluster-mode job.

    python -m dryad_b200.runtime.vertex_host --service http://127.0.0.1:PORT --id w3.0 --job-dir DIR

Role of the reference's VertexHost.exe (DryadVertex/VertexHost/vertexHost/main.cpp:110-298 — argument
parsing, `--startfrompn`, exit code) with its controller loop
(DryadVertex/VertexHost/system/dprocess/src/dvertexpncontrol.cpp: long-poll the process-node
mailbox for a DVertexCommandBlock, Start / Terminate, status blocks back) and of the managed wrapper
vertex that runs the user's code between channel readers and writers
(DryadVertex/VertexHost/system/managedwrapper/ManagedWrapperVertex.cpp).

One host process serves many vertices one after another (process reuse, as the reference's
`--startfrompn` hosts do).  For every START command it
  1. opens the input channels named in the command — block-stream files written by the producers, or
     this host's own cache when the producer ran here (ProcessService/Cache.cs role),
  2. runs the vertex's operator code from the job package on them,
  3. writes one channel file per output port as an uncommitted version, renames it into place, and —
     for a vertex that ends in an output table — writes its table partition as a version file that only
     the graph manager's commit makes visible,
  4. reports a status block: state, per-input-channel byte counts and error codes, output sizes.
An input that cannot be read is reported with the index of the failing channel so that the graph
manager re-executes its producer (DrGraph.cpp:392-415).
"""
from __future__ import annotations

import argparse
import os
import pickle
import sys
import time
import traceback
from typing import Any, Dict, List, Optional, Tuple

CHAN_DIR = "channels"
DYN_UID_BASE = 1_000_000          # stage ids of vertices created at run time (sched/dynamic.py)


def channel_path(job_dir: str, uid: int, part: int, port: int, version: int, kind: str = "chan") -> str:
    """<job dir>/channels/<stage>_<partition>_<port>.v<version>.<chan|pkl>"""
    return os.path.join(job_dir, CHAN_DIR, f"{uid}_{part}_{port}.v{version}.{kind}")


def write_ports(job_dir: str, uid: int, part: int, version: int, ports, compression) -> List[Tuple[str, int, int]]:
    """Write every output port of a vertex through ONE multi-port writer (channel/factory.VertexWriter,
    the reference's DryadLinqVertexWriter): per port (path, stored bytes, records).  Records that the wire
    format carries go out as block streams (optionally gzip, the intermediate-data compression setting);
    a port whose records it cannot carry, or that holds raw text, is pickled instead.  Files appear under
    their final names only when complete (write to *.tmp, rename)."""
    from ..api.errors import DryadLinqException
    from ..channel.factory import VertexWriter
    os.makedirs(os.path.join(job_dir, CHAN_DIR), exist_ok=True)
    finals = [channel_path(job_dir, uid, part, k, version, "chan") for k in range(len(ports))]
    w = VertexWriter([f"file://{f}.tmp" for f in finals], compression)
    out: List[Tuple[str, int, int]] = []
    pickled: Dict[int, Any] = {}
    for k, payload in enumerate(ports):
        raw = getattr(payload, "is_text_partition", False)                    # raw text partition: keep the bytes
        if raw:
            pickled[k] = payload
            continue
        try:
            w.write_item_sequence(payload.to_records() if hasattr(payload, "to_records") else payload, k)
        except DryadLinqException:
            pickled[k] = payload                  # TypeNotSerializable: user classes without a serializer
    w.close()
    for k, payload in enumerate(ports):
        tmp = finals[k] + ".tmp"
        if k in pickled:
            os.remove(tmp)
            final = channel_path(job_dir, uid, part, k, version, "pkl")
            import cloudpickle                    # classes defined inside a function travel by value
            with open(final + ".tmp", "wb") as f:
                cloudpickle.dump(payload, f, protocol=pickle.HIGHEST_PROTOCOL)
            os.replace(final + ".tmp", final)
            out.append((final, os.path.getsize(final), _count(payload)))
        else:
            os.replace(tmp, finals[k])
            out.append((finals[k], getattr(w.ports[k], "stored_bytes", 0), w.ports[k].records))
    return out


def read_channel(job_dir: str, uid: int, part: int, port: int, version: int):
    """-> (payload, stored bytes); FileNotFoundError / a decode error mean the channel is unreadable."""
    from ..channel.factory import ChannelReader
    p = channel_path(job_dir, uid, part, port, version, "chan")
    if os.path.exists(p):
        return list(ChannelReader(f"file://{p}")), os.path.getsize(p)
    p = channel_path(job_dir, uid, part, port, version, "pkl")
    with open(p, "rb") as f:
        return pickle.load(f), os.path.getsize(p)


def _count(payload) -> int:
    if getattr(payload, "is_text_partition", False) and getattr(payload, "_lines", 1) is None:
        if getattr(payload, "_data", b"") is None:
            return payload.nbytes // 32 + 1           # not read yet: an estimate is enough for placement weights
        return payload.line_count()
    try:
        return len(payload)
    except TypeError:
        return 0


class _NoReport:
    job_dir = None

    def job_start(self, *a, **k): ...
    def vertex_event(self, *a, **k): ...
    def job_stop(self, *a, **k): ...


class _LazyPorts:
    def __init__(self, run, uid, part):
        self.run, self.uid, self.part = run, uid, part

    def __getitem__(self, port):
        return self.run.load_input(self.uid, self.part, port)


class _LazyStage:
    """Stands where the in-process executor keeps `out[stage uid]`: a list of partitions, each a list
    of ports — here every access opens the channel on demand."""

    def __init__(self, run, uid, nparts):
        self.run, self.uid, self.nparts = run, uid, nparts

    def __len__(self):
        return self.nparts

    def __getitem__(self, part):
        if part < 0:
            part += self.nparts
        if not 0 <= part < self.nparts:
            raise IndexError(part)
        return _LazyPorts(self.run, self.uid, part)

    def __iter__(self):
        return (self[i] for i in range(self.nparts))


class _Run:
    """ONE vertex attempt: which channel versions it reads, what a run-time rewrite substituted, which
    inputs and outputs are in-process FIFOs (members of a sub-graph vertex run concurrently, so nothing
    of this may live on the host object)."""

    def __init__(self, host, member: dict, subst: Dict[str, Optional[str]], internal=None):
        self.host = host
        self.uid, self.part = int(member["uid"]), int(member["part"])
        self.gmv, self.version = int(member["gmv"]), int(member["version"])
        self.subst = subst
        self.internal = internal
        self.versions: Dict[Tuple[int, int], int] = {}
        self.input_index: Dict[Tuple[int, int], int] = {}
        for k, (u, sp, ver) in enumerate(member["inputs"]):
            self.versions[(int(u), int(sp))] = int(ver)
            self.input_index[(int(u), int(sp))] = k
        self.input_bytes = [0] * len(member["inputs"])
        self.cache_hits = 0
        self.remote_reads = 0
        self.fifo_in: Dict[Tuple[int, int], str] = {}        # (producer uid, part) -> fifo name
        self.fifo_out: List[str] = []
        self._fifo_payload: Dict[Tuple[int, int], Any] = {}

    def load_input(self, uid: int, part: int, port: int):
        from .executor import InputReadError
        host = self.host
        hops = 0
        while f"{uid}/{part}" in self.subst and hops < 64:         # run-time rewrites: follow the splice chain
            t = self.subst[f"{uid}/{part}"]
            if t is None:
                return []                                          # folded into an internal vertex's output
            uid, part, port = int(t.split("/")[0]), int(t.split("/")[1]), 0
            hops += 1
        src = host.nodes.get(uid)
        if src is not None and src.type.value == "InputTable":
            # a stored (or packaged) table is read where it is consumed: no channel file between the source and its reader
            payload = host.ctx._read_table_partition(src.params["table"], part)
            idx = self.input_index.get((uid, part), -1)
            if idx >= 0:
                self.input_bytes[idx] += payload.nbytes if getattr(payload, "is_text_partition", False) else 0
            return payload
        name = self.fifo_in.get((uid, part))
        if name is not None:
            if (uid, part) not in self._fifo_payload:
                self._fifo_payload[(uid, part)] = host.read_fifo(name, uid, part)
            return self._fifo_payload[(uid, part)]
        ver = self.versions.get((uid, part))
        if ver is None:
            raise InputReadError(uid, part, f"stage {uid}[{part}] is not an input of this vertex")
        key = f"{uid}_{part}_{port}.v{ver}"
        idx = self.input_index.get((uid, part), -1)
        n = host._cached.get(key)
        if n is not None:
            import ctypes as C
            buf = C.create_string_buffer(n)
            got = host._hostlib.cache_get(host._cache, key.encode(), buf, n)
            if got == n:
                self.cache_hits += 1
                if idx >= 0:
                    self.input_bytes[idx] += n
                return pickle.loads(buf.raw)
        try:
            if host.remote_reads:
                payload, size = host.fetch_channel(uid, part, port, ver)
                self.remote_reads += 1
            else:
                payload, size = read_channel(host.job_dir, uid, part, port, ver)
        except Exception as ex:  # noqa: BLE001  (missing, truncated or corrupt: all "cannot read")
            raise InputReadError(uid, part, f"cannot read channel {key}: {type(ex).__name__}: {ex}") from ex
        if idx >= 0:
            self.input_bytes[idx] += size
        return payload


class VertexHost:
    def __init__(self, service_uri: str, host_id: str, job_dir: str):
        from . import jobpackage
        from .executor import JobExecutor
        from .process_service import MailboxClient
        self.id, self.job_dir = host_id, job_dir
        self.client = MailboxClient(service_uri, f"vertex host {host_id}")
        self.ctx, self.plan = jobpackage.load(job_dir)
        self.ex = JobExecutor(self.ctx, self.plan, reporter=_NoReport())
        self.nodes = {n.uid: n for n in self.plan.nodes}
        self.backend = self.ex.backend
        self.roots = {r.uid: r for r in self.plan.roots}
        self.compression = self.ctx.intermediate_data_compression_scheme
        self._handoff: Dict[str, Any] = {}
        # channels produced on ANOTHER computer are fetched from that node's daemon over HTTP, in ranges
        # (managedchannel HttpReader + ProcessService file server); on one box a shared directory serves as well
        self.remote_reads = bool(getattr(self.ctx, "remote_channel_reads", False))
        # channel cache with spill: outputs of vertices that ran HERE are served from memory
        from .. import _native as N
        self._hostlib = N.host()
        low = int(os.environ.get("DRYAD_CACHE_LOW_WATER", str(64 << 20)))
        high = int(os.environ.get("DRYAD_CACHE_HIGH_WATER", str(128 << 20)))
        spill = os.path.join(job_dir, "spill", host_id)
        os.makedirs(spill, exist_ok=True)
        self._cache = self._hostlib.cache_create(spill.encode(), low, high)
        self._cached: Dict[str, int] = {}
        import threading
        self._cache_lock = threading.Lock()

    # ---- channels -------------------------------------------------------------------------------
    def _cache_put(self, key: str, payload) -> None:
        try:
            b = pickle.dumps(payload, protocol=pickle.HIGHEST_PROTOCOL)
        except Exception:  # noqa: BLE001
            return
        with self._cache_lock:
            if len(b) <= (32 << 20) and self._hostlib.cache_put(self._cache, key.encode(), b, len(b)) == 0:
                self._cached[key] = len(b)

    def fetch_channel(self, uid: int, part: int, port: int, version: int, piece: int = 4 << 20):
        """Read a channel through the process service's file server: ranged GETs into a scratch file."""
        scratch = os.path.join(self.job_dir, "spill", self.id)
        for kind in ("chan", "pkl"):
            name = os.path.basename(channel_path(self.job_dir, uid, part, port, version, kind))
            local = os.path.join(scratch, f"fetch-{name}")
            try:
                size = 0
                with open(local, "wb") as f:
                    while True:
                        data = self.client.fetch_file(f"{CHAN_DIR}/{name}", size, size + piece - 1)
                        f.write(data)
                        size += len(data)
                        if len(data) < piece:
                            break
            except FileNotFoundError:
                if os.path.exists(local):
                    os.remove(local)
                continue
            try:
                if kind == "chan":
                    from ..channel.factory import ChannelReader
                    return list(ChannelReader(f"file://{local}")), size
                with open(local, "rb") as f:
                    return pickle.load(f), size
            finally:
                os.remove(local)
        raise FileNotFoundError(f"channel {uid}_{part}_{port}.v{version} is not served by {self.client.uri}")

    def write_fifo(self, name: str, payload) -> None:
        """Stream one output to a member of the same sub-graph vertex through a bounded in-process FIFO
        (fifo://32/..., channelfifo.cpp: the writer blocks while 32 blocks are outstanding)."""
        from ..api.errors import DryadLinqException
        from ..channel.factory import ChannelWriter
        w = ChannelWriter(f"fifo://32/{name}", timeout=3600.0)
        try:
            for rec in (payload.to_records() if hasattr(payload, "to_records") else payload):
                w.write(rec)
        except DryadLinqException:
            self._handoff[name] = payload          # records the wire format cannot carry: handed over as objects
        finally:
            w.close()

    def read_fifo(self, name: str, uid: int, part: int):
        from ..channel.factory import ChannelReader
        from .executor import InputReadError
        try:
            got = list(ChannelReader(f"fifo://32/{name}", timeout=3600.0))
        except (BrokenPipeError, TimeoutError) as ex:
            raise InputReadError(uid, part, f"aborted: the producer of FIFO {name} failed ({type(ex).__name__})") from ex
        return self._handoff.pop(name, got)

    def _apply_widths(self, w: dict) -> None:
        """Widths decided at run time by the dynamic distributor."""
        for k, n in w.get("widths", {}).items():
            self.nodes[int(k)].partition_count = int(n)
        for k, n in w.get("nports", {}).items():
            d = self.nodes[int(k)]
            tail = d.members[-1] if getattr(d, "members", None) else d
            tail.params["nports"] = int(n)
            d.n_out_ports = int(n)
        for k, n in w.get("pcount", {}).items():
            d = self.nodes[int(k)]
            (d.members[-1] if getattr(d, "members", None) else d).params["pcount"] = int(n)

    # ---- one vertex -----------------------------------------------------------------------------
    def run_member(self, run: _Run) -> dict:
        """Run one vertex attempt; returns its result record (what the status block carries)."""
        import copy
        from ..channel import store
        from ..channel.factory import abort_fifo
        from .executor import InputReadError
        from .vertex_ops import VertexEnv
        uid, part, version = run.uid, run.part, run.version
        internal = run.internal
        res = {"uid": uid, "part": part, "gmv": run.gmv, "version": version, "ok": False, "n_in": 0, "n_out": 0,
               "seconds": 0.0, "hits": 0, "bytes_in": [], "bad": -1, "message": "", "out_bytes": 0, "out_sizes": [],
               "part_file": None, "deterministic": False}
        ex = copy.copy(self.ex)                                    # its `out` is this attempt's view of the channels
        ex.out = {n.uid: _LazyStage(run, n.uid, n.partition_count) for n in self.plan.nodes}
        t0 = time.perf_counter()
        try:
            env = VertexEnv(self.ctx, uid, part, version)
            if internal is not None:
                # an internal vertex of a run-time rewrite: fold / copy exactly the channels named in the command
                from . import vertex_ops
                subst, run.subst = run.subst, {}
                # producers of the plan are read at the consumer's port, outputs of other internal vertices at port 0
                payloads = [run.load_input(u, sp, int(internal[2]) if u < DYN_UID_BASE else 0) for (u, sp) in run.input_index]
                run.subst = subst
                ins = [payloads]
                if internal[0] == "copy":
                    ports = [payloads[0]]
                else:
                    ports = [vertex_ops.internal_combine(self.nodes[int(internal[1])], [self.backend.to_records(p) for p in payloads], env)]
            elif self.nodes[uid].type.value == "InputTable":
                # a source is read where it is consumed; its vertex only reports the partition's size and
                # record count (estimated for data that has not been read), which the manager uses for
                # placement weights and for sizing dynamic distributors
                table = self.nodes[uid].params["table"]
                parts_ = table.partitions
                if hasattr(parts_, "nbytes"):
                    size = parts_.nbytes(part)
                    res["n_out"] = size // 32 + 1
                else:
                    payload = self.ctx._read_table_partition(table, part)
                    if getattr(payload, "is_text_partition", False):
                        size = payload.nbytes
                    else:
                        try:
                            size = len(pickle.dumps(payload, protocol=pickle.HIGHEST_PROTOCOL))
                        except Exception:  # noqa: BLE001
                            size = 64 * _count(payload)
                    res["n_out"] = _count(payload)
                res["out_bytes"], res["out_sizes"], res["ok"] = size, [size], True
                res["seconds"], res["remote_reads"] = time.perf_counter() - t0, 0
                return res
            else:
                node = self.nodes[uid]
                ins = ex._inputs(node, part)
                if self.ctx.fault_injector is not None:
                    self.ctx.fault_injector(node, part, version)
                ports = self.backend.run_vertex(node, part, ins, env)
            res["n_in"] = sum(self.backend.count(p) for inp in ins for p in inp)
            for name in run.fifo_out:                              # streamed edges first: the consumer is waiting
                self.write_fifo(name, ports[0])
            if run.fifo_out and not self._needs_file(uid):
                res["n_out"] = sum(_count(p) for p in ports)      # streamed edges keep nothing
            else:
                for k, (path, size, nrec) in enumerate(write_ports(self.job_dir, uid, part, version, ports, self.compression)):
                    self._cache_put(f"{uid}_{part}_{k}.v{version}", ports[k])
                    res["out_sizes"].append(size)
                    res["out_bytes"] += size
                    res["n_out"] += nrec
            root = self.roots.get(uid)
            if root is not None:
                onode = root if root.type.value == "OutputTable" else root.members[-1]
                uri = onode.params.get("uri") or temp_output_uri(self.job_dir, onode.params["qexpr_id"])
                if store.supports_part_commit(uri):
                    tmp, final, size = store.write_part(uri, part, ports[0], onode.params.get("compression")
                                                        or self.ctx.output_data_compression_scheme)
                    tmp_v = f"{final}.v{version}.tmp"
                    os.replace(tmp, tmp_v)
                    res["part_file"] = [tmp_v, final, size]
            res["ok"] = True
        except BaseException as ex_:  # noqa: BLE001  (vertex code is user code; SystemExit must pass)
            if isinstance(ex_, (SystemExit, KeyboardInterrupt)) or getattr(self, "raise_errors", False):
                raise                                              # single-vertex replay under a debugger
            for name in run.fifo_out:
                abort_fifo(name)                                   # the consumers of the streamed edges fail fast
            res["message"] = "".join(traceback.format_exception_only(type(ex_), ex_)).strip()[:3000]
            if isinstance(ex_, InputReadError):
                res["bad"] = run.input_index.get((ex_.producer_uid, ex_.producer_part), -1)
            res["deterministic"] = getattr(ex_, "error_code", None) is not None and not getattr(ex_, "transient", False)
        res["seconds"] = time.perf_counter() - t0
        res["hits"] = run.cache_hits
        res["remote_reads"] = run.remote_reads
        res["bytes_in"] = list(run.input_bytes)
        return res

    def _is_fifo_edge(self, child, parent) -> bool:
        op = next((o for c, o in parent.children if c is child), None)
        return (getattr(child.channel_type, "value", "") == "MemoryFIFO" and op is not None
                and not self.ex.edge_is_cross_product(parent, child, op) and child.partition_count == parent.partition_count)

    def _needs_file(self, uid: int) -> bool:
        """False when every consumer of the stage reads it through a FIFO (a streamed edge keeps nothing)."""
        node = self.nodes[uid]
        return uid in self.roots or not node.parents or any(not self._is_fifo_edge(node, p) for p in node.parents)

    def run_command(self, cmd):
        """-> status block.  A plain START runs one vertex; a `subgraph` START runs the members of a start
        clique concurrently, wired by in-process FIFOs (DryadSubGraphVertex, subgraphvertex.cpp)."""
        import json
        from .command import CMD_START, STATE_COMPLETED, STATE_FAILED, VertexCommand
        extra = cmd.args[3:]
        subst: Dict[str, Optional[str]] = {}
        for a in extra:
            if a.startswith("subst="):
                subst = json.loads(a[6:])
            elif a.startswith("widths="):
                self._apply_widths(json.loads(a[7:]))
        if extra and extra[0] == "subgraph":
            results = self.run_subgraph(json.loads(extra[1]), subst)
            ok = all(r["ok"] for r in results)
            slim = [{k: r[k] for k in ("uid", "part", "gmv", "version", "ok", "n_in", "n_out", "seconds", "hits", "bad",
                                        "out_bytes", "part_file", "deterministic", "remote_reads")} | {"message": r["message"][:600]}
                    for r in results]
            return VertexCommand(CMD_START, cmd.vertex_id, cmd.version, state=STATE_COMPLETED if ok else STATE_FAILED,
                                 args=["subgraph", json.dumps(slim)])
        internal = extra[1:4] if extra and extra[0] == "internal" else None     # [kind, consumer uid, port]
        member = {"uid": cmd.vertex_id, "part": int(cmd.args[1]), "gmv": int(cmd.args[2]), "version": cmd.version,
                  "inputs": [u.split("://", 1)[1].split("/") for u, _len in cmd.inputs]}     # chan://<uid>/<part>/<version>
        r = self.run_member(_Run(self, member, subst, internal))
        outputs = [(f"port{k}", sz) for k, sz in enumerate(r["out_sizes"])]
        if r["part_file"]:
            outputs.append((f"{r['part_file'][0]}|{r['part_file'][1]}", r["part_file"][2]))
        chans = [(b, 1 if k == r["bad"] else 0) for k, b in enumerate(r["bytes_in"])]
        return VertexCommand(CMD_START, r["uid"], r["version"], state=STATE_COMPLETED if r["ok"] else STATE_FAILED,
                             exit_code=0 if r["ok"] else (2 if r["deterministic"] else 1), message=r["message"],
                             outputs=outputs, channel_status=chans,
                             args=[str(r["part"]), str(r["gmv"]), str(r["n_in"]), str(r["n_out"]), f"{r['seconds']:.6f}", str(r["hits"]),
                                   str(r["remote_reads"])])

    def run_subgraph(self, members: List[dict], subst) -> List[dict]:
        import threading
        from ..channel.factory import drop_fifo
        runs = [_Run(self, m, subst) for m in members]
        here = {(r.uid, r.part): r for r in runs}
        names: List[str] = []
        for r in runs:
            node = self.nodes[r.uid]
            for child, _op in node.children:
                p = here.get((child.uid, r.part))
                if p is not None and self._is_fifo_edge(child, node):
                    name = f"{child.uid}_{r.part}_to_{r.uid}_v{r.version}"
                    r.fifo_in[(child.uid, r.part)] = name
                    p.fifo_out.append(name)
                    names.append(name)
        from ..channel.factory import ensure_fifo
        for n in names:
            ensure_fifo(n, 32)              # exists before any member runs: a producer that fails first can still abort it
        results: List[Optional[dict]] = [None] * len(runs)

        def work(i):
            results[i] = self.run_member(runs[i])

        threads = [threading.Thread(target=work, args=(i,), name=f"member-{r.uid}.{r.part}") for i, r in enumerate(runs)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        for n in names:
            drop_fifo(n)
            self._handoff.pop(n, None)
        return [r for r in results if r is not None]

    # ---- command loop ---------------------------------------------------------------------------
    def serve(self, idle_timeout: float = 300.0) -> int:
        from .command import CMD_TERMINATE, STATE_COMPLETED, VertexCommand
        seen = 0
        got = None
        self.client.set(self.id, "hello", "ready", str(os.getpid()).encode())
        last = time.monotonic()
        while True:
            if got is None:
                try:
                    got = self.client.get(self.id, "cmd", seen, 2.0)
                except ConnectionError:
                    return 3                       # the graph manager is gone: nothing left to work for
            if got is None or got[0] <= seen:
                got = None
                if time.monotonic() - last > idle_timeout:
                    return 4
                continue
            seen = got[0]
            last = time.monotonic()
            cmd = VertexCommand.parse(got[2])
            got = None
            if cmd is None:
                continue
            if cmd.command == CMD_TERMINATE:
                return 0
            st = self.run_command(cmd)
            try:
                got = self.client.set(self.id, "status", "completed" if st.state == STATE_COMPLETED else "failed",
                                      st.serialize(), wait="cmd", seen=seen, wait_timeout=2.0)
            except ConnectionError:
                return 3

    def close(self):
        self.client.close()
        if self._cache:
            self._hostlib.cache_destroy(self._cache)
            self._cache = None


def temp_output_uri(job_dir: str, qexpr_id: int) -> str:
    return f"partfile://{os.path.join(job_dir, 'out', f'q{qexpr_id}.pt')}"


def main(argv: Optional[List[str]] = None) -> int:
    ap = argparse.ArgumentParser(prog="vertex_host")
    ap.add_argument("--service", required=True, help="URI of the process service whose mailbox carries the commands")
    ap.add_argument("--id", required=True, help="this host's mailbox name")
    ap.add_argument("--job-dir", required=True)
    ap.add_argument("--idle-timeout", type=float, default=300.0)
    a = ap.parse_args(argv)
    host = VertexHost(a.service, a.id, a.job_dir)
    prof = None
    if os.environ.get("DRYAD_HOST_PROFILE"):           # profile the vertices this host runs: <job dir>/log/host-<id>.prof
        import cProfile
        prof = cProfile.Profile()
        prof.enable()
    try:
        return host.serve(a.idle_timeout)
    finally:
        if prof is not None:
            prof.disable()
            prof.dump_stats(os.path.join(a.job_dir, "log", f"host-{a.id}.prof"))
        host.close()


if __name__ == "__main__":
    sys.exit(main())
