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
        raw = hasattr(payload, "data") and hasattr(payload, "to_records")     # raw text partition: keep the bytes
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
    if hasattr(payload, "data") and getattr(payload, "_lines", 1) is None:
        return payload.data.count(b"\n")
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
    def __init__(self, host, uid, part):
        self.host, self.uid, self.part = host, uid, part

    def __getitem__(self, port):
        return self.host.load_input(self.uid, self.part, port)


class _LazyStage:
    """Stands where the in-process executor keeps `out[stage uid]`: a list of partitions, each a list
    of ports — here every access opens the channel on demand."""

    def __init__(self, host, uid, nparts):
        self.host, self.uid, self.nparts = host, uid, nparts

    def __len__(self):
        return self.nparts

    def __getitem__(self, part):
        if part < 0:
            part += self.nparts
        if not 0 <= part < self.nparts:
            raise IndexError(part)
        return _LazyPorts(self.host, self.uid, part)

    def __iter__(self):
        return (self[i] for i in range(self.nparts))


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
        self._versions: Dict[Tuple[int, int], int] = {}
        self._subst: Dict[str, Optional[str]] = {}
        self._input_index: Dict[Tuple[int, int], int] = {}
        self._input_bytes: List[int] = []
        self._cache_hits = 0
        # channel cache with spill: outputs of vertices that ran HERE are served from memory
        from .. import _native as N
        self._hostlib = N.host()
        low = int(os.environ.get("DRYAD_CACHE_LOW_WATER", str(64 << 20)))
        high = int(os.environ.get("DRYAD_CACHE_HIGH_WATER", str(128 << 20)))
        spill = os.path.join(job_dir, "spill", host_id)
        os.makedirs(spill, exist_ok=True)
        self._cache = self._hostlib.cache_create(spill.encode(), low, high)
        self._cached: Dict[str, int] = {}

    # ---- channels -------------------------------------------------------------------------------
    def load_input(self, uid: int, part: int, port: int):
        from .executor import InputReadError
        hops = 0
        while f"{uid}/{part}" in self._subst and hops < 64:       # run-time rewrites: follow the splice chain
            t = self._subst[f"{uid}/{part}"]
            if t is None:
                return []                                          # folded into an internal vertex's output
            uid, part, port = int(t.split("/")[0]), int(t.split("/")[1]), 0
            hops += 1
        ver = self._versions.get((uid, part))
        if ver is None:
            raise InputReadError(uid, part, f"stage {uid}[{part}] is not an input of this vertex")
        key = f"{uid}_{part}_{port}.v{ver}"
        idx = self._input_index.get((uid, part), -1)
        n = self._cached.get(key)
        if n is not None:
            import ctypes as C
            buf = C.create_string_buffer(n)
            got = self._hostlib.cache_get(self._cache, key.encode(), buf, n)
            if got == n:
                self._cache_hits += 1
                if idx >= 0:
                    self._input_bytes[idx] += n
                return pickle.loads(buf.raw)
        try:
            payload, size = read_channel(self.job_dir, uid, part, port, ver)
        except Exception as ex:  # noqa: BLE001  (missing, truncated or corrupt: all "cannot read")
            raise InputReadError(uid, part, f"cannot read channel {key}: {type(ex).__name__}: {ex}") from ex
        if idx >= 0:
            self._input_bytes[idx] += size
        return payload

    def _cache_put(self, key: str, payload) -> None:
        try:
            b = pickle.dumps(payload, protocol=pickle.HIGHEST_PROTOCOL)
        except Exception:  # noqa: BLE001
            return
        if len(b) <= (32 << 20) and self._hostlib.cache_put(self._cache, key.encode(), b, len(b)) == 0:
            self._cached[key] = len(b)

    # ---- one vertex -----------------------------------------------------------------------------
    def run_vertex(self, cmd):
        from ..channel import store
        from .command import CMD_START, STATE_COMPLETED, STATE_FAILED, VertexCommand
        from .executor import InputReadError
        from .vertex_ops import VertexEnv
        uid, version = cmd.vertex_id, cmd.version
        part, gmv = int(cmd.args[1]), int(cmd.args[2])
        extra = cmd.args[3:]
        internal = extra[1:4] if extra and extra[0] == "internal" else None     # [kind, consumer uid, port]
        self._subst = {}
        for a in extra:
            if a.startswith("subst="):
                import json
                self._subst = json.loads(a[6:])
            elif a.startswith("widths="):
                import json
                w = json.loads(a[7:])                   # widths decided at run time by the dynamic distributor
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
        node = self.nodes[uid] if internal is None else None
        self._versions, self._input_index = {}, {}
        for k, (u, _len) in enumerate(cmd.inputs):                     # chan://<uid>/<part>/<version>
            a = u.split("://", 1)[1].split("/")
            self._versions[(int(a[0]), int(a[1]))] = int(a[2])
            self._input_index[(int(a[0]), int(a[1]))] = k
        self._input_bytes = [0] * len(cmd.inputs)
        self._cache_hits = 0
        self.ex.out = {n.uid: _LazyStage(self, n.uid, n.partition_count) for n in self.plan.nodes}
        t0 = time.perf_counter()
        outputs: List[Tuple[str, int]] = []
        try:
            env = VertexEnv(self.ctx, uid, part, version)
            if internal is not None:
                # an internal vertex of a run-time rewrite: fold / copy exactly the channels named in the command
                from . import vertex_ops
                subst, self._subst = self._subst, {}
                # producers of the plan are read at the consumer's port, outputs of other internal vertices at port 0
                payloads = [self.load_input(u, s, int(internal[2]) if u < DYN_UID_BASE else 0) for (u, s) in self._input_index]
                self._subst = subst
                ins = [payloads]
                if internal[0] == "copy":
                    ports = [payloads[0]]
                else:
                    ports = [vertex_ops.internal_combine(self.nodes[int(internal[1])], [self.backend.to_records(p) for p in payloads], env)]
            else:
                ins = self.ex._inputs(node, part)
                if self.ctx.fault_injector is not None:
                    self.ctx.fault_injector(node, part, version)
                ports = self.backend.run_vertex(node, part, ins, env)
            n_in = sum(self.backend.count(p) for inp in ins for p in inp)
            n_out = 0
            for k, (path, size, nrec) in enumerate(write_ports(self.job_dir, uid, part, version, ports, self.compression)):
                self._cache_put(f"{uid}_{part}_{k}.v{version}", ports[k])
                outputs.append((path, size))
                n_out += nrec
            root = self.roots.get(uid)
            if root is not None:
                onode = root if root.type.value == "OutputTable" else root.members[-1]
                uri = onode.params.get("uri") or temp_output_uri(self.job_dir, onode.params["qexpr_id"])
                if store.supports_part_commit(uri):
                    tmp, final, size = store.write_part(uri, part, ports[0], onode.params.get("compression")
                                                        or self.ctx.output_data_compression_scheme)
                    tmp_v = f"{final}.v{version}.tmp"
                    os.replace(tmp, tmp_v)
                    outputs.append((f"{tmp_v}|{final}", size))
            dt = time.perf_counter() - t0
            return VertexCommand(CMD_START, uid, version, state=STATE_COMPLETED, outputs=outputs,
                                 args=[str(part), str(gmv), str(n_in), str(n_out), f"{dt:.6f}", str(self._cache_hits)],
                                 channel_status=[(b, 0) for b in self._input_bytes])
        except BaseException as ex:  # noqa: BLE001  (vertex code is user code; SystemExit must pass)
            if isinstance(ex, (SystemExit, KeyboardInterrupt)):
                raise
            text = "".join(traceback.format_exception_only(type(ex), ex)).strip()
            chans = [(b, 0) for b in self._input_bytes]
            if isinstance(ex, InputReadError):
                k = self._input_index.get((ex.producer_uid, ex.producer_part), -1)
                if k >= 0:
                    chans[k] = (chans[k][0], 1)
            deterministic = getattr(ex, "code", None) is not None and not getattr(ex, "transient", False)
            return VertexCommand(CMD_START, uid, version, state=STATE_FAILED, exit_code=2 if deterministic else 1,
                                 message=text[:4000], channel_status=chans,
                                 args=[str(part), str(gmv), "0", "0", f"{time.perf_counter() - t0:.6f}", "0"])

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
            st = self.run_vertex(cmd)
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
    try:
        return host.serve(a.idle_timeout)
    finally:
        host.close()


if __name__ == "__main__":
    sys.exit(main())
