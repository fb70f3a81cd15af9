"""Node daemon: process table, versioned key/value mailbox and file server over HTTP.

Reference: ProcessService/ProcessService.cs:52-752 (Launch, BlockOnStatus, SetValue, Kill),
ProcessService/HttpServer.cs:42-707 (PUT create/launch/setvalue/kill, GET long-poll
`?key&version&timeout`, ranged file GET), headers X-Dryad-ValueVersion / X-Dryad-ProcessStatus
(ClusterInterface/Process.cs:236-240).  The mailbox itself is the native one (csrc/host/runtime.cpp);
this module is its HTTP face so a worker on another process can long-poll its command.
"""
from __future__ import annotations

import os
import subprocess
import threading
import urllib.parse
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Dict, Optional

from .services import Mailbox

H_VERSION, H_STATUS, H_EXIT = "X-Dryad-ValueVersion", "X-Dryad-ProcessStatus", "X-Dryad-ProcessExitCode"


class ProcessService:
    def __init__(self, root_dir: str, host: str = "127.0.0.1", port: int = 0):
        self.root = os.path.abspath(root_dir)
        os.makedirs(self.root, exist_ok=True)
        self.mailboxes: Dict[str, Mailbox] = {}
        self.procs: Dict[str, subprocess.Popen] = {}
        self._lock = threading.Lock()
        svc = self

        class Handler(BaseHTTPRequestHandler):
            protocol_version = "HTTP/1.1"       # keep-alive: a worker keeps ONE connection for all its polls and reports

            def log_message(self, *a):
                pass

            def setup(self):
                super().setup()
                import socket
                self.request.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)   # small messages: no 40 ms delayed-ACK stalls

            def end_headers(self):
                if not any(h.lower().startswith(b"content-length") for h in getattr(self, "_headers_buffer", [])):
                    self.send_header("Content-Length", "0")
                super().end_headers()

            def _mb(self, pid: str) -> Mailbox:
                with svc._lock:
                    m = svc.mailboxes.get(pid)
                    if m is None:
                        m = svc.mailboxes[pid] = Mailbox()
                    return m

            def do_PUT(self):
                u = urllib.parse.urlparse(self.path)
                q = urllib.parse.parse_qs(u.query)
                parts = u.path.strip("/").split("/")
                body = self.rfile.read(int(self.headers.get("Content-Length", 0) or 0))
                if len(parts) == 3 and parts[0] == "process" and parts[2] == "setvalue":
                    mb = self._mb(parts[1])
                    v = mb.set(q["key"][0], q.get("status", [""])[0], body)
                    got = None
                    if "wait" in q:
                        # report-and-wait: the same exchange long-polls another key (a worker hands in its
                        # status and receives its next command in one round trip)
                        got = mb.block_on_status(q["wait"][0], int(q.get("version", ["0"])[0]), float(q.get("timeout", ["1"])[0]))
                    self.send_response(200)
                    self.send_header(H_VERSION, str(got[0] if got else v))
                    if got:
                        self.send_header(H_STATUS, got[1])
                        self.send_header("X-Dryad-Waited", "1")
                        self.send_header("Content-Length", str(len(got[2])))
                    self.end_headers()
                    if got:
                        self.wfile.write(got[2])
                elif len(parts) == 3 and parts[0] == "process" and parts[2] == "launch":
                    cmd = body.decode().split("\x00")
                    log = open(os.path.join(svc.root, f"process-{parts[1]}-stdout.txt"), "wb")
                    with svc._lock:
                        svc.procs[parts[1]] = subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT, cwd=svc.root)
                    self.send_response(200)
                    self.end_headers()
                elif len(parts) == 3 and parts[0] == "process" and parts[2] == "kill":
                    with svc._lock:
                        p = svc.procs.get(parts[1])
                    if p is not None and p.poll() is None:
                        p.kill()
                    self.send_response(200)
                    self.end_headers()
                else:
                    self.send_response(404)
                    self.end_headers()

            def do_GET(self):
                u = urllib.parse.urlparse(self.path)
                q = urllib.parse.parse_qs(u.query)
                parts = u.path.strip("/").split("/")
                if len(parts) == 3 and parts[0] == "process" and parts[2] == "status":
                    got = self._mb(parts[1]).block_on_status(q["key"][0], int(q.get("version", ["0"])[0]),
                                                             float(q.get("timeout", ["1"])[0]))
                    with svc._lock:
                        p = svc.procs.get(parts[1])
                    self.send_response(200 if got else 204)
                    if got:
                        self.send_header(H_VERSION, str(got[0]))
                        self.send_header(H_STATUS, got[1])
                    if p is not None and p.poll() is not None:
                        self.send_header(H_EXIT, str(p.returncode))
                    if got:
                        self.send_header("Content-Length", str(len(got[2])))
                    self.end_headers()
                    if got:
                        self.wfile.write(got[2])
                elif parts and parts[0] == "file":
                    path = os.path.normpath(os.path.join(svc.root, *parts[1:]))
                    if not path.startswith(svc.root) or not os.path.isfile(path):
                        self.send_response(404)
                        self.end_headers()
                        return
                    size = os.path.getsize(path)
                    start, end = 0, size - 1
                    rng = self.headers.get("Range")
                    if rng and rng.startswith("bytes="):
                        a, _, b = rng[6:].partition("-")
                        start = int(a or 0)
                        end = int(b) if b else size - 1
                    with open(path, "rb") as f:
                        f.seek(start)
                        data = f.read(end - start + 1)
                    self.send_response(206 if rng else 200)
                    self.send_header("Content-Length", str(len(data)))
                    self.end_headers()
                    self.wfile.write(data)
                else:
                    self.send_response(404)
                    self.end_headers()

        class Server(ThreadingHTTPServer):
            # every worker of a job long-polls and reports at the same instant: the default listen
            # backlog of 5 drops connections (1 s SYN retransmits) from 8 ranks on
            request_queue_size = 256
            daemon_threads = True

        self.httpd = Server((host, port), Handler)
        self.port = self.httpd.server_address[1]
        self.uri = f"http://{host}:{self.port}"
        self._thread = threading.Thread(target=self.httpd.serve_forever, daemon=True)
        self._thread.start()

    def close(self):
        self.httpd.shutdown()
        self.httpd.server_close()
        for p in self.procs.values():
            if p.poll() is None:
                p.kill()
        for m in self.mailboxes.values():
            m.close()


class MailboxClient:
    """A process's side of the mailbox protocol: ONE keep-alive HTTP connection to the node daemon
    for all its long-polls and reports (ClusterInterface/HttpScheduler.cs:34-126 is the GM-side
    counterpart in the reference; the vertex host's side is dvertexpncontrol.cpp's status loop)."""

    def __init__(self, uri: str, who: str = "client"):
        self.uri, self.who = uri, who
        self._conn = None

    def request(self, method: str, path: str, body: bytes = b"", timeout: float = 30.0):
        """One exchange; reconnects once when the connection was dropped.  Raises ConnectionError when
        the daemon cannot be reached at all."""
        import http.client
        import socket
        u = urllib.parse.urlparse(self.uri)
        for attempt in (0, 1):
            try:
                if self._conn is None:
                    self._conn = http.client.HTTPConnection(u.hostname, u.port, timeout=timeout)
                    self._conn.connect()
                    self._conn.sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
                self._conn.sock.settimeout(timeout)
                self._conn.request(method, path, body=body or None,
                                   headers={"Content-Length": str(len(body))} if body else {})
                resp = self._conn.getresponse()
                return resp.status, resp.headers, resp.read()
            except (http.client.HTTPException, OSError) as ex:
                try:
                    self._conn.close()
                except Exception:  # noqa: BLE001
                    pass
                self._conn = None
                if attempt:
                    raise ConnectionError(f"{self.who}: the process service at {self.uri} is unreachable ({ex})") from ex

    def get(self, pid: str, key: str, seen: int, timeout: float):
        """Long-poll: (version, status, data) of `key` once its version exceeds `seen`, else None."""
        q = urllib.parse.urlencode({"key": key, "version": seen, "timeout": timeout})
        status, headers, data = self.request("GET", f"/process/{pid}/status?{q}", b"", timeout + 10.0)
        if status != 200:
            return None
        return int(headers.get(H_VERSION, "0")), headers.get(H_STATUS, ""), data

    def set(self, pid: str, key: str, status: str, data: bytes = b"", wait: Optional[str] = None, seen: int = 0,
            wait_timeout: float = 5.0, timeout: float = 30.0):
        """Set a value; with `wait`, the same exchange long-polls another key and returns its
        (version, status, data) when it changed within wait_timeout, else None."""
        q = {"key": key, "status": status}
        if wait:
            q.update({"wait": wait, "version": seen, "timeout": wait_timeout})
        st, headers, body = self.request("PUT", f"/process/{pid}/setvalue?{urllib.parse.urlencode(q)}", data,
                                         timeout + (wait_timeout if wait else 0.0))
        if wait and st == 200 and headers.get("X-Dryad-Waited") == "1":
            return int(headers.get(H_VERSION, "0")), headers.get(H_STATUS, ""), body
        return None

    def fetch_file(self, rel_path: str, start: int = 0, end: Optional[int] = None, timeout: float = 30.0) -> bytes:
        """Ranged GET from the daemon's file server (a channel file that lives on another computer)."""
        import http.client
        u = urllib.parse.urlparse(self.uri)
        c = http.client.HTTPConnection(u.hostname, u.port, timeout=timeout)
        try:
            hdr = {"Range": f"bytes={start}-{'' if end is None else end}"} if (start or end is not None) else {}
            c.request("GET", "/file/" + rel_path.lstrip("/"), headers=hdr)
            r = c.getresponse()
            data = r.read()
            if r.status not in (200, 206):
                raise FileNotFoundError(rel_path)
            return data
        finally:
            c.close()

    def close(self):
        if self._conn is not None:
            try:
                self._conn.close()
            except Exception:  # noqa: BLE001
                pass
            self._conn = None


def main(argv=None) -> int:
    """Daemon main (ProcessService/Program.cs): serve the mailbox, process table and file server of one node
    until SIGTERM / SIGINT.  Prints `DRYAD_PROCESS_SERVER_URI=<uri>` on stdout once it listens.

        python -m dryad_b200.runtime.process_service --root DIR [--host 127.0.0.1] [--port 0]
    """
    import argparse
    import signal
    import sys
    ap = argparse.ArgumentParser(prog="process_service")
    ap.add_argument("--root", required=True, help="directory served by the file server; process logs go here")
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=0)
    a = ap.parse_args(argv)
    svc = ProcessService(a.root, a.host, a.port)
    stop = threading.Event()
    signal.signal(signal.SIGTERM, lambda *_: stop.set())
    signal.signal(signal.SIGINT, lambda *_: stop.set())
    print(f"DRYAD_PROCESS_SERVER_URI={svc.uri}", flush=True)
    try:
        while not stop.wait(0.5):
            pass
    finally:
        svc.close()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
